"""Drop-in for ``matcha.models.components.flow_matching`` backed by the sm_100a decoder kernels.

Mirrors /root/reference/matcha/models/components/flow_matching.py:
  BaseConditionalFlowMatching.forward / solve_ode_euler / compute_loss      (:42-151)
  ConditionalFlowMatching.__init__ (builds ``self.estimator = Decoder(...)``)  (:162-189)
Same names, argument meaning and results; ``solve_ode_euler`` hands the whole loop to one C call.
``compute_loss`` returns the forward value only (the decoder kernels are inference kernels; the
backward pass of training is outside this path); its phi_t construction and masked-MSE
reduction are fused into the first / last kernel of the decoder eval (csrc/decoder.cu).
"""
import torch

from ._lib import check, load_library
from .decoder import Decoder
from .monotonic_align import _stream_ptr


class BaseConditionalFlowMatching(torch.nn.Module):
    def __init__(self, n_feats, cfm_params, n_spks=1, spk_emb_dim=128):
        super().__init__()
        self.n_feats = n_feats
        self.n_spks = n_spks
        self.spk_emb_dim = spk_emb_dim
        self.solver = cfm_params.solver if hasattr(cfm_params, "solver") else "euler"
        self.sigma_min = cfm_params.sigma_min if hasattr(cfm_params, "sigma_min") else 1e-4
        self.estimator = None

    @torch.inference_mode()
    def forward(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None):
        z = torch.randn_like(mu) * temperature                                    # flow_matching.py:59
        return self.solve_ode_euler(z, self._host_t_span(n_timesteps, mu.device), mu, mask, spks, cond)

    def _host_t_span(self, n_timesteps, device):
        """torch.linspace(0, 1, n + 1, device=mu.device) (:62), read back ONCE per (n, device) and kept on the host: the C entry
        point takes the (t_k, dt_k) grid as host floats, so handing solve_euler a device tensor costs a device synchronisation
        on every call (one pipeline bubble per synthesised batch)."""
        key = (int(n_timesteps), str(device))
        cache = self.__dict__.setdefault("_t_span_host", {})
        if key not in cache:
            cache[key] = torch.linspace(0, 1, n_timesteps + 1, device=device).to("cpu", torch.float32)
        return cache[key]

    def solve_ode_euler(self, x, t_span, mu, mask, spks, cond):
        return self.estimator.solve_euler(x, t_span, mu, mask)                    # :67-104 in one call

    def compute_loss(self, x1, mask, mu, spks=None, cond=None, *, t=None, z=None):
        """flow_matching.py:106-151, forward value.  ``t`` [B,1,1] / ``z`` may be injected (tests share
        them with the oracle); otherwise they are drawn in the reference's RNG order (t, then z).
        One C call (matcha_b200_cfm_loss_forward): in the 16-bit modes phi_t is formed inside the decoder's input-packing
        kernel and the squared error inside the final projection's epilogue; u_pred / u_target are never stored."""
        if not x1.is_cuda:
            raise RuntimeError("matcha_b200.CFM needs CUDA tensors (sm_100a); there is no CPU fallback")
        dev = x1.device
        batch_size = mu.shape[0]
        if t is None:
            t = torch.rand([batch_size, 1, 1], device=dev, dtype=mu.dtype)        # :130
        if z is None:
            z = torch.randn_like(x1)                                              # :133
        sums, phi_t, _ = self.estimator.loss_forward(x1, z, t.reshape(batch_size), mu, mask, self.sigma_min)   # :138-148
        loss = (sums[0] / (sums[1] * x1.shape[1])).float()                       # :149
        return loss, phi_t.to(x1.dtype)


class ConditionalFlowMatching(BaseConditionalFlowMatching):
    def __init__(self, in_channels, out_channel, cfm_params, decoder_params, n_spks=1, spk_emb_dim=64, precision="fp32"):
        super().__init__(n_feats=in_channels, cfm_params=cfm_params, n_spks=n_spks, spk_emb_dim=spk_emb_dim)
        if n_spks > 1:
            in_channels = in_channels + spk_emb_dim
        self.estimator = Decoder(in_channels=in_channels, out_channels=out_channel, precision=precision, **decoder_params)

    def load_state_dict(self, state_dict, strict=True):
        self.estimator.load_state_dict(state_dict, strict=strict)
        return self


CFM = ConditionalFlowMatching   # the name matcha_tts.py imports
