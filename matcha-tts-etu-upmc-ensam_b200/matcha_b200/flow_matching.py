"""Drop-in for ``matcha.models.components.flow_matching`` backed by the sm_100a decoder kernels.

Mirrors /root/reference/matcha/models/components/flow_matching.py:
  BaseConditionalFlowMatching.forward / solve_ode_euler / compute_loss      (:42-151)
  ConditionalFlowMatching.__init__ (builds ``self.estimator = Decoder(...)``)  (:162-189)
Same names, argument meaning and results; ``solve_ode_euler`` hands the whole loop to one C call.
``compute_loss`` returns the forward value only (the decoder kernels are inference kernels; the
backward pass of training is outside this path).
"""
import torch
import torch.nn.functional as F

from .decoder import Decoder


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
        t_span = torch.linspace(0, 1, n_timesteps + 1, device=mu.device)          # :62
        return self.solve_ode_euler(z, t_span, mu, mask, spks, cond)

    def solve_ode_euler(self, x, t_span, mu, mask, spks, cond):
        return self.estimator.solve_euler(x, t_span, mu, mask)                    # :67-104 in one call

    def compute_loss(self, x1, mask, mu, spks=None, cond=None):
        batch_size = mu.shape[0]
        t = torch.rand([batch_size, 1, 1], device=mu.device, dtype=mu.dtype)      # :130 (RNG order: t, then z)
        z = torch.randn_like(x1)                                                  # :133
        phi_t = (1 - (1 - self.sigma_min) * t) * z + t * x1                       # :138
        u_target = x1 - (1 - self.sigma_min) * z                                  # :142
        u_pred = self.estimator(phi_t, mask, mu, t.squeeze(), cond)               # :145
        loss = F.mse_loss(u_pred, u_target, reduction="sum")                      # :148
        loss = loss / (torch.sum(mask) * u_target.shape[1])                       # :149
        return loss, phi_t


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
