"""``MatchaTTS`` with every device step on the sm_100a kernels: the host-side mirror of
/root/reference/matcha/models/matcha_tts.py for the two flows BASELINE configs[3] / [4] name.

  * ``synthesise(x, x_lengths, n_timesteps, temperature=1.0, length_scale=1.0)``   matcha_tts.py:178-249
      text encoder -> durations -> generate_path / mu_y -> CFM Euler solve -> denormalised mel
  * ``forward(x, x_lengths, y, y_lengths)``                                         matcha_tts.py:251-324
      text encoder -> log-prior -> MAS -> logw_ / mu_y -> duration, prior and flow-matching losses (FORWARD VALUES:
      the kernels are inference kernels, the backward pass of training is outside this path)
Same argument meaning and result layout as the reference; ``state_dict`` keys are the reference's (``encoder.*``,
``decoder.estimator.*``).  The ``out_size`` random cut (:290-312) is mirrored too: offsets are drawn on the host with Python's
``random`` exactly as the reference does, and the cut is applied to ``y``, ``attn`` and ``mu_y`` (cutting ``mu_y`` is the same as
forming it from the cut ``attn``: every column of ``attn`` is one-hot).  No CPU fallback.
"""
import datetime as dt
import math
import random
from collections import OrderedDict
from types import SimpleNamespace

import torch

from . import alignment
from .flow_matching import CFM
from .text_encoder import TextEncoder, default_params

_DECODER_PARAMS = dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)


class MatchaTTS(torch.nn.Module):
    """``MatchaTTS(n_vocab, out_channels=80, hidden_channels=192)`` as in the reference's simplified constructor
    (matcha_tts.py:108-176); ``precision`` selects the arithmetic: "fp32" (SIMT everywhere), "fp16" (tensor cores in the
    encoder and the decoder) or "bf16" (decoder operands in bf16, encoder in fp16)."""

    def __init__(self, n_vocab, out_channels=80, hidden_channels=192, data_statistics=None, prior_loss=True,
                 use_precomputed_durations=False, precision="fp16"):
        super().__init__()
        if (out_channels, hidden_channels) != (80, 192):
            raise NotImplementedError("matcha_b200.MatchaTTS implements the reference's configured sizes only (80 mel bins, 192 channels)")
        self.n_vocab, self.n_spks, self.n_feats = n_vocab, 1, out_channels
        self.prior_loss = prior_loss
        self.use_precomputed_durations = use_precomputed_durations
        ep, dp = default_params()
        self.encoder = TextEncoder("RoPE Encoder", ep, dp, n_vocab, precision="fp32" if precision == "fp32" else "fp16")
        self.decoder = CFM(in_channels=2 * out_channels, out_channel=out_channels,
                           cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4), decoder_params=_DECODER_PARAMS,
                           precision=precision)
        data_statistics = data_statistics or {"mel_mean": 0.0, "mel_std": 1.0}
        self.mel_mean, self.mel_std = float(data_statistics["mel_mean"]), float(data_statistics["mel_std"])

    def initialize_weights(self, seed=0):
        self.encoder.initialize_weights(seed)
        self.decoder.estimator.initialize_weights(seed + 1)
        return self

    def load_state_dict(self, state_dict, strict=True):
        enc = OrderedDict((k[len("encoder."):], v) for k, v in state_dict.items() if k.startswith("encoder."))
        dec = OrderedDict((k, v) for k, v in state_dict.items() if k.startswith("decoder."))
        self.encoder.load_state_dict(enc, strict)
        self.decoder.load_state_dict(dec, strict)
        return self

    def state_dict(self, *a, **k):
        sd = OrderedDict(("encoder." + k_, v) for k_, v in self.encoder.state_dict().items())
        sd.update(("decoder.estimator." + k_, v) for k_, v in self.decoder.estimator.state_dict().items())
        return sd

    @torch.inference_mode()
    def synthesise(self, x, x_lengths, n_timesteps, temperature=1.0, length_scale=1.0):
        t0 = dt.datetime.now()
        mu_x, logw, x_mask = self.encoder(x, x_lengths)                                          # :212
        pre = alignment.synthesis_prestep(mu_x, logw, x_mask, length_scale)                      # :214-227
        y_max_length = pre["y_max_length"]
        mu_y, y_mask = pre["mu_y"], pre["y_mask"]
        encoder_outputs = mu_y[:, :, :y_max_length]
        decoder_outputs = self.decoder(mu_y, y_mask, n_timesteps, temperature)[:, :, :y_max_length]   # :230-231
        t = (dt.datetime.now() - t0).total_seconds()
        rtf = t * 22050 / (decoder_outputs.shape[-1] * 256)
        return {"encoder_outputs": encoder_outputs, "decoder_outputs": decoder_outputs, "attn": pre["attn"][:, :, :y_max_length],      # as the reference writes it (:243): the cut lands on the T_x axis
                "mel": decoder_outputs * self.mel_std + self.mel_mean, "mel_lengths": pre["y_lengths"], "rtf": rtf}

    @torch.no_grad()
    def forward(self, x, x_lengths, y, y_lengths, out_size=None, cond=None, durations=None, *, t=None, z=None):
        """Returns (dur_loss, prior_loss, diff_loss, attn) as the reference does; ``t`` / ``z`` may be injected for the
        flow-matching loss (tests share them with the oracle)."""
        dev = y.device
        mu_x, logw, x_mask = self.encoder(x, x_lengths)                                          # :271
        T_y = y.shape[-1]
        y_lengths = y_lengths.to(dev)
        y_mask = (torch.arange(T_y, device=dev).unsqueeze(0) < y_lengths.unsqueeze(1)).unsqueeze(1).to(x_mask.dtype)
        if self.use_precomputed_durations:
            r = alignment.expand(durations.squeeze(1).to(dev), mu_x, x_lengths, y_lengths, y=y, want_attn=True, want_logw=True,
                                 want_prior=bool(self.prior_loss))                               # :278-279
            attn, logw_, mu_y, sums = r["attn"], r["logw"].unsqueeze(1), r["mu_y"], r["prior_sums"]
            prior = (sums[0] / (sums[1] * self.n_feats)).float() if self.prior_loss else 0
        else:
            r = alignment.align(mu_x, y, x_lengths, y_lengths, prior_loss=self.prior_loss)       # :281-291, 314-321
            attn, logw_, mu_y, prior = r["attn"], r["logw_"], r["mu_y"], r["prior_loss"]
        dur_loss = torch.sum((logw - logw_) ** 2) / torch.sum(x_lengths.to(dev))                 # utils/model.py:117-135
        if out_size is not None:                                                                 # :290-312
            max_offset = (y_lengths - out_size).clamp(0)
            offset_ranges = list(zip([0] * max_offset.shape[0], max_offset.cpu().numpy()))
            out_offset = [int(random.choice(range(start, end))) if end > start else 0 for start, end in offset_ranges]
            attn_cut = torch.zeros(attn.shape[0], attn.shape[1], out_size, dtype=attn.dtype, device=dev)
            y_cut = torch.zeros(y.shape[0], self.n_feats, out_size, dtype=y.dtype, device=dev)
            mu_cut = torch.zeros(y.shape[0], self.n_feats, out_size, dtype=mu_y.dtype, device=dev)
            y_cut_lengths = []
            yl_host = y_lengths.cpu().tolist()
            for i, off in enumerate(out_offset):
                n = out_size + min(yl_host[i] - out_size, 0)
                y_cut_lengths.append(n)
                y_cut[i, :, :n] = y[i, :, off:off + n]
                attn_cut[i, :, :n] = attn[i, :, off:off + n]
                mu_cut[i, :, :n] = mu_y[i, :, off:off + n]
            y_cut_lengths = torch.tensor(y_cut_lengths, device=dev)
            y_mask = (torch.arange(out_size, device=dev).unsqueeze(0) < y_cut_lengths.unsqueeze(1)).unsqueeze(1).to(x_mask.dtype)
            attn, y, mu_y = attn_cut, y_cut, mu_cut
            if self.prior_loss:                                                                  # :319-322 on the cut tensors
                prior = torch.sum(0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask) / (torch.sum(y_mask) * self.n_feats)
        diff_loss, _ = self.decoder.compute_loss(x1=y, mask=y_mask, mu=mu_y, cond=cond, t=t, z=z)   # :317
        return dur_loss, prior, diff_loss, attn
