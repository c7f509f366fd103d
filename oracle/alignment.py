"""oracle/alignment.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU (torch fp32) restatement of the code either side of MAS and the decoder in the reference's
``MatchaTTS``.  Each function cites the reference lines it follows.
Parity pinned: tests/test_oracle_alignment.py checks these against tests/golden/alignment_golden.npz,
which tests/golden/make_alignment_golden.py produced by running the reference's own
``MatchaTTS.forward`` / ``MatchaTTS.synthesise`` (unmodified, on CPU, with import stubs for the absent
pytorch_lightning / matplotlib and oracle/_ref's compiled core.pyx) and recording the tensors at
these seams.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import math

import numpy as np
import torch

from . import mas as omas


def sequence_mask(length, max_length=None):
    """utils/model.py:13-34"""
    if max_length is None:
        max_length = int(length.max())
    x = torch.arange(max_length, dtype=length.dtype, device=length.device)
    return x.unsqueeze(0) < length.unsqueeze(1)


def fix_len_compatibility(length, num_downsamplings_in_unet=2):
    """utils/model.py:37-57"""
    f = 2 ** num_downsamplings_in_unet
    return int(math.ceil(float(length) / f) * f)


def generate_path(duration, mask):
    """utils/model.py:77-114 (build_alignment_path): duration [B,T_x], mask [B,T_x,T_y]."""
    b, t_x, t_y = mask.shape
    cum_duration = torch.cumsum(duration, 1)
    cum_flat = cum_duration.view(b * t_x)
    path = sequence_mask(cum_flat, t_y).to(mask.dtype).view(b, t_x, t_y)
    shifted = torch.nn.functional.pad(path, [0, 0, 1, 0, 0, 0])[:, :-1]
    return (path - shifted) * mask


def log_prior(mu_x, y):
    """matcha_tts.py:277-282; mu_x [B,C,T_x], y [B,C,T_y] -> [B,T_x,T_y]."""
    n_feats = mu_x.shape[1]
    const = -0.5 * math.log(2 * math.pi) * n_feats
    factor = -0.5 * torch.ones(mu_x.shape, dtype=mu_x.dtype, device=mu_x.device)
    y_square = torch.matmul(factor.transpose(1, 2), y ** 2)
    y_mu_double = torch.matmul(2.0 * (factor * mu_x).transpose(1, 2), y)
    mu_square = torch.sum(factor * (mu_x ** 2), 1).unsqueeze(-1)
    return y_square - y_mu_double + mu_square + const


def maximum_path(value, mask, use_ref=False):
    """monotonic_align/__init__.py:40-55 through the C restatement (or oracle/_ref's core.pyx)."""
    fn = omas.maximum_path_ref if use_ref else omas.maximum_path_oracle
    path = fn(value.detach().numpy().astype(np.float32), mask.detach().numpy().astype(np.float32))
    return torch.from_numpy(path).to(value.dtype)


def align(mu_x, y, x_lengths, y_lengths, use_ref=False):
    """matcha_tts.py:270-287, 314-315, 320-321 without the out_size cut.
    Returns dict(log_prior, attn, logw_ [B,1,T_x], mu_y, prior_loss)."""
    T_x, T_y = mu_x.shape[-1], y.shape[-1]
    x_mask = sequence_mask(x_lengths, T_x).unsqueeze(1).to(mu_x.dtype)
    y_mask = sequence_mask(y_lengths, T_y).unsqueeze(1).to(x_mask)
    attn_mask = x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)
    lp = log_prior(mu_x, y)
    attn = maximum_path(lp, attn_mask.squeeze(1), use_ref=use_ref)
    logw_ = torch.log(1e-8 + torch.sum(attn.unsqueeze(1), -1)) * x_mask
    mu_y = torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    prior_loss = torch.sum(0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask)
    prior_loss = prior_loss / (torch.sum(y_mask) * mu_x.shape[1])
    return {"log_prior": lp, "attn": attn, "logw_": logw_, "mu_y": mu_y, "prior_loss": prior_loss}


def duration_loss(logw, logw_, lengths):
    """utils/model.py:117-135"""
    return torch.sum((logw - logw_) ** 2) / torch.sum(lengths)


def synthesis_prestep(mu_x, logw, x_mask, length_scale=1.0):
    """matcha_tts.py:218-229"""
    w = torch.exp(logw) * x_mask
    w_ceil = torch.ceil(w) * length_scale
    y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()
    y_max_length = int(y_lengths.max())
    y_max_length_ = fix_len_compatibility(y_max_length)
    y_mask = sequence_mask(y_lengths, y_max_length_).unsqueeze(1).to(x_mask.dtype)
    attn_mask = x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)
    attn = generate_path(w_ceil.squeeze(1), attn_mask.squeeze(1)).unsqueeze(1)
    mu_y = torch.matmul(attn.squeeze(1).transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    return {"mu_y": mu_y, "y_mask": y_mask, "attn": attn, "y_lengths": y_lengths, "y_max_length": y_max_length}


def cfm_mix(x1, z, t, sigma_min=1e-4):
    """flow_matching.py:138, 142: (phi_t, u_target) for t [B,1,1]."""
    phi_t = (1 - (1 - sigma_min) * t) * z + t * x1
    u = x1 - (1 - sigma_min) * z
    return phi_t, u


def cfm_mse(u_pred, u, mask):
    """flow_matching.py:148-149"""
    loss = torch.nn.functional.mse_loss(u_pred, u, reduction="sum")
    return loss / (torch.sum(mask) * u.shape[1])
