"""Device versions of the code either side of MAS and the decoder in ``MatchaTTS``.

Mirrors, with the reference's names and argument meaning:
  * ``log_prior(mu_x, y)``                 the no_grad block of MatchaTTS.forward    matcha_tts.py:277-282
  * ``generate_path(duration, mask)``      utils/model.py:77-114 (build_alignment_path)
  * ``align(mu_x, y, x_lengths, y_lengths)``  log-prior -> MAS -> logw_ / mu_y / prior loss
                                           matcha_tts.py:277-287, 314-315, 320-321 (no ``out_size`` cut)
  * ``synthesis_prestep(mu_x, logw, x_mask, length_scale)``                          matcha_tts.py:218-229
All tensors are CUDA fp32; everything is asynchronous on the current stream except where the
reference itself synchronises (``y_lengths.max()`` in synthesis).  No CPU fallback.
"""
import ctypes
import math

import torch

from ._lib import check, load_library
from .monotonic_align import PATH_F32, _stream_ptr, lengths_from_mask


def _f32c(t):
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()


def _i32(t, device):
    return t.to(device=device, dtype=torch.int32).contiguous()


def _ptr(t):
    return t.data_ptr() if t is not None else None


def log_prior(mu_x, y, x_lengths=None, y_lengths=None):
    """mu_x [B,C,T_x], y [B,C,T_y] -> log N(y; mu_x, I) summed over channels, [B,T_x,T_y]
    (matcha_tts.py:277-282).  With lengths, tiles wholly outside an item's valid block are left
    unwritten (MAS never reads them)."""
    if not mu_x.is_cuda:
        raise RuntimeError("matcha_b200.alignment needs CUDA tensors (sm_100a); there is no CPU fallback")
    mu_x, y = _f32c(mu_x), _f32c(y)
    B, C, T_x = mu_x.shape
    T_y = y.shape[-1]
    assert y.shape[:2] == (B, C)
    dev = mu_x.device
    out = torch.empty((B, T_x, T_y), dtype=torch.float32, device=dev)
    tx = _i32(x_lengths, dev) if x_lengths is not None else None
    ty = _i32(y_lengths, dev) if y_lengths is not None else None
    if B == 0:
        return out
    with torch.cuda.device(dev):
        check(load_library().matcha_b200_align_log_prior(out.data_ptr(), mu_x.data_ptr(), y.data_ptr(), _ptr(tx), _ptr(ty),
                                                         B, C, T_x, T_y, _stream_ptr(dev)), "align_log_prior")
    return out


def maximum_path_with_durations(value, t_xs, t_ys, max_neg_val=-1e9):
    """MAS returning (path fp32 [B,T_x,T_y], durations int32 [B,T_x]); durations come from the
    backtrack's start frames, i.e. ``path.sum(-1)`` without re-reading the path."""
    value = _f32c(value)
    assert value.is_cuda and value.dim() == 3
    B, T_x, T_y = value.shape
    dev = value.device
    path = torch.empty_like(value)
    dur = torch.empty((B, T_x), dtype=torch.int32, device=dev)
    if B == 0:
        return path, dur
    tx, ty = _i32(t_xs, dev), _i32(t_ys, dev)        # named: the buffers must outlive the launch
    with torch.cuda.device(dev):
        check(load_library().matcha_b200_mas_align(
            path.data_ptr(), PATH_F32, dur.data_ptr(), value.data_ptr(), None,
            tx.data_ptr(), ty.data_ptr(), B, T_x, T_y,
            ctypes.c_float(max_neg_val), None, _stream_ptr(dev)), "mas_align")
    return path, dur


def expand(durations, mu_x=None, x_lengths=None, y_lengths=None, T_y=None, y=None,
           want_attn=False, want_mu_y=True, want_logw=False, want_prior=False):
    """Duration expansion (one kernel).  durations [B,T_x] int32 or float; mu_x [B,C,T_x].
    Returns a dict with the requested ``attn`` [B,T_x,T_y], ``mu_y`` [B,C,T_y], ``logw`` [B,T_x],
    ``prior_sums`` (fp64 [2]: numerator of matcha_tts.py:320 and sum(y_mask))."""
    assert durations.is_cuda and durations.dim() == 2
    dev = durations.device
    is_int = durations.dtype == torch.int32
    d = durations.contiguous() if is_int else _f32c(durations)
    B, T_x = d.shape
    if y is not None:
        y = _f32c(y)
        T_y = y.shape[-1]
    assert T_y is not None
    C = 1
    if mu_x is not None:
        mu_x = _f32c(mu_x)
        C = mu_x.shape[1]
        assert mu_x.shape == (B, C, T_x)
    res = {}
    attn = torch.empty((B, T_x, T_y), dtype=torch.float32, device=dev) if want_attn else None
    mu_y = torch.empty((B, C, T_y), dtype=torch.float32, device=dev) if (want_mu_y and mu_x is not None) else None
    logw = torch.empty((B, T_x), dtype=torch.float32, device=dev) if want_logw else None
    prior = torch.empty(2, dtype=torch.float64, device=dev) if want_prior else None
    tx = _i32(x_lengths, dev) if x_lengths is not None else None
    ty = _i32(y_lengths, dev) if y_lengths is not None else None
    if B > 0:
        with torch.cuda.device(dev):
            check(load_library().matcha_b200_align_expand(
                d.data_ptr(), 1 if is_int else 0, _ptr(mu_x), _ptr(y), _ptr(tx), _ptr(ty),
                _ptr(attn), _ptr(mu_y), _ptr(logw), _ptr(prior), B, C, T_x, T_y, _stream_ptr(dev)), "align_expand")
    elif prior is not None:
        prior.zero_()
    res["attn"], res["mu_y"], res["logw"], res["prior_sums"] = attn, mu_y, logw, prior
    return res


def generate_path(duration, mask):
    """utils/model.py:77-114: duration [B,T_x] (>= 0), mask [B,T_x,T_y] -> 0/1 path in ``mask``'s dtype.
    The mask must be the outer product of two prefix masks (what matcha_tts.py:223 builds)."""
    assert mask.is_cuda and mask.dim() == 3
    t_xs, t_ys = lengths_from_mask(mask)
    r = expand(duration.to(mask.device), None, t_xs, t_ys, T_y=mask.shape[-1], want_attn=True, want_mu_y=False)
    return r["attn"].to(mask.dtype)


def align(mu_x, y, x_lengths, y_lengths, prior_loss=True):
    """The alignment part of the training step (matcha_tts.py:277-287, 314-315, 320-321):
    returns dict(attn [B,T_x,T_y], durations int32 [B,T_x], logw_ [B,1,T_x], mu_y [B,C,T_y],
    prior_loss scalar or 0).  ``out_size`` cutting (:290-312) is host-side slicing and stays with the caller."""
    dev = mu_x.device
    tx, ty = _i32(x_lengths, dev), _i32(y_lengths, dev)
    lp = log_prior(mu_x, y, tx, ty)
    attn, dur = maximum_path_with_durations(lp, tx, ty)
    r = expand(dur, mu_x, tx, ty, y=y, want_logw=True, want_prior=bool(prior_loss))
    out = {"attn": attn, "durations": dur, "logw_": r["logw"].unsqueeze(1), "mu_y": r["mu_y"], "prior_loss": 0}
    if prior_loss:
        s = r["prior_sums"]
        out["prior_loss"] = (s[0] / (s[1] * mu_x.shape[1])).float()
    return out


def fix_len_compatibility(length, num_downsamplings_in_unet=2):
    """utils/model.py:37-57 on a Python int."""
    f = 2 ** num_downsamplings_in_unet
    return int(math.ceil(int(length) / f) * f)


def synthesis_prestep(mu_x, logw, x_mask, length_scale=1.0):
    """matcha_tts.py:218-229: durations from the predicted log-durations, then mu_y / attn / y_mask
    for the decoder.  mu_x [B,C,T_x], logw and x_mask [B,1,T_x].  Returns dict(mu_y [B,C,T_y_], y_mask
    [B,1,T_y_], attn [B,1,T_x,T_y_], y_lengths [B] int64, y_max_length int)."""
    w = torch.exp(logw) * x_mask                                                   # :218
    w_ceil = torch.ceil(w) * length_scale                                          # :219
    y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()               # :220
    y_max_length = int(y_lengths.max())                                            # :221 (host sync, as the reference)
    T_y = fix_len_compatibility(y_max_length)                                      # :222
    x_lengths = x_mask.sum(dim=(1, 2)).to(torch.int32)
    r = expand(w_ceil.squeeze(1), mu_x, x_lengths, y_lengths, T_y=T_y, want_attn=True)
    y_mask = (torch.arange(T_y, device=mu_x.device).unsqueeze(0) < y_lengths.unsqueeze(1)).unsqueeze(1).to(x_mask.dtype)
    return {"mu_y": r["mu_y"], "y_mask": y_mask, "attn": r["attn"].unsqueeze(1), "y_lengths": y_lengths,
            "y_max_length": y_max_length}
