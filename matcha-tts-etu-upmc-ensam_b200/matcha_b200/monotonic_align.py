"""Drop-in for ``matcha.utils.monotonic_align`` backed by the sm_100a MAS kernel.

Mirrors /root/reference/matcha/utils/monotonic_align/__init__.py:
  * ``maximum_path(value, mask)``                         (__init__.py:40-55)
  * ``compute_batch_alignments(paths, values, t_xs, t_ys, max_neg_val=-1e9)``
                                                          (core.pyx:101-128, bound as maximum_path_c)
Same argument meaning, same result (bit-exact 0/1 path, ties broken toward token x-1 as in
core.pyx:73).  Differences, all deliberate:
  * the work happens on the GPU and stays asynchronous on the current CUDA stream when the
    inputs are CUDA tensors (the reference synchronises through ``.cpu()``);
  * lengths outside 1 <= t_x <= t_y are undefined behaviour in the reference; here the item's
    path is all zeros (and ``check_lengths=True`` raises);
  * there is no CPU fallback: without the built library or a CUDA device these functions raise.
"""
import ctypes

import numpy as np
import torch

from ._lib import check, load_library

ERR_UNSUPPORTED = -2       # MB200_ERR_UNSUPPORTED (include/matcha_b200.h)

PATH_I32, PATH_F32 = 0, 1


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("matcha_b200.monotonic_align needs a CUDA device (sm_100a); there is no CPU fallback")


def _fused_mask_fits(value):
    out = (ctypes.c_int * 8)()
    B, T_x, T_y = value.shape
    rc = load_library().matcha_b200_mas_query_launch(max(B, 1), T_x, T_y, 1, 1, ctypes.byref(out))
    if rc == 0:
        return True
    if rc == ERR_UNSUPPORTED:          # the fused-mask plan does not fit on chip: the caller forms value*mask first
        return False
    check(rc, "mas_query_launch")      # any other code is a real error, not "does not fit"
    return False


def lengths_from_mask(mask):
    """t_x[b] = int(sum_x mask[b,x,0]), t_y[b] = int(sum_y mask[b,0,y]) on the device (__init__.py:52-53)."""
    lib = load_library()
    assert mask.is_cuda and mask.dim() == 3
    m = mask if (mask.dtype == torch.float32 and mask.is_contiguous()) else mask.float().contiguous()
    B, T_x, T_y = m.shape
    t = torch.empty((2, max(B, 1)), dtype=torch.int32, device=m.device)
    with torch.cuda.device(m.device):
        check(lib.matcha_b200_mas_lengths_from_mask(m.data_ptr(), B, T_x, T_y, t[0].data_ptr(), t[1].data_ptr(),
                                                    _stream_ptr(m.device)), "mas_lengths_from_mask")
    return t[0, :B], t[1, :B]


def maximum_path_from_lengths(value, t_xs, t_ys, mask=None, out_dtype=torch.float32, max_neg_val=-1e9,
                              check_lengths=False, out=None):
    """Device-level entry: fp32 ``value`` [B,T_x,T_y] (CUDA, contiguous), int32 lengths [B] (CUDA).
    ``mask`` (fp32, same shape) multiplies the scores inside the kernel when given.
    Returns the alignment as float32 or int32, every element written by the kernel."""
    lib = load_library()
    assert value.is_cuda and value.dtype == torch.float32 and value.is_contiguous() and value.dim() == 3
    assert t_xs.dtype == torch.int32 and t_ys.dtype == torch.int32 and t_xs.is_cuda and t_ys.is_cuda
    assert out_dtype in (torch.float32, torch.int32)
    B, T_x, T_y = value.shape
    if out is not None:
        assert out.is_cuda and out.is_contiguous() and out.shape == value.shape and out.dtype == out_dtype
        path = out
    else:
        path = torch.empty((B, T_x, T_y), dtype=out_dtype, device=value.device)
    if B == 0 or T_x == 0 or T_y == 0:
        return path.zero_()
    mptr = None
    if mask is not None:
        assert mask.is_cuda and mask.dtype == torch.float32 and mask.is_contiguous() and mask.shape == value.shape
        mptr = mask.data_ptr()
    status = torch.zeros(1, dtype=torch.int32, device=value.device) if check_lengths else None
    t_xs, t_ys = t_xs.contiguous(), t_ys.contiguous()      # named: a temporary would be freed before the kernel reads it
    with torch.cuda.device(value.device):
        check(lib.matcha_b200_mas_compute_batch_alignments(
            path.data_ptr(), PATH_F32 if out_dtype == torch.float32 else PATH_I32,
            value.data_ptr(), mptr, t_xs.data_ptr(), t_ys.data_ptr(),
            B, T_x, T_y, ctypes.c_float(max_neg_val),
            status.data_ptr() if status is not None else None, _stream_ptr(value.device)),
            "mas_compute_batch_alignments")
    if status is not None:
        bad = int(status.item())
        if bad:
            raise ValueError(f"maximum_path: {bad} item(s) violate 1 <= t_x <= t_y (undefined in the reference)")
    return path


def maximum_path(value, mask, mask_mode="full", check_lengths=False):
    """Same contract as the reference ``maximum_path`` (__init__.py:40-55).

    value: [b, t_x, t_y] log-likelihoods; mask: [b, t_x, t_y].  Returns the 0/1 alignment with
    ``value``'s dtype on ``value``'s device.

    mask_mode="full"   scores are ``value*mask`` for an arbitrary mask, as the reference computes.
    mask_mode="prefix" the caller asserts the mask is the outer product of two prefix masks
                       (what matcha_tts.py:271 builds); only its first row/column are read, which
                       saves one full pass over HBM.
    """
    _require_cuda()
    assert value.dim() == 3 and mask.shape == value.shape
    assert mask_mode in ("full", "prefix")
    in_device, in_dtype = value.device, value.dtype
    dev = in_device if value.is_cuda else torch.device("cuda", torch.cuda.current_device())
    v = value.detach().to(dev)
    m = mask.detach().to(dev)
    t_xs, t_ys = lengths_from_mask(m)
    kernel_mask = None
    if v.dtype == torch.float32 and m.dtype == torch.float32:
        v = v.contiguous()
        if mask_mode == "full":
            kernel_mask = m.contiguous()
            if not _fused_mask_fits(v):
                # shape too large for the fused-mask kernel variant: form value*mask first
                scores = torch.empty_like(v)
                with torch.cuda.device(dev):
                    check(load_library().matcha_b200_mas_apply_mask(
                        scores.data_ptr(), v.data_ptr(), kernel_mask.data_ptr(), v.numel(), _stream_ptr(dev)),
                        "mas_apply_mask")
                v, kernel_mask = scores, None
    else:
        # off the hot path: form value*mask in the caller's dtype exactly as __init__.py:45,48 do
        v = ((v * m) if mask_mode == "full" else v).float().contiguous()
    path = maximum_path_from_lengths(v, t_xs, t_ys, mask=kernel_mask, out_dtype=torch.float32,
                                     check_lengths=check_lengths)
    return path.to(device=in_device, dtype=in_dtype)


def compute_batch_alignments(paths, values, t_xs, t_ys, max_neg_val=-1e9):
    """core.pyx:101-107 with the same argument order.  Accepts either numpy arrays (host buffers:
    copied to the device and back inside the call through a sliced H2D / kernel / D2H pipeline,
    synchronous, like the reference; ``paths`` must be zero-initialised by the caller exactly as the
    reference requires, __init__.py:49) or CUDA tensors (asynchronous on the current stream, every
    element of ``paths`` written).  ``paths`` is int32 (or float32 for CUDA tensors); ``values`` is
    left untouched (the reference clobbers it)."""
    _require_cuda()
    lib = load_library()
    if isinstance(values, np.ndarray):
        assert paths.dtype == np.int32 and values.dtype == np.float32
        assert paths.flags.c_contiguous and values.flags.c_contiguous and paths.shape == values.shape
        t_xs = np.ascontiguousarray(t_xs, dtype=np.int32)
        t_ys = np.ascontiguousarray(t_ys, dtype=np.int32)
        B, T_x, T_y = values.shape
        return check(lib.matcha_b200_mas_maximum_path_host(
            paths.ctypes.data, values.ctypes.data, t_xs.ctypes.data, t_ys.ctypes.data,
            B, T_x, T_y, ctypes.c_float(max_neg_val)), "mas_maximum_path_host")
    assert values.is_cuda and paths.is_cuda and paths.shape == values.shape and paths.is_contiguous()
    assert paths.dtype in (torch.int32, torch.float32)
    maximum_path_from_lengths(values, t_xs.to(values.device, torch.int32), t_ys.to(values.device, torch.int32),
                              out_dtype=paths.dtype, max_neg_val=max_neg_val, out=paths)
    return 0


# the reference binds the core under this name (__init__.py:5)
maximum_path_c = compute_batch_alignments
