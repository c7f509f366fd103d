"""oracle/mas.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Python access to the two CPU checkers for monotonic alignment search:

* ``oracle_*``  : the plain-C restatement in ``oracle/mas_oracle.c`` (built to
  ``oracle/libmas_oracle.so``), loaded with ctypes.
* ``ref_*``     : the reference's own ``core.pyx`` compiled into ``oracle/_ref/`` by
  ``oracle/Makefile`` (``compute_batch_alignments``, core.pyx:101-128).

``maximum_path_*`` restate the reference wrapper
(/root/reference/matcha/utils/monotonic_align/__init__.py:40-55) on numpy arrays.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  Parity pinned: see tests/test_oracle_mas.py.
"""
import ctypes
import glob
import importlib.util
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_ref_core = None


def _oracle_lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libmas_oracle.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run __graft_entry__.build() or `make -C oracle`")
        lib = ctypes.CDLL(path)
        lib.mas_oracle_batch.restype = ctypes.c_int
        lib.mas_oracle_batch.argtypes = [
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int]
        lib.mas_oracle_max_threads.restype = ctypes.c_int
        _lib = lib
    return _lib


def oracle_max_threads():
    return int(_oracle_lib().mas_oracle_max_threads())


def oracle_compute_batch_alignments(paths, values, t_xs, t_ys, max_neg_val=-1e9, n_threads=0):
    """Same contract as core.pyx:101-107: int32 paths (pre-zeroed), fp32 values (clobbered
    with cumulative scores), int32 lengths.  Returns the number of out-of-domain items."""
    assert paths.dtype == np.int32 and values.dtype == np.float32
    assert paths.flags.c_contiguous and values.flags.c_contiguous and paths.shape == values.shape
    t_xs = np.ascontiguousarray(t_xs, dtype=np.int32)
    t_ys = np.ascontiguousarray(t_ys, dtype=np.int32)
    B, T_x, T_y = values.shape
    return int(_oracle_lib().mas_oracle_batch(
        paths.ctypes.data, values.ctypes.data, t_xs.ctypes.data, t_ys.ctypes.data,
        B, T_x, T_y, ctypes.c_float(max_neg_val), int(n_threads)))


def ref_available():
    return bool(glob.glob(os.path.join(_HERE, "_ref", "core*.so")))


def ref_compute_batch_alignments(paths, values, t_xs, t_ys, max_neg_val=-1e9):
    """The reference's compiled core.pyx, called exactly as __init__.py:54 calls it."""
    global _ref_core
    if _ref_core is None:
        hits = glob.glob(os.path.join(_HERE, "_ref", "core*.so"))
        if not hits:
            raise RuntimeError("oracle/_ref/core*.so missing: run `make -C oracle` where /root/reference exists")
        spec = importlib.util.spec_from_file_location("core", hits[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _ref_core = mod
    _ref_core.compute_batch_alignments(paths, values, t_xs, t_ys, max_neg_val)


def set_omp_threads(n):
    """Sets the OpenMP team size the compiled core.pyx (prange, core.pyx:121) and the C restatement will use, on the
    libgomp the process has loaded, and returns the value now in force.  Needed because launchers such as
    torch.distributed.run export OMP_NUM_THREADS=1 to every rank."""
    os.environ["OMP_NUM_THREADS"] = str(int(n))
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(int(n))
        return int(gomp.omp_get_max_threads())
    except OSError:
        return int(n)


def maximum_path_ref_torch(value, mask):
    """The reference WRAPPER on torch tensors, restated line by line from
    /root/reference/matcha/utils/monotonic_align/__init__.py:40-55 (value*mask, .cpu().numpy() copies, np.zeros paths,
    the two mask reductions, the compiled core, torch.from_numpy back to the caller's device and dtype)."""
    import torch
    value = value * mask                                                        # :45
    device, dtype = value.device, value.dtype                                   # :46-47
    value = value.data.cpu().numpy().astype(np.float32)                         # :48
    path = np.zeros_like(value).astype(np.int32)                                # :49
    mask = mask.data.cpu().numpy()                                              # :50
    t_x_max = mask.sum(1)[:, 0].astype(np.int32)                                # :52
    t_y_max = mask.sum(2)[:, 0].astype(np.int32)                                # :53
    ref_compute_batch_alignments(path, value, t_x_max, t_y_max)                 # :54
    return torch.from_numpy(path).to(device=device, dtype=dtype)                # :55


def _lengths_from_mask(mask):
    # __init__.py:52-53
    t_x = mask.sum(1)[:, 0].astype(np.int32)
    t_y = mask.sum(2)[:, 0].astype(np.int32)
    return t_x, t_y


def maximum_path_ref(value, mask):
    """__init__.py:40-55 on numpy arrays, backed by the compiled reference core."""
    value = (value * mask).astype(np.float32)
    path = np.zeros_like(value).astype(np.int32)
    t_x, t_y = _lengths_from_mask(mask)
    ref_compute_batch_alignments(path, value, t_x, t_y)
    return path


def maximum_path_oracle(value, mask, n_threads=0):
    """Same wrapper, backed by the C restatement."""
    value = np.ascontiguousarray((value * mask).astype(np.float32))
    path = np.zeros(value.shape, dtype=np.int32)
    t_x, t_y = _lengths_from_mask(mask)
    oracle_compute_batch_alignments(path, value, t_x, t_y, n_threads=n_threads)
    return path
