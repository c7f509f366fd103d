"""CPU tests of the boundary: the library loads, exports every symbol include/matcha_b200.h
declares, validates arguments, and plans launches -- no compute without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from matcha_b200._lib import LibraryNotBuilt, load_library
    try:
        return load_library()
    except LibraryNotBuilt:
        import __graft_entry__ as g
        g.build()
        return load_library()


def test_header_symbols_all_exported(lib):
    from matcha_b200._lib import DECLARED_SYMBOLS
    header = open(os.path.join(ROOT, "include", "matcha_b200.h")).read()
    declared = set(re.findall(r"\b(matcha_b200_\w+)\s*\(", header))
    assert declared == set(DECLARED_SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None


def test_error_strings_and_version(lib):
    assert lib.matcha_b200_version().decode().count(".") == 2
    assert lib.matcha_b200_error_string(0) == b"ok"
    for code in (-1, -2, -3, -4):
        assert len(lib.matcha_b200_error_string(code)) > 4


def test_argument_validation_needs_no_gpu(lib):
    f = lib.matcha_b200_mas_compute_batch_alignments
    assert f(None, 0, None, None, None, None, 4, 0, 8, ctypes.c_float(-1e9), None, None) == -1
    assert f(None, 0, None, None, None, None, 4, 8, 8, ctypes.c_float(-1e9), None, None) == -1
    assert f(None, 0, None, None, None, None, 0, 8, 8, ctypes.c_float(-1e9), None, None) == 0
    buf = (ctypes.c_int * 8)()
    assert f(buf, 5, buf, None, buf, buf, 1, 1, 2, ctypes.c_float(-1e9), None, None) == -1   # bad dtype
    assert lib.matcha_b200_mas_lengths_from_mask(None, 2, 4, 4, None, None, None) == -1
    assert lib.matcha_b200_mas_maximum_path_host(None, None, None, None, 0, 4, 4, ctypes.c_float(-1e9)) == 0


@pytest.mark.parametrize("shape", [(16, 128, 800), (64, 400, 2000), (8, 50, 200), (4, 20, 64), (2, 450, 520),
                                   (1024, 200, 1000)])
@pytest.mark.parametrize("has_mask", [0, 1])
def test_launch_plan(lib, shape, has_mask):
    out = (ctypes.c_int * 8)()
    assert lib.matcha_b200_mas_query_launch(*shape, has_mask, 1, ctypes.byref(out)) == 0
    block, smem, r, nstage, use_tma, ndp = out[1], out[2], out[3], out[4], out[5], out[7]
    assert r % 2 == 1 and 1 <= r <= 13 and 1 <= nstage <= 8 and (use_tma & 1) == 1
    assert block == 32 * (ndp + 5) and block <= 1024
    assert 32 * r * ndp >= shape[1], "one wavefront pass must cover every token"
    assert smem <= 232448


def test_unsupported_shape_is_reported(lib):
    out = (ctypes.c_int * 8)()
    assert lib.matcha_b200_mas_query_launch(1, 1400, 4000, 0, 1, ctypes.byref(out)) == -2


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from matcha_b200 import monotonic_align
    with pytest.raises(RuntimeError):
        monotonic_align.maximum_path(torch.zeros(1, 2, 3), torch.ones(1, 2, 3))
