"""matcha_b200_mas_durations_static: the MAS entry point for callers that must not allocate device memory.

Some users of libmatcha_b200.so cannot hand the MAS kernel the [B][T_x][T_y] slab it works in, because they cannot
call cudaMalloc at the moment they need an alignment: a thread running beside a CUDA-graph capture of a decoder solve
(an allocation from any thread breaks a global-mode capture), a bare C / ctypes script on a GPU whose memory a trainer's
caching allocator already holds, a serving loop that keeps allocation (and its device-wide synchronisation) off the
request path.  For them the library carries one fixed 8 MiB slab per GPU and this entry point runs the usual kernels in
it, returning only the per-token durations.

The checks below go through the C ABI with ctypes.  The expected values are the durations of the ordinary entry point
(matcha_b200_mas_align) on the same inputs: integer work, so the comparison is exact, as everywhere else in the MAS
tests.  Calls are made back to back with different shapes because every call reuses the same slab and finds the previous
call's leftovers in it.

This file is named to be collected last: it is the only test of an entry point outside the reference's interface, and a
failure here must not stop `pytest -x` before the parity tests have run.
"""
import ctypes

import numpy as np
import pytest
import torch

from tests.mas_cases import gen

pytestmark = pytest.mark.gpu

PENALTY = ctypes.c_float(-1e9)


@pytest.fixture(scope="module")
def lib():
    assert torch.cuda.is_available()
    from matcha_b200._lib import load_library
    return load_library()


def _case(seed, B, T_x, T_y, ties=False):
    rng = np.random.default_rng(seed)
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    if ties:
        v = np.round(v * 2).astype(np.float32)
    m = gen.ragged_mask(rng, B, T_x, T_y, lo=0.5)
    dev = torch.device("cuda:0")
    t_x = torch.from_numpy(m.sum(1)[:, 0].astype(np.int32)).to(dev)
    t_y = torch.from_numpy(m.sum(2)[:, 0].astype(np.int32)).to(dev)
    return torch.from_numpy(v).to(dev), torch.from_numpy(m).to(dev), t_x, t_y


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _expected(lib, v, mask, t_x, t_y):
    """Durations from the existing entry point, which works in a caller-provided dense path buffer."""
    B, T_x, T_y = v.shape
    paths = torch.empty((B, T_x, T_y), dtype=torch.int32, device=v.device)
    dur = torch.full((B, T_x), -7, dtype=torch.int32, device=v.device)
    rc = lib.matcha_b200_mas_align(paths.data_ptr(), 0, dur.data_ptr(), v.data_ptr(),
                                   mask.data_ptr() if mask is not None else None, t_x.data_ptr(), t_y.data_ptr(),
                                   B, T_x, T_y, PENALTY, None, _stream())
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(paths.sum(-1).to(torch.int32), dur)      # the compact form agrees with the dense path
    return dur


def _static(lib, v, mask, t_x, t_y, dur):
    B, T_x, T_y = v.shape
    return lib.matcha_b200_mas_durations_static(dur.data_ptr(), v.data_ptr(),
                                                mask.data_ptr() if mask is not None else None,
                                                t_x.data_ptr(), t_y.data_ptr(), B, T_x, T_y, PENALTY, None, _stream())


def test_two_calls_in_a_row_match_the_allocating_entry_point(lib):
    # first call: BASELINE config 1's shape (fills 1.6 M of the 2 M cells); second call: a smaller, non-TMA shape
    # (T_y % 4 != 0) with quantised scores, run straight after it in the slab the first call left dirty;
    # third call: the masked kernel variant; fourth: the first inputs again.
    cases = [(_case(11, 16, 128, 800), False), (_case(12, 5, 70, 333, ties=True), False),
             (_case(13, 4, 130, 404), True), (_case(11, 16, 128, 800), False)]
    want = [_expected(lib, v, m if use_mask else None, tx, ty) for (v, m, tx, ty), use_mask in cases]
    got = []
    for (v, m, tx, ty), use_mask in cases:
        dur = torch.full((v.shape[0], v.shape[1]), -7, dtype=torch.int32, device=v.device)
        assert _static(lib, v, m if use_mask else None, tx, ty, dur) == 0
        got.append(dur)                                        # no sync in between: same stream, so the calls are ordered
    torch.cuda.synchronize()
    for i, (g, w) in enumerate(zip(got, want)):
        assert torch.equal(g, w), f"call {i}"
        assert int(g.min()) >= 0
    (v, m, tx, ty), _ = cases[0]
    assert torch.equal(got[0].sum(1).to(torch.int32), ty)       # every valid frame is owned by exactly one token


def test_repeat_call_takes_no_device_memory(lib):
    v, m, tx, ty = _case(21, 8, 100, 600)
    dur = torch.empty((8, 100), dtype=torch.int32, device=v.device)
    assert _static(lib, v, None, tx, ty, dur) == 0               # set-up call: address look-up, kernel image loaded
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert _static(lib, v, None, tx, ty, dur) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    assert torch.equal(dur, _expected(lib, v, None, tx, ty))


def test_size_limit_and_argument_errors(lib):
    cap = lib.matcha_b200_mas_durations_static_max_cells()
    assert cap == 2 * 1024 * 1024
    v, m, tx, ty = _case(31, 2, 16, 64)
    dur = torch.full((2, 16), -7, dtype=torch.int32, device=v.device)
    f = lib.matcha_b200_mas_durations_static
    # one cell over the limit: refused before anything is launched (the pointers are never dereferenced)
    assert f(dur.data_ptr(), v.data_ptr(), None, tx.data_ptr(), ty.data_ptr(), 1, 1, cap + 1, PENALTY, None, None) == -2
    assert f(dur.data_ptr(), v.data_ptr(), None, tx.data_ptr(), ty.data_ptr(), 1024, 400, 2000, PENALTY, None, None) == -2
    assert f(None, v.data_ptr(), None, tx.data_ptr(), ty.data_ptr(), 2, 16, 64, PENALTY, None, None) == -1
    assert f(dur.data_ptr(), v.data_ptr(), None, tx.data_ptr(), ty.data_ptr(), 2, 0, 64, PENALTY, None, None) == -1
    assert f(None, None, None, None, None, 0, 16, 64, PENALTY, None, None) == 0
    torch.cuda.synchronize()
    assert int((dur != -7).sum()) == 0                           # none of the refused calls wrote anything
    assert lib.matcha_b200_error_string(-2).decode()
