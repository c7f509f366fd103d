"""GPU test of the tcgen05/TMEM implicit-GEMM kernel in isolation (csrc/gemm_bf16.cuh) against a plain
PyTorch fp32 reference of the same op on the same bf16-rounded operands."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def run_gemm(a0, a1, taps, w, bias, act, res, mask, want_f32=True, want_bf16=False):
    from matcha_b200._lib import check, load_library
    lib = load_library()
    B, T, c0 = a0.shape
    c1 = a1.shape[2] if a1 is not None else 0
    N = w.shape[0]
    out = torch.full((B, T, N), float("nan"), device="cuda") if want_f32 else None
    outb = torch.zeros((B, T, N), device="cuda", dtype=torch.bfloat16) if want_bf16 else None
    tap_arr = (ctypes.c_int * len(taps))(*taps)
    p = lambda t: t.data_ptr() if t is not None else None
    check(lib.matcha_b200_debug_gemm_bf16(p(a0), c0, p(a1), c1, len(taps), tap_arr, B, T, p(w), N, p(bias), act, p(res),
                                          p(out), p(outb), p(mask), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
          "debug_gemm_bf16")
    torch.cuda.synchronize()
    return out, outb


def reference(a0, a1, taps, w, bias, act, res):
    a = a0.float() if a1 is None else torch.cat([a0.float(), a1.float()], 2)
    B, T, C = a.shape
    cols = []
    for off in taps:
        sh = torch.zeros_like(a)
        lo, hi = max(0, -off), min(T, T - off)
        if hi > lo:
            sh[:, lo:hi] = a[:, lo + off:hi + off]
        cols.append(sh)
    A = torch.cat(cols, 2)                                  # [B, T, ntaps*C]
    y = A @ w.float().t()
    if bias is not None:
        y = y + bias
    if act == 1:
        y = torch.nn.functional.gelu(y)
    if res is not None:
        y = y + res
    return y


def rel(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


@pytest.mark.parametrize("B,T,c0,c1,taps,N", [
    (2, 200, 256, 0, [0], 256), (1, 128, 256, 0, [0], 256), (3, 130, 256, 256, [-1, 0, 1], 256),
    (2, 64, 192, 0, [-1, 0, 1], 256), (2, 333, 256, 0, [0], 768), (2, 100, 1024, 0, [0], 256), (1, 5, 256, 0, [-1, 0, 1], 1024)])
def test_gemm_matches_torch(B, T, c0, c1, taps, N):
    g = torch.Generator(device="cuda").manual_seed(B * 7 + T)
    a0 = torch.randn(B, T, c0, device="cuda", generator=g).bfloat16()
    a1 = torch.randn(B, T, c1, device="cuda", generator=g).bfloat16() if c1 else None
    w = (torch.randn(N, len(taps) * (c0 + c1), device="cuda", generator=g) * 0.05).bfloat16()
    out, _ = run_gemm(a0, a1, taps, w, None, 0, None, None)
    want = reference(a0, a1, taps, w, None, 0, None)
    assert torch.isfinite(out).all()
    assert rel(out, want) < 1e-5


def test_gemm_epilogue_bias_gelu_residual_masked_bf16():
    g = torch.Generator(device="cuda").manual_seed(5)
    B, T, N = 2, 150, 256
    a0 = torch.randn(B, T, 256, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, 256, device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(B, T, N, device="cuda", generator=g)
    mask = torch.ones(B, T, device="cuda"); mask[1, 90:] = 0
    out, outb = run_gemm(a0, None, [0], w, bias, 1, res, mask, want_bf16=True)
    want = reference(a0, None, [0], w, bias, 1, res)
    assert rel(out, want) < 1e-5
    assert torch.equal(outb, (want * mask[:, :, None]).bfloat16()) or rel(outb.float(), (want * mask[:, :, None])) < 4e-3
    assert outb[1, 90:].abs().sum() == 0
