"""GPU test of the fused feed-forward kernel (csrc/ffn_fused.cuh) in isolation against a plain PyTorch fp32
reference of the same op (transformer.py:105-188 + the residual of :366) on the same 16-bit-rounded operands.
Tolerances: the hidden activation is rounded to 16 bits inside the kernel (rel 2^-11 for fp16, 2^-8 for bf16) and
the GELU carries tanh.approx's 2^-11 relative error, so the FFN term is good to ~5e-4 (fp16) / 4e-3 (bf16)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def run_ffn(a, w1, b1, w2, b2, res, mask, f16, want_f32=True, want_16=True, ln=None):
    from matcha_b200._lib import check, load_library
    lib = load_library()
    B, T, _ = a.shape
    out = torch.full((B, T, 256), float("nan"), device="cuda") if want_f32 else None
    out16 = torch.zeros((B, T, 256), device="cuda", dtype=a.dtype) if want_16 else None
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    ln_in, ln_g, ln_b = ln if ln is not None else (None, None, None)
    check(lib.matcha_b200_debug_ffn(p(a), p(w1), p(b1), p(w2), p(b2), p(res), p(out), p(out16), p(mask), B, T, int(f16),
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), p(ln_in), p(ln_g), p(ln_b)), "debug_ffn")
    torch.cuda.synchronize()
    return out, out16


def reference(a, w1, b1, w2, b2, res):
    h = torch.nn.functional.gelu(a.float() @ w1.float().t() + b1)
    return res + h @ w2.float().t() + b2, h @ w2.float().t() + b2


def rel(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


@pytest.mark.parametrize("B,T,f16", [(2, 200, True), (1, 128, True), (3, 130, True), (2, 5, True), (4, 400, True),
                                     (2, 200, False), (1, 333, False)])
def test_ffn_matches_torch(B, T, f16):
    dt = torch.float16 if f16 else torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(B * 11 + T)
    a = torch.randn(B, T, 256, device="cuda", generator=g).to(dt)
    w1 = (torch.randn(1024, 256, device="cuda", generator=g) * 0.09).to(dt)      # Kaiming scale: pre-activations ~ N(0, 2)
    w2 = (torch.randn(256, 1024, device="cuda", generator=g) * 0.044).to(dt)
    b1 = torch.randn(1024, device="cuda", generator=g) * 0.5
    b2 = torch.randn(256, device="cuda", generator=g)
    res = torch.randn(B, T, 256, device="cuda", generator=g)
    mask = torch.ones(B, T, device="cuda")
    mask[B - 1, (2 * T) // 3:] = 0
    out, out16 = run_ffn(a, w1, b1, w2, b2, res, mask, f16)
    want, want_term = reference(a, w1, b1, w2, b2, res)
    assert torch.isfinite(out).all()
    tol = 6e-4 if f16 else 4e-3
    assert rel(out - res, want_term) < tol                    # the feed-forward term itself
    assert rel(out, want) < tol
    assert rel(out16.float(), want * mask[:, :, None]) < (1e-3 if f16 else 6e-3)
    assert out16[B - 1, (2 * T) // 3:].abs().sum() == 0


def test_ffn_only_16bit_output_and_large_activations():
    g = torch.Generator(device="cuda").manual_seed(3)
    B, T = 2, 257
    a = (torch.randn(B, T, 256, device="cuda", generator=g) * 3).half()
    w1 = (torch.randn(1024, 256, device="cuda", generator=g) * 0.09).half()
    w2 = (torch.randn(256, 1024, device="cuda", generator=g) * 0.044).half()
    b1 = torch.zeros(1024, device="cuda")
    b2 = torch.zeros(256, device="cuda")
    res = torch.zeros(B, T, 256, device="cuda")
    out, out16 = run_ffn(a, w1, b1, w2, b2, res, None, True, want_f32=False)
    want, _ = reference(a, w1, b1, w2, b2, res)
    assert out is None
    assert rel(out16.float(), want) < 1e-3


@pytest.mark.parametrize("B,T", [(2, 200), (1, 131), (3, 64)])
def test_ffn_with_layernorm_folded_in(B, T):
    """norm3 + FeedForward + residual (transformer.py:352-366): the kernel builds its A operand from the fp32 rows."""
    g = torch.Generator(device="cuda").manual_seed(T)
    x = torch.randn(B, T, 256, device="cuda", generator=g) * 1.7 + 0.3
    ln_g = 1 + 0.2 * torch.randn(256, device="cuda", generator=g)
    ln_b = 0.1 * torch.randn(256, device="cuda", generator=g)
    w1 = (torch.randn(1024, 256, device="cuda", generator=g) * 0.09).half()
    w2 = (torch.randn(256, 1024, device="cuda", generator=g) * 0.044).half()
    b1 = torch.randn(1024, device="cuda", generator=g) * 0.5
    b2 = torch.randn(256, device="cuda", generator=g)
    dummy = torch.zeros(B, T, 256, device="cuda", dtype=torch.float16)
    out, _ = run_ffn(dummy, w1, b1, w2, b2, x, None, True, want_16=False, ln=(x, ln_g, ln_b))
    a = torch.nn.functional.layer_norm(x, (256,), ln_g, ln_b, 1e-5)
    want, want_term = reference(a.half(), w1, b1, w2, b2, x)
    assert rel(out - x, want_term) < 8e-4
    assert rel(out, want) < 6e-4


# ------------------------------------------------------------------ OUTPROJ variant: the whole tail of the transformer block
def run_outproj(att, wo, bo, x1, ln_g, ln_b, w1, b1, w2, b2, mask, f16):
    from matcha_b200._lib import check, load_library
    lib = load_library()
    B, T, _ = att.shape
    out = torch.full((B, T, 256), float("nan"), device="cuda")
    out16 = torch.zeros((B, T, 256), device="cuda", dtype=att.dtype)
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    lib.matcha_b200_debug_ffn_outproj.restype = ctypes.c_int
    check(lib.matcha_b200_debug_ffn_outproj(p(att), p(wo), p(bo), p(x1), p(ln_g), p(ln_b), p(w1), p(b1), p(w2), p(b2), p(out), p(out16),
                                            p(mask), B, T, int(f16), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
          "debug_ffn_outproj")
    torch.cuda.synchronize()
    return out, out16


@pytest.mark.parametrize("B,T,f16", [(2, 200, True), (1, 128, True), (3, 130, True), (2, 5, True), (4, 400, True), (8, 800, True),
                                     (2, 200, False), (1, 333, False)])
def test_outproj_norm3_ffn_matches_torch(B, T, f16):
    """transformer.py:320-368 after the attention core: x2 = x1 + to_out[0](att); out = x2 + FF(norm3(x2)), against
    plain torch fp32 on the same 16-bit-rounded att / weights.  The row means are offset on purpose (LayerNorm's
    cancellation case)."""
    dt = torch.float16 if f16 else torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(B * 13 + T)
    att = torch.randn(B, T, 256, device="cuda", generator=g).to(dt)
    wo = (torch.randn(256, 256, device="cuda", generator=g) * 0.09).to(dt)
    bo = torch.randn(256, device="cuda", generator=g) * 0.3
    x1 = torch.randn(B, T, 256, device="cuda", generator=g) * 2.0 + 3.0 * torch.randn(B, T, 1, device="cuda", generator=g)
    ln_g = 1.0 + 0.1 * torch.randn(256, device="cuda", generator=g)
    ln_b = 0.1 * torch.randn(256, device="cuda", generator=g)
    w1 = (torch.randn(1024, 256, device="cuda", generator=g) * 0.09).to(dt)
    w2 = (torch.randn(256, 1024, device="cuda", generator=g) * 0.044).to(dt)
    b1 = torch.randn(1024, device="cuda", generator=g) * 0.5
    b2 = torch.randn(256, device="cuda", generator=g)
    mask = torch.ones(B, T, device="cuda")
    mask[B - 1, (2 * T) // 3:] = 0
    out, out16 = run_outproj(att, wo, bo, x1, ln_g, ln_b, w1, b1, w2, b2, mask, f16)
    x2 = x1 + att.float() @ wo.float().t() + bo
    a = torch.nn.functional.layer_norm(x2, (256,), ln_g, ln_b, eps=1e-5)
    a16 = a.to(dt).float()                                    # the kernel rounds norm3's output to the operand format
    h = torch.nn.functional.gelu(a16 @ w1.float().t() + b1)
    want = x2 + h @ w2.float().t() + b2
    assert torch.isfinite(out).all()
    tol = 6e-4 if f16 else 4e-3
    assert rel(out, want) < tol
    assert rel(out - x2, want - x2) < (2e-3 if f16 else 1.5e-2)       # the feed-forward term alone
    assert rel(out16.float(), want * mask[:, :, None]) < (1e-3 if f16 else 6e-3)
    assert out16[B - 1, (2 * T) // 3:].abs().sum() == 0
