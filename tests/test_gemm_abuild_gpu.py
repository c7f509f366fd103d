"""GPU tests of the A-building GEMMs (csrc/gemm_abuild.cuh) in isolation against plain PyTorch fp32 references of the
same chains: GroupNorm(8) + Mish + mask (+ time bias | + res_conv and LayerNorm) followed by the conv / projection,
with the operand rounded to 16 bits where the kernel rounds it (decoder.py:59-66,80-85; transformer.py:316-325)."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def gsums(h):
    B, T, _ = h.shape
    g = h.double().view(B, T, 8, 32)
    return torch.stack([g.sum((1, 3)), (g * g).sum((1, 3))], -1).contiguous()       # [B][8][2]


def gn_mish(h, gamma, beta, mask):
    y = F.group_norm(h.transpose(1, 2), 8, gamma, beta, 1e-5)
    return (F.mish(y) * mask[:, None, :]).transpose(1, 2)


def run(mode, h, gamma, beta, mask, tb, r, ln_g, ln_b, w, N, bias, f16=True):
    from matcha_b200._lib import check, load_library
    lib = load_library()
    B, T, _ = h.shape
    gs = gsums(h)
    x1 = torch.full((B, T, 256), float("nan"), device="cuda") if mode == 1 else None
    out16 = torch.zeros(B, T, N, device="cuda", dtype=w.dtype) if mode == 1 else None
    out32 = torch.full((B, T, N), float("nan"), device="cuda") if mode == 2 else None
    gst = torch.zeros(B, 8, 2, device="cuda", dtype=torch.float64) if mode == 2 else None
    check(lib.matcha_b200_debug_abuild(mode, _p(h), _p(gs), _p(gamma), _p(beta), _p(mask), _p(tb), _p(r), _p(ln_g), _p(ln_b),
                                       _p(x1), _p(w), N, _p(bias), _p(out16), _p(out32), _p(gst), B, T, int(f16),
                                       ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "debug_abuild")
    torch.cuda.synchronize()
    return x1, out16, out32, gst


def rel(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


def make(B, T, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    h = torch.randn(B, T, 256, device="cuda", generator=g) * 1.3 + 0.2
    gamma = 1 + 0.2 * torch.randn(256, device="cuda", generator=g)
    beta = 0.1 * torch.randn(256, device="cuda", generator=g)
    mask = torch.ones(B, T, device="cuda")
    mask[B - 1, (3 * T) // 4:] = 0
    return g, h, gamma, beta, mask


@pytest.mark.parametrize("B,T", [(2, 200), (1, 128), (3, 131), (2, 5), (4, 400)])
def test_qkv_with_groupnorm_mish_res_layernorm(B, T):
    g, h, gamma, beta, mask = make(B, T, 100 + T)
    r = torch.randn(B, T, 256, device="cuda", generator=g)
    ln_g = 1 + 0.2 * torch.randn(256, device="cuda", generator=g)
    ln_b = 0.1 * torch.randn(256, device="cuda", generator=g)
    w = (torch.randn(768, 256, device="cuda", generator=g) * 0.06).half()
    x1, out16, _, _ = run(1, h, gamma, beta, mask, None, r, ln_g, ln_b, w, 768, None)
    want_x1 = gn_mish(h, gamma, beta, mask) + r
    assert rel(x1, want_x1) < 2e-6
    a = F.layer_norm(want_x1, (256,), ln_g, ln_b, 1e-5).half()
    want = a.float() @ w.float().t()
    assert rel(out16.float(), want) < 1e-3


@pytest.mark.parametrize("B,T", [(2, 200), (1, 128), (3, 131), (2, 5), (4, 400), (1, 257)])
def test_conv3_with_groupnorm_mish_time_bias(B, T):
    g, h, gamma, beta, mask = make(B, T, 200 + T)
    tb = torch.randn(1, 256, device="cuda", generator=g) * 0.5
    w = (torch.randn(256, 3, 256, device="cuda", generator=g) * 0.04).half()          # [out][tap][in]
    bias = torch.randn(256, device="cuda", generator=g)
    _, _, out, gst = run(2, h, gamma, beta, mask, tb, None, None, None, w.view(256, 768), 256, bias)
    a = ((gn_mish(h, gamma, beta, mask) + tb[0]) * mask[:, :, None]).half().float()
    want = F.conv1d(a.transpose(1, 2), w.float().permute(0, 2, 1).contiguous(), bias, padding=1).transpose(1, 2)
    assert torch.isfinite(out).all()
    assert rel(out, want) < 2e-4                     # the operand itself is rounded where Mish(GN(.)) differs in the last ulp
    assert rel(gst, gsums(want)) < 1e-4
