"""GPU tests of the decoder's self-attention kernels (tcgen05 `attention_tc_kernel` and the mma.sync
`attention_mma_kernel`) against a plain torch fp32 reference of the same op on the same 16-bit-rounded
q|k|v rows: softmax(q k^T / 8 + mask01) v per head (diffusers Attention + SDPA with an additive float mask,
transformer.py:251-259, 320-325)."""
import ctypes

import pytest
import torch

from matcha_b200._lib import check, load_library

pytestmark = pytest.mark.gpu
H, D = 4, 64


def run_attention(qkv16, mask, f16, impl):
    lib = load_library()
    B, T, _ = qkv16.shape
    out = torch.full((B, T, H * D), float("nan"), device="cuda", dtype=qkv16.dtype)
    check(lib.matcha_b200_debug_attention(qkv16.data_ptr(), mask.data_ptr(), out.data_ptr(), B, T, int(f16), impl,
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "attention")
    torch.cuda.synchronize()
    return out


def reference(qkv16, mask):
    B, T, _ = qkv16.shape
    q, k, v = (x.float().view(B, T, H, D).transpose(1, 2) for x in qkv16.split(H * D, dim=-1))
    logits = q @ k.transpose(-1, -2) * 0.125 + mask[:, None, None, :]        # additive 0/1 mask
    return (torch.softmax(logits, -1) @ v).transpose(1, 2).reshape(B, T, H * D)


def make(B, T, dtype, seed, scale=1.0, ragged=True):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    qkv = (torch.randn(B, T, 3 * H * D, device="cuda", generator=g) * scale).to(dtype)
    mask = torch.ones(B, T, device="cuda")
    if ragged:
        for b in range(B):
            mask[b, int(T * (0.55 + 0.45 * b / max(B - 1, 1))):] = 0
    return qkv, mask


def rel(a, b):
    return (torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double())).item()


SHAPES = [(2, 800), (3, 400), (2, 100), (1, 130), (2, 128), (1, 257), (2, 16), (1, 1000), (1, 7),
          (32, 800), (40, 400), (80, 130)]          # more tiles than persistent CTAs, ragged tiles included


@pytest.mark.parametrize("impl", [0, 1], ids=["tcgen05", "mma_sync"])
@pytest.mark.parametrize("B,T", SHAPES)
def test_attention_fp16_matches_torch(impl, B, T):
    qkv, mask = make(B, T, torch.float16, 100 + T)
    got, want = run_attention(qkv, mask, True, impl), reference(qkv, mask)
    assert torch.isfinite(got.float()).all()
    assert rel(got.float(), want) < 1.5e-3, (impl, B, T, rel(got.float(), want))
    assert (got.float() - want).abs().max().item() < 2e-2


@pytest.mark.parametrize("impl", [0, 1], ids=["tcgen05", "mma_sync"])
def test_attention_bf16_and_peaked_logits(impl):
    qkv, mask = make(2, 400, torch.bfloat16, 7, scale=3.0)        # logits ~ N(0, 9*8): near one-hot softmax rows
    got, want = run_attention(qkv, mask, False, impl), reference(qkv, mask)
    assert torch.isfinite(got.float()).all() and rel(got.float(), want) < 1e-2


def test_attention_kernels_agree_and_are_deterministic():
    qkv, mask = make(4, 800, torch.float16, 3)
    a, b = run_attention(qkv, mask, True, 0), run_attention(qkv, mask, True, 0)
    assert torch.equal(a, b)
    c = run_attention(qkv, mask, True, 1)
    assert rel(a.float(), c.float()) < 1e-3
