"""GPU parity tests of the text encoder (csrc/text_encoder.cu through matcha_b200.text_encoder.TextEncoder):
against the golden outputs of the reference's own TextEncoder (tests/golden/text_encoder_golden.npz) and against
the torch restatement in fp64 on the same seeded inputs, at the bench shape and at tile-edge shapes.
Tolerance: rel-L2 <= 1e-5 in fp32 mode (the north_star's fp32 bar), <= 2e-3 in the fp16 tensor-core mode."""
import os

import numpy as np
import pytest
import torch

from oracle import text_encoder as oenc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = {"fp32": 1e-5, "fp16": 2e-3}


def rel_l2(a, b):
    return float(torch.linalg.norm((a.double().cpu() - b.double().cpu())) / torch.linalg.norm(b.double().cpu()))


def close_logw(got, want, precision):
    """log-durations: rel-L2 under the mode's bar; a batch with only a handful of valid tokens has a handful of O(0.1)
    numbers in it, for which the fp16 mode is held to an absolute 5e-3 instead (the fp32 bar stays relative)."""
    if rel_l2(got, want) < TOL[precision]:
        return True
    return precision == "fp16" and float((got.double().cpu() - want.double().cpu()).abs().max()) < 5e-3


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "text_encoder_golden.npz"))


@pytest.fixture(scope="module", params=["fp32", "fp16"])
def encoder(golden, request):
    from matcha_b200.text_encoder import TextEncoder
    n_vocab = int(golden["n_vocab"][0])
    m = TextEncoder(n_vocab=n_vocab, precision=request.param)
    m.load_state_dict(oenc.make_weights(int(golden["weight_seed"][0]), "rich", n_vocab))
    return m


@pytest.fixture(scope="module")
def weights64(golden):
    w = oenc.make_weights(int(golden["weight_seed"][0]), "rich", int(golden["n_vocab"][0]))
    return oenc.cast_weights({k: v.cuda() for k, v in w.items()}, torch.float64)


@pytest.mark.parametrize("name", ["ragged_T50", "full_T33", "one_token", "long_T160"])
def test_matches_reference_golden(golden, encoder, name):
    ids = torch.from_numpy(golden[name + "/ids"]).cuda()
    lengths = torch.from_numpy(golden[name + "/lengths"]).cuda()
    encoder.use_graph = False
    mu, logw, mask = encoder(ids, lengths)
    assert mu.dtype == torch.float32 and mu.is_cuda and mu.shape == golden[name + "/mu"].shape
    assert torch.equal(mask.cpu(), torch.from_numpy(golden[name + "/mask"]))
    assert rel_l2(mu, torch.from_numpy(golden[name + "/mu"])) < TOL[encoder.precision]
    assert close_logw(logw, torch.from_numpy(golden[name + "/logw"]), encoder.precision)
    pad = mask[:, 0, :] == 0
    assert float(mu.transpose(1, 2)[pad].abs().sum()) == 0.0 and float(logw[:, 0, :][pad].abs().sum()) == 0.0


@pytest.mark.parametrize("B,T,frac", [(32, 128, 0.7), (3, 31, 0.5), (2, 257, 0.6), (5, 64, 1.0), (1, 1, 1.0), (4, 400, 0.7)])
def test_matches_fp64_oracle(encoder, weights64, B, T, frac):
    """Bench shape (32 x 128 ragged), rows not a multiple of the 32-row tile, several attention key blocks, a single token,
    and T_text = 400 (BASELINE configs[1]'s upper end)."""
    g = torch.Generator().manual_seed(B * 1000 + T)
    ids = torch.randint(0, encoder.vocab_size, (B, T), generator=g).cuda()
    lens = torch.tensor([T] + [max(1, int(T * (frac + (1 - frac) * float(torch.rand((), generator=g))))) for _ in range(B - 1)])
    encoder.use_graph = False
    mu, logw, mask = encoder(ids, lens.cuda())
    omu, ologw, omask = oenc.text_encoder_forward(weights64, ids, lens.cuda())
    assert torch.equal(mask.double(), omask)
    assert rel_l2(mu, omu) < TOL[encoder.precision]
    assert close_logw(logw, ologw, encoder.precision)


def test_graph_replay_equals_plain_launches(encoder):
    g = torch.Generator().manual_seed(9)
    B, T = 6, 70
    encoder.use_graph = False
    batches = []
    for _ in range(3):
        ids = torch.randint(0, encoder.vocab_size, (B, T), generator=g).cuda()
        lens = torch.tensor([T, 60, 50, 44, 33, 7]).cuda()
        batches.append((ids, lens, encoder(ids, lens)))
    encoder.use_graph = True
    encoder._graphs.clear(); encoder._seen.clear()
    for ids, lens, want in batches:                     # 1st call plain, 2nd captures, 3rd replays
        got = encoder(ids, lens)
        assert all(torch.equal(a, b) for a, b in zip(got, want))
    assert len(encoder._graphs) == 1
    encoder.use_graph = False


def test_int32_ids_and_cpu_lengths(encoder):
    g = torch.Generator().manual_seed(4)
    ids = torch.randint(0, encoder.vocab_size, (2, 20), generator=g)
    a = encoder(ids.cuda(), torch.tensor([20, 11]).cuda())
    b = encoder(ids.int().cuda(), torch.tensor([20, 11]))          # lengths on the host, int32 ids
    assert all(torch.equal(x, y) for x, y in zip(a, b))


def test_check_inputs_raises_like_the_reference_embedding(encoder):
    # nn.Embedding raises IndexError on an id outside [0, n_vocab) (text_encoder.py:377-379); the kernel clamps, so the
    # Python mirror validates when asked to (check_inputs / MB200_CHECK_INPUTS=1)
    ids = torch.zeros(2, 12, dtype=torch.int64)
    lens = torch.tensor([12, 7])
    encoder.check_inputs = True
    try:
        encoder(ids.cuda(), lens)                                       # in range: passes
        bad = ids.clone(); bad[1, 3] = encoder.vocab_size
        with pytest.raises(IndexError):
            encoder(bad.cuda(), lens)
        bad[1, 3] = -1
        with pytest.raises(IndexError):
            encoder(bad.cuda(), lens)
        with pytest.raises(ValueError):
            encoder(ids.cuda(), torch.tensor([13, 7]))
    finally:
        encoder.check_inputs = False
    out = encoder(bad.cuda(), lens)                                     # default: clamped, no error (documented deviation)
    assert all(bool(torch.isfinite(o).all()) for o in out)


def test_c_abi_argument_checks(encoder):
    import ctypes
    from matcha_b200._lib import load_library
    lib = load_library()
    ws = torch.empty(int(lib.matcha_b200_text_encoder_workspace_bytes(1, 8, 0)), dtype=torch.uint8, device="cuda")
    assert lib.matcha_b200_text_encoder_forward(None, None, None, None, 178, None, None, None, None, None, 1, 8, 0, ws.data_ptr(), ws.numel(), None) == -1
    assert lib.matcha_b200_text_encoder_workspace_bytes(0, 8, 0) == -1
    assert lib.matcha_b200_text_encoder_workspace_bytes(1, 8, 1) == -1          # bf16 is not a text-encoder mode
    assert lib.matcha_b200_text_encoder_blob_floats(178) == sum(int(np.prod(s)) for s in oenc.param_shapes(178).values()) + 3


# ---------------------------------------------------------------------------------------------------------------------
# kernel level: the tcgen05 implicit GEMM (te_gemm_tc_kernel) against torch in fp64 on the fp16-rounded operands
def _P(t):
    import ctypes
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("B,T", [(2, 50), (3, 128), (2, 300), (1, 1)])
@pytest.mark.parametrize("case", ["prenet_conv", "prenet_proj", "qkv", "outproj_ln", "ffn1", "ffn2_ln", "dp1", "dp2_proj"])
def test_tc_gemm_kernel(B, T, case):
    import torch.nn.functional as F
    from matcha_b200._lib import load_library
    lib = load_library()
    lib.matcha_b200_debug_te_gemm_tc.restype = __import__("ctypes").c_int
    #            Cin  N   taps relu_pre mask_pre res   ln    relu_post mask_post f32   f16   mask16 dp
    cfg = {"prenet_conv": (192, 192, 5, 0, 0, False, True, 1, 0, False, True, 1, False),
           "prenet_proj": (192, 192, 1, 0, 0, True, False, 0, 1, True, True, 0, False),
           "qkv": (192, 576, 1, 0, 0, False, False, 0, 0, True, False, 0, False),
           "outproj_ln": (192, 192, 1, 0, 0, True, True, 0, 0, True, True, 1, False),
           "ffn1": (192, 768, 3, 1, 0, False, False, 0, 0, False, True, 0, False),
           "ffn2_ln": (768, 192, 3, 0, 1, True, True, 0, 1, True, True, 0, False),
           "dp1": (192, 256, 3, 1, 0, False, True, 0, 0, False, True, 1, False),
           "dp2_proj": (256, 256, 3, 1, 0, False, True, 0, 0, True, False, 0, True)}[case]
    Cin, N, taps, relu_pre, mask_pre, use_res, ln, relu_post, mask_post, want32, want16, mask16, dp = cfg
    g = torch.Generator().manual_seed(B * 100 + T + Cin)
    dev = "cuda"
    lens = torch.tensor([T] + [max(1, int(T * 0.6))] * (B - 1))
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    a16 = torch.randn(B, T, Cin, generator=g).to(dev).half()
    w16 = (torch.randn(N, taps, Cin, generator=g) / (taps * Cin) ** 0.5).to(dev).half()
    bias = (0.1 * torch.randn(N, generator=g)).to(dev)
    res = torch.randn(B, T, N, generator=g).to(dev) if use_res else None
    lg = (1 + 0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    lb = (0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    dpw = torch.randn(N, generator=g).to(dev) if dp else None
    dpb = torch.randn(4, generator=g).to(dev) if dp else None
    out32 = torch.full((B, T, N), float("nan"), device=dev) if want32 else None
    out16 = torch.full((B, T, N), float("nan"), device=dev, dtype=torch.float16) if want16 else None
    dpo = torch.full((B, T), float("nan"), device=dev) if dp else None
    rc = lib.matcha_b200_debug_te_gemm_tc(_P(a16), Cin, _P(mask), taps, B, T, _P(w16), _P(bias), N, relu_pre, mask_pre, _P(res), _P(lg), _P(lb),
                                          relu_post, mask_post, _P(out32), _P(out16), mask16, _P(dpw), _P(dpb), _P(dpo), None, None, None, None)
    torch.cuda.synchronize()
    assert rc == 1
    v = F.conv1d(a16.double().transpose(1, 2), w16.double().permute(0, 2, 1), bias.double(), padding=taps // 2).transpose(1, 2)
    if relu_pre: v = torch.relu(v)
    if mask_pre: v = v * mask[..., None]
    if use_res: v = v + res.double()
    if ln: v = F.layer_norm(v, (N,), lg.double(), lb.double(), eps=1e-5)
    if relu_post: v = torch.relu(v)
    if mask_post: v = v * mask[..., None]
    if want32:
        assert rel_l2(out32, v) < 2e-6               # same fp16 operands, fp32 accumulation: only summation order differs
    if want16:
        v16 = v * mask[..., None] if mask16 else v
        assert rel_l2(out16.float(), v16) < 6e-4     # + one rounding to fp16
    if dp:
        d = ((v * mask[..., None]) @ dpw.double() + dpb[0].double()) * mask
        assert rel_l2(dpo, d) < 2e-6


def test_tc_gemm_rope_epilogue_and_attention():
    """The q|k|v projection with RoPE in its epilogue followed by attention with rope_done = 1 (the fp16-mode pair) against
    torch in fp64 on the same fp16-rounded operands."""
    import ctypes
    from matcha_b200._lib import load_library
    from matcha_b200.text_encoder import rope_tables
    lib = load_library()
    B, T = 3, 150
    g = torch.Generator().manual_seed(12)
    dev = "cuda"
    lens = torch.tensor([T, 97, 33])
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    a16 = torch.randn(B, T, 192, generator=g).to(dev).half()
    w16 = (torch.randn(576, 192, generator=g) / 192 ** 0.5).to(dev).half()
    bias = (0.1 * torch.randn(576, generator=g)).to(dev)
    cos, sin = (t.to(dev) for t in rope_tables(256))
    qkv = torch.full((B, T, 576), float("nan"), device=dev)
    rc = lib.matcha_b200_debug_te_gemm_tc(_P(a16), 192, _P(mask), 1, B, T, _P(w16), _P(bias), 576, 0, 0, None, None, None, 0, 0, _P(qkv), None, 0,
                                          None, None, None, _P(cos), _P(sin), None, None)
    torch.cuda.synchronize()
    assert rc == 1
    ref = a16.double() @ w16.double().t() + bias.double()
    q, k, v = (ref[..., i * 192:(i + 1) * 192].view(B, T, 2, 96).transpose(1, 2) for i in range(3))
    q, k = oenc.rope(q), oenc.rope(k)
    want = torch.cat([t_.transpose(1, 2).reshape(B, T, 192) for t_ in (q, k, v)], -1)
    assert rel_l2(qkv, want) < 2e-6


@pytest.mark.parametrize("B,T", [(2, 50), (32, 128), (1, 1), (3, 33), (2, 257)])
def test_attention_kernel(B, T):
    """te_attention_kernel on its own (RoPE inside, the fp32-mode form) against torch in fp64."""
    from matcha_b200._lib import load_library
    lib = load_library()
    g = torch.Generator().manual_seed(B + T)
    dev = "cuda"
    qkv = torch.randn(B, T, 576, generator=g).to(dev)
    lens = torch.tensor([T] + [max(1, int(T * 0.6))] * (B - 1))
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    cos, sin = (t.contiguous() for t in oenc.rope_tables(max(T, 8), dev))
    out = torch.full((B, T, 192), float("nan"), device=dev)
    assert lib.matcha_b200_debug_te_attention(_P(qkv), _P(mask), _P(cos), _P(sin), _P(out), B, T, None) == 1
    torch.cuda.synchronize()
    d = qkv.double()
    q, k, v = (d[..., i * 192:(i + 1) * 192].view(B, T, 2, 96).transpose(1, 2) for i in range(3))
    sc = oenc.rope(q) @ oenc.rope(k).transpose(-1, -2) / 96 ** 0.5
    sc = sc.masked_fill((mask[:, None, :, None] * mask[:, None, None, :]) == 0, -1e4)
    want = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B, T, 192)
    assert rel_l2(out, want) < 2e-6


@pytest.mark.parametrize("case", ["prenet_conv", "prenet_proj", "qkv", "outproj_ln", "ffn1", "ffn2_ln", "dp1", "dp2_proj"])
def test_simt_gemm_kernel(case):
    """te_gemm_kernel (fp32 SIMT implicit GEMM with its epilogues) against torch in fp64."""
    import torch.nn.functional as F
    from matcha_b200._lib import load_library
    lib = load_library()
    #            Cin  N   taps mask_in relu_pre mask_pre res   ln    relu_post mask_post dp
    cfg = {"prenet_conv": (192, 192, 5, 1, 0, 0, False, True, 1, 0, False), "prenet_proj": (192, 192, 1, 0, 0, 0, True, False, 0, 1, False),
           "qkv": (192, 576, 1, 0, 0, 0, False, False, 0, 0, False), "outproj_ln": (192, 192, 1, 0, 0, 0, True, True, 0, 0, False),
           "ffn1": (192, 768, 3, 1, 1, 0, False, False, 0, 0, False), "ffn2_ln": (768, 192, 3, 0, 0, 1, True, True, 0, 1, False),
           "dp1": (192, 256, 3, 1, 1, 0, False, True, 0, 0, False), "dp2_proj": (256, 256, 3, 1, 1, 0, False, True, 0, 0, True)}[case]
    Cin, N, taps, mask_in, relu_pre, mask_pre, use_res, ln, relu_post, mask_post, dp = cfg
    B, T = 3, 70
    g = torch.Generator().manual_seed(Cin + N)
    dev = "cuda"
    a = torch.randn(B, T, Cin, generator=g).to(dev)
    lens = torch.tensor([T, 41, 9])
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    w = (torch.randn(N, taps, Cin, generator=g) / (taps * Cin) ** 0.5).to(dev)
    bias = (0.1 * torch.randn(N, generator=g)).to(dev)
    res = torch.randn(B, T, N, generator=g).to(dev) if use_res else None
    lg = (1 + 0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    lb = (0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    dpw = torch.randn(N, generator=g).to(dev) if dp else None
    dpb = torch.randn(4, generator=g).to(dev) if dp else None
    out = torch.full((B, T, N), float("nan"), device=dev)
    dpo = torch.full((B, T), float("nan"), device=dev) if dp else None
    rc = lib.matcha_b200_debug_te_gemm(_P(a), Cin, _P(mask), mask_in, taps, B, T, _P(w), _P(bias), N, relu_pre, mask_pre, _P(res), _P(lg), _P(lb),
                                       relu_post, mask_post, _P(out), _P(dpw), _P(dpb), _P(dpo), None)
    torch.cuda.synchronize()
    assert rc == 1
    x = a.double() * (mask[..., None].double() if mask_in else 1.0)
    v = F.conv1d(x.transpose(1, 2), w.double().permute(0, 2, 1), bias.double(), padding=taps // 2).transpose(1, 2)
    if relu_pre: v = torch.relu(v)
    if mask_pre: v = v * mask[..., None]
    if use_res: v = v + res.double()
    if ln: v = F.layer_norm(v, (N,), lg.double(), lb.double(), eps=1e-5)
    if relu_post: v = torch.relu(v)
    if mask_post: v = v * mask[..., None]
    assert rel_l2(out, v) < 2e-6
    if dp:
        assert rel_l2(dpo, ((v * mask[..., None]) @ dpw.double() + dpb[0].double()) * mask) < 2e-6


@pytest.mark.parametrize("B,T", [(2, 50), (32, 128), (1, 1), (3, 65), (2, 257)])
def test_attention_mma_kernel(B, T):
    """te_attention_mma_kernel (fp16 mode: mma.sync QK^T / PV on fp16 q|k|v rows, RoPE already applied) against torch in fp64 on
    the same fp16-rounded inputs.  P is rounded to fp16 for the PV product and the output is stored in fp16: 1.5e-3."""
    from matcha_b200._lib import load_library
    lib = load_library()
    g = torch.Generator().manual_seed(7 * B + T)
    dev = "cuda"
    qkv16 = torch.randn(B, T, 576, generator=g).to(dev).half()
    lens = torch.tensor([T] + [max(1, int(T * 0.6))] * (B - 1))
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    out16 = torch.full((B, T, 192), float("nan"), device=dev, dtype=torch.float16)
    assert lib.matcha_b200_debug_te_attention_mma(_P(qkv16), _P(mask), _P(out16), B, T, None) == 1
    torch.cuda.synchronize()
    d = qkv16.double()
    q, k, v = (d[..., i * 192:(i + 1) * 192].view(B, T, 2, 96).transpose(1, 2) for i in range(3))
    sc = q @ k.transpose(-1, -2) / 96 ** 0.5
    sc = sc.masked_fill((mask[:, None, :, None] * mask[:, None, None, :]) == 0, -1e4)
    want = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B, T, 192)
    assert torch.isfinite(out16).all()
    assert rel_l2(out16.float(), want) < 1.5e-3
