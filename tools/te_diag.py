"""Kernel-level diagnosis of the text-encoder kernels against torch on the same GPU (fp64 reference)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import torch, torch.nn.functional as F
from matcha_b200._lib import load_library
from oracle import text_encoder as oenc
lib = load_library()
dev = torch.device("cuda:0")
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
def rel(a, b): return float(torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double()))

def gemm_case(B, T, Cin, N, ntaps, mask_in, relu_pre, mask_pre, use_res, ln, relu_post, mask_post, dp=False, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(B, T, Cin, generator=g).to(dev)
    lens = torch.tensor([T] + [max(1, int(T * 0.6))] * (B - 1))
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    w = (torch.randn(N, ntaps, Cin, generator=g) / (ntaps * Cin) ** 0.5).to(dev)
    bias = (0.1 * torch.randn(N, generator=g)).to(dev)
    res = torch.randn(B, T, N, generator=g).to(dev) if use_res else None
    lg = (1 + 0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    lb = (0.1 * torch.randn(N, generator=g)).to(dev) if ln else None
    dpw = torch.randn(N, generator=g).to(dev) if dp else None
    dpb = torch.randn(4, generator=g).to(dev) if dp else None
    out = torch.full((B, T, N), float("nan"), device=dev)
    dpo = torch.full((B, T), float("nan"), device=dev) if dp else None
    rc = lib.matcha_b200_debug_te_gemm(P(a), Cin, P(mask), int(mask_in), ntaps, B, T, P(w), P(bias), N, int(relu_pre), int(mask_pre), P(res),
                                       P(lg), P(lb), int(relu_post), int(mask_post), P(out), P(dpw), P(dpb), P(dpo), None)
    torch.cuda.synchronize()
    # reference in fp64
    x = a.double() * (mask[..., None].double() if mask_in else 1.0)
    wc = w.double().permute(0, 2, 1)                                   # [N][Cin][taps]
    v = F.conv1d(x.transpose(1, 2), wc, bias.double(), padding=ntaps // 2).transpose(1, 2)
    if relu_pre: v = torch.relu(v)
    if mask_pre: v = v * mask[..., None]
    if use_res: v = v + res.double()
    if ln: v = F.layer_norm(v, (N,), lg.double(), lb.double(), eps=1e-5)
    if relu_post: v = torch.relu(v)
    if mask_post: v = v * mask[..., None]
    msg = f"gemm B={B} T={T} Cin={Cin} N={N} taps={ntaps} min={mask_in} rp={relu_pre} mp={mask_pre} res={use_res} ln={ln} rq={relu_post} mq={mask_post}: rc={rc} rel={rel(out, v):.2e}"
    if dp:
        d = ((v * mask[..., None]) @ dpw.double() + dpb[0].double()) * mask
        msg += f" dp rel={rel(dpo, d):.2e}"
    print(msg, flush=True)

def attn_case(B, T, seed=0):
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn(B, T, 576, generator=g).to(dev)
    lens = torch.tensor([T] + [max(1, int(T * 0.6))] * (B - 1))
    mask = (torch.arange(T)[None] < lens[:, None]).float().to(dev)
    cos, sin = oenc.rope_tables(max(T, 8), dev)
    out = torch.full((B, T, 192), float("nan"), device=dev)
    rc = lib.matcha_b200_debug_te_attention(P(qkv), P(mask), P(cos.contiguous()), P(sin.contiguous()), P(out), B, T, None)
    torch.cuda.synchronize()
    d = qkv.double()
    q, k, v = (d[..., i * 192:(i + 1) * 192].view(B, T, 2, 96).transpose(1, 2) for i in range(3))
    q, k = oenc.rope(q), oenc.rope(k)
    sc = q @ k.transpose(-1, -2) / 96 ** 0.5
    am = (mask[:, None, :, None] * mask[:, None, None, :])
    sc = sc.masked_fill(am == 0, -1e4)
    o = (torch.softmax(sc, -1) @ v).transpose(1, 2).reshape(B, T, 192)
    print(f"attention B={B} T={T}: rc={rc} rel={rel(out, o):.2e}", flush=True)

print("PDL", os.environ.get("MB200_PDL", "1"))
for (B, T) in ((2, 50), (32, 128), (64, 128)):
    gemm_case(B, T, 192, 192, 5, True, False, False, False, True, True, False)        # prenet conv
    gemm_case(B, T, 192, 192, 1, False, False, False, True, False, False, True)       # prenet projection
    gemm_case(B, T, 192, 576, 1, False, False, False, False, False, False, False)     # qkv
    gemm_case(B, T, 192, 192, 1, False, False, False, True, True, False, False)       # out proj + LN1
    gemm_case(B, T, 192, 768, 3, True, True, False, False, False, False, False)       # ffn1
    gemm_case(B, T, 768, 192, 3, False, False, True, True, True, False, True)         # ffn2 + LN2
    gemm_case(B, T, 192, 256, 3, True, True, False, False, True, False, False)        # dp1
    gemm_case(B, T, 256, 256, 3, True, True, False, False, True, False, False, dp=True)  # dp2 + proj
    attn_case(B, T)
attn_case(1, 1); attn_case(3, 33); attn_case(2, 257)
