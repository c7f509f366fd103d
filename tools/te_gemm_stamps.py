"""globaltimer stamps of one te_gemm_tc_kernel launch (CTA 0): python tools/te_gemm_stamps.py"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import torch
from matcha_b200._lib import load_library
lib = load_library()
dev = "cuda"
P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
B, T = 32, 128
g = torch.Generator().manual_seed(0)
mask = torch.ones(B, T, device=dev)
for name, (Cin, N, taps, use_res, ln) in {"qkv-like K=192 N=576": (192, 576, 1, False, False), "outproj K=192 res+LN": (192, 192, 1, True, True),
                                          "ffn1 K=576 N=768": (192, 768, 3, False, False), "ffn2 K=2304 res+LN": (768, 192, 3, True, True),
                                          "prenet K=960 LN": (192, 192, 5, False, True)}.items():
    a16 = torch.randn(B, T, Cin, generator=g).to(dev).half()
    w16 = (torch.randn(N, taps, Cin, generator=g) / (taps * Cin) ** 0.5).to(dev).half()
    bias = torch.zeros(N, device=dev)
    res = torch.randn(B, T, N, generator=g).to(dev) if use_res else None
    lg, lb = (torch.ones(N, device=dev), torch.zeros(N, device=dev)) if ln else (None, None)
    out32 = torch.empty(B, T, N, device=dev); out16 = torch.empty(B, T, N, device=dev, dtype=torch.float16)
    st = torch.zeros(8, dtype=torch.int64, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(3):
        torch.cuda.synchronize(); e0.record()
        rc = lib.matcha_b200_debug_te_gemm_tc(P(a16), Cin, P(mask), taps, B, T, P(w16), P(bias), N, 0, 0, P(res), P(lg), P(lb), 0, 0, P(out32), P(out16), 0,
                                              None, None, None, None, None, P(st), None)
        e1.record(); torch.cuda.synchronize()
    s = st.cpu().tolist()
    print(f"{name:24s} rc={rc} event {e0.elapsed_time(e1)*1e3:6.1f} us | ns from CTA start: wait-done {s[1]-s[0]:6d} first-stage {s[2]-s[0]:6d} acc-ready {s[3]-s[0]:6d} epilogue-done {s[4]-s[0]:6d} cta-end {s[5]-s[0]:6d}", flush=True)

# attention
from oracle import text_encoder as oenc
for (B2, T2) in ((32, 128), (1, 128)):
    qkv = torch.randn(B2, T2, 576, generator=g).to(dev)
    m2 = torch.ones(B2, T2, device=dev)
    cos, sin = (t.contiguous() for t in oenc.rope_tables(256, dev))
    out = torch.empty(B2, T2, 192, device=dev)
    st = torch.zeros(16, dtype=torch.int64, device=dev)
    for it in range(3):
        rc = lib.matcha_b200_debug_te_attention_stamps(P(qkv), P(m2), P(cos), P(sin), P(out), B2, T2, P(st), None)
        torch.cuda.synchronize()
    s = st.cpu().tolist()
    print(f"attention B={B2} T={T2} rc={rc} ns from start: q-loaded {s[1]-s[0]} | blk0 tile-ready {s[2]-s[0]} done {s[3]-s[0]} | blk1 ready {s[4]-s[0]} done {s[5]-s[0]} | blk2 ready {s[6]-s[0]} done {s[7]-s[0]} | end {s[8]-s[0]}", flush=True)
