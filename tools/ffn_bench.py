"""Micro-benchmark of the fused feed-forward kernel against the two-GEMM path (profiling aid)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200._lib import check, load_library
lib = load_library()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
for B, T in [(32, 800), (32, 400), (8, 800)]:
    a = torch.randn(B, T, 256, device="cuda").half()
    w1 = (torch.randn(1024, 256, device="cuda") * 0.09).half()
    w2 = (torch.randn(256, 1024, device="cuda") * 0.044).half()
    b1, b2 = torch.randn(1024, device="cuda"), torch.randn(256, device="cuda")
    res = torch.randn(B, T, 256, device="cuda")
    mask = torch.ones(B, T, device="cuda")
    out = torch.empty(B, T, 256, device="cuda")
    out16 = torch.empty(B, T, 256, device="cuda", dtype=torch.float16)
    hid = torch.empty(B, T, 1024, device="cuda", dtype=torch.float16)
    tap = (ctypes.c_int * 1)(0)
    def fused():
        check(lib.matcha_b200_debug_ffn(p(a), p(w1), p(b1), p(w2), p(b2), p(res), p(out), p(out16), p(mask), B, T, 1, st, p(res), p(b2), p(b2)), "ffn")
    def split():
        check(lib.matcha_b200_debug_gemm_bf16(p(a), 256, None, 0, 1, tap, B, T, p(w1), 1024, p(b1), 1, None, None, p(hid), None, st), "g1")
        check(lib.matcha_b200_debug_gemm_bf16(p(hid), 1024, None, 0, 1, tap, B, T, p(w2), 256, p(b2), 0, p(res), p(out), p(out16), p(mask), st), "g2")
    for name, fn in (("fused", fused), ("split", split)):
        for _ in range(3): fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): fn()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 50
        fl = 4.0 * B * T * 256 * 1024
        print(f"B={B} T={T} {name}: {us:8.1f} us  {fl/us/1e6:7.1f} TFLOP/s")
        if name == "fused":
            buf = (ctypes.c_longlong * 64)(); lib.matcha_b200_debug_ffn_stamps(buf)
            u = lambda i: (buf[i] - buf[0]) / 1965.0
            print(f"   stamps us: A landed {u(1):.2f}  acc2 complete {u(2):.2f}  end {u(3):.2f}")
            for c in range(8):
                print(f"   chunk {c}: acc1 ready {u(8+5*c):6.2f}  tmem read {u(9+5*c):6.2f}  gelu {u(10+5*c):6.2f}  H free {u(11+5*c):6.2f}  H written {u(12+5*c):6.2f}")
