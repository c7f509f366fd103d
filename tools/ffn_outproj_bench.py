"""Micro-benchmark of the OUTPROJ variant of the fused feed-forward kernel with its clock64 stamps (profiling aid)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200._lib import check, load_library
lib = load_library()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
for B, T in ([(8, int(os.environ['FB_T']))] if 'FB_T' in os.environ else [(8, 800), (32, 800), (8, 400)]):
    att = torch.randn(B, T, 256, device="cuda").half()
    wo = (torch.randn(256, 256, device="cuda") * 0.09).half()
    w1 = (torch.randn(1024, 256, device="cuda") * 0.09).half()
    w2 = (torch.randn(256, 1024, device="cuda") * 0.044).half()
    bo, b1, b2 = torch.randn(256, device="cuda"), torch.randn(1024, device="cuda"), torch.randn(256, device="cuda")
    g_, b_ = torch.ones(256, device="cuda"), torch.zeros(256, device="cuda")
    x1 = torch.randn(B, T, 256, device="cuda")
    mask = torch.ones(B, T, device="cuda")
    out = torch.empty(B, T, 256, device="cuda")
    out16 = torch.empty(B, T, 256, device="cuda", dtype=torch.float16)
    def run():
        check(lib.matcha_b200_debug_ffn_outproj(p(att), p(wo), p(bo), p(x1), p(g_), p(b_), p(w1), p(b1), p(w2), p(b2), p(out), p(out16),
                                                p(mask), B, T, 1, st), "ffn_outproj")
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 50
    buf = (ctypes.c_longlong * 64)(); lib.matcha_b200_debug_ffn_stamps(buf)
    u = lambda i: (buf[i] - buf[0]) / 1965.0
    print(f"B={B} T={T}: {us:8.1f} us | stamps us: x1 issued {u(4):.2f}  g0 done {u(5):.2f}  pass1 done {u(6):.2f}  merged {u(7):.2f}  A written {u(1):.2f}  "
          f"acc1(0) ready {u(8):.2f}  acc1(7) ready {u(8+35):.2f}  acc2 complete {u(2):.2f}  end {u(3):.2f}")
