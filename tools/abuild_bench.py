"""Micro-benchmark of the A-building GEMMs (gemm_abuild.cuh) through the debug entry (profiling aid)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200._lib import check, load_library
lib = load_library()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
for mode in (1, 2):
    for B, T in ([(int(os.environ['AB_B']), int(os.environ['AB_T']))] if 'AB_B' in os.environ else [(32, 400), (32, 800)]):
        h = torch.randn(B, T, 256, device="cuda")
        g4 = h.double().view(B, T, 8, 32)
        gs = torch.stack([g4.sum((1, 3)), (g4 * g4).sum((1, 3))], -1).contiguous()
        gamma, beta = torch.ones(256, device="cuda"), torch.zeros(256, device="cuda")
        mask = torch.ones(B, T, device="cuda")
        tb = torch.randn(1, 256, device="cuda")
        r = torch.randn(B, T, 256, device="cuda")
        N = 768 if mode == 1 else 256
        K = 256 if mode == 1 else 768
        w = (torch.randn(N, K, device="cuda") * 0.05).half()
        bias = torch.randn(N, device="cuda")
        x1 = torch.empty(B, T, 256, device="cuda")
        out16 = torch.empty(B, T, N, device="cuda", dtype=torch.float16)
        out32 = torch.empty(B, T, N, device="cuda")
        gst = torch.zeros(B, 8, 2, device="cuda", dtype=torch.float64)
        def run():
            check(lib.matcha_b200_debug_abuild(mode, p(h), p(gs), p(gamma), p(beta), p(mask), p(tb), p(r), p(gamma), p(beta), p(x1), p(w), N,
                                               p(bias), p(out16), p(out32), p(gst), B, T, 1, st), "abuild")
        for _ in range(3): run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        print(f"mode {mode} B={B} T={T}: {e0.elapsed_time(e1) * 50:8.1f} us")
