"""Micro-benchmark of the tcgen05 implicit-GEMM kernel through the debug entry (profiling aid)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200._lib import check, load_library
lib = load_library()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
B, T = int(os.environ.get('GB_B', '32')), int(os.environ.get('GB_T', '800'))
for (c0, taps, N, with_res, with_b16, with_f32) in [(64, [0], 256, 0, 1, 0), (64, [0], 256, 0, 0, 1), (256, [0], 256, 0, 0, 1), (1024, [0], 256, 0, 0, 1), (256, [-1, 0, 1], 256, 0, 0, 1),
                                          (256, [0], 256, 1, 1, 1), (256, [0], 1024, 0, 1, 0), (256, [0], 768, 0, 1, 0), (1024, [0], 256, 1, 1, 1)]:
    a = torch.randn(B, T, c0, device="cuda").bfloat16()
    w = (torch.randn(N, len(taps) * c0, device="cuda") * 0.05).bfloat16()
    res = torch.randn(B, T, N, device="cuda") if with_res else None
    out = torch.empty(B, T, N, device="cuda") if with_f32 else None
    outb = torch.empty(B, T, N, device="cuda", dtype=torch.bfloat16) if with_b16 else None
    tp = (ctypes.c_int * len(taps))(*taps)
    p = lambda t: t.data_ptr() if t is not None else None
    def run():
        check(lib.matcha_b200_debug_gemm_bf16(p(a), c0, None, 0, len(taps), tp, B, T, p(w), N, None, 0, p(res), p(out), p(outb), None, st), "gemm")
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    fl = 2.0 * B * T * N * c0 * len(taps)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 8)(); lib.matcha_b200_debug_gemm_stamps(buf)
    st_ = [(buf[i] - buf[0]) / 1965.0 for i in range(8)]
    print(f"   stamps us: first-stage {st_[1]:.2f} acc-ready {st_[2]:.2f} chunk0-done {st_[3]:.2f} | chunk1: ldtm {st_[4]:.2f} transposed {st_[5]:.2f} math {st_[6]:.2f} stored {st_[7]:.2f}")
    print(f"K={c0*len(taps):5d} N={N:5d} res={with_res} b16={with_b16} f32={with_f32}: {us:8.1f} us  {fl/us/1e6:7.1f} TFLOP/s")
