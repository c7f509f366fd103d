"""Micro-benchmark of the self-attention kernels through the debug entry (profiling aid)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200._lib import check, load_library
lib = load_library()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for B, T in [(32, 800), (32, 400), (4, 800)]:
    qkv = torch.randn(B, T, 768, device="cuda").half()
    mask = torch.ones(B, T, device="cuda")
    out = torch.empty(B, T, 256, device="cuda", dtype=torch.float16)
    for impl in (0, 1):
        run = lambda: check(lib.matcha_b200_debug_attention(qkv.data_ptr(), mask.data_ptr(), out.data_ptr(), B, T, 1, impl, st), "attn")
        for _ in range(3): run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 50
        fl = 4.0 * B * 4 * T * T * 64
        print(f"B={B} T={T} impl={'tcgen05' if impl == 0 else 'mma.sync'}: {us:8.1f} us  {fl/us/1e6:7.1f} TFLOP/s")
