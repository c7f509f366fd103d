"""MAS kernel at BASELINE configs[1] with and without the per-token durations output (N > 1 headline step uses it)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
from matcha_b200._lib import load_library
lib = load_library()
B, T_x, T_y = 1024, 400, 2000
dev = torch.device("cuda")
g = torch.Generator(device=dev); g.manual_seed(1000)
value = torch.randn(B, T_x, T_y, device=dev, generator=g)
rng = np.random.default_rng(1000)
ty = rng.integers(int(0.7 * T_y), T_y + 1, B).astype(np.int32); ty[0] = T_y
tx = np.minimum((ty * (T_x / T_y)).astype(np.int32), T_x).clip(1); tx[0] = T_x
t_x, t_y = torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev)
path = torch.empty(B, T_x, T_y, device=dev)
dur = torch.empty(B, T_x, dtype=torch.int32, device=dev)
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def run(with_dur, dtype):
    rc = lib.matcha_b200_mas_align(path.data_ptr(), dtype, dur.data_ptr() if with_dur else None, value.data_ptr(), None,
                                   t_x.data_ptr(), t_y.data_ptr(), B, T_x, T_y, ctypes.c_float(-1e9), None, st())
    assert rc == 0, rc
for dtype in (0, 1):
    for with_dur in (False, True, False, True):
        for _ in range(5): run(with_dur, dtype)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(20): run(with_dur, dtype)
        e1.record(); torch.cuda.synchronize()
        print(f"path_dtype={dtype} durations={with_dur}: {e0.elapsed_time(e1) / 20:.4f} ms/step")
