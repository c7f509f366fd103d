"""Profiling aid: prints the clock64 stamps (in us at 1.965 GHz) of CTA 0's first utterance."""
import ctypes, os, sys
os.environ["MB200_MAS_DEBUG"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from matcha_b200 import monotonic_align as ma
from matcha_b200._lib import load_library
B, T_x, T_y = (int(x) for x in sys.argv[1:4])
dev = torch.device("cuda:0")
tx, ty = np.full(B, T_x, np.int32), np.full(B, T_y, np.int32)
v = torch.randn(B, T_x, T_y, device=dev); out = torch.empty_like(v)
txd, tyd = torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev)
for _ in range(3):
    ma.maximum_path_from_lengths(v, txd, tyd, out=out)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
load_library().matcha_b200_mas_debug_read(buf)
t0 = buf[0]
names = {0: "start", 20: "bt_start", 21: "bt_end", 22: "wr_begin", 23: "wr_zero1_end", 24: "wr_bfull", 25: "wr_end"}
for i in range(1, 20, 2):
    if buf[i]:
        names[i] = f"dp{(i-1)//2}_first"; names[i + 1] = f"dp{(i-1)//2}_last"
for i in sorted(names):
    if buf[i]:
        print(f"{names[i]:14s} {(buf[i]-t0)/1965.0:9.2f} us")
