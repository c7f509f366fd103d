"""Runs the MAS kernel a few times on one shape (profiling aid): python tools/mas_one.py B T_x T_y [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from matcha_b200 import monotonic_align as ma
from bench import ragged_lengths
B, T_x, T_y = (int(x) for x in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda:0")
tx, ty = ragged_lengths(np.random.default_rng(0), B, T_x, T_y)
v = torch.randn(B, T_x, T_y, device=dev); out = torch.empty_like(v)
txd, tyd = torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev)
for _ in range(iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ma.maximum_path_from_lengths(v, txd, tyd, out=out); e1.record(); torch.cuda.synchronize()
    print("ms", e0.elapsed_time(e1))
