"""Times the MAS kernel (device-resident, C ABI) over BASELINE.json configs[0]/[1] shapes and
prints one JSON line per shape: ms, alignments/s, algorithmic GB/s and fraction of measured HBM peak."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
from matcha_b200 import monotonic_align as ma
from bench import measured_peaks, ragged_lengths

dev = torch.device("cuda:0")
peak, _ = measured_peaks()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
shapes = [(16, 128, 800)] + [(B, tx, ty) for (tx, ty) in ((128, 800), (200, 1000), (400, 2000)) for B in (64, 128, 256, 512, 1024)]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) for x in s.split("x")) for s in sys.argv[1:]]
for (B, T_x, T_y) in shapes:
    rng = np.random.default_rng(B + T_x)
    for ragged in (True, False):
        if ragged:
            tx, ty = ragged_lengths(rng, B, T_x, T_y)
        else:
            tx, ty = np.full(B, T_x, np.int32), np.full(B, T_y, np.int32)
        v = torch.randn(B, T_x, T_y, device=dev)
        out = torch.empty_like(v)
        txd, tyd = torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev)
        for _ in range(3):
            ma.maximum_path_from_lengths(v, txd, tyd, out=out)
        ts = []
        for _ in range(10):
            flush.zero_()      # L2 flush between timed iterations (256 MB > 126 MB L2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ma.maximum_path_from_lengths(v, txd, tyd, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        alg = 4.0 * float((tx.astype(np.int64) * ty).sum()) + 4.0 * B * T_x * T_y
        print(json.dumps({"B": B, "T_x": T_x, "T_y": T_y, "ragged": ragged, "ms": round(ms, 4), "min_ms": round(min(ts), 4),
                          "alignments_per_s": round(B / ms * 1e3), "GBps": round(alg / ms / 1e6, 1),
                          "frac_hbm": round(alg / ms / 1e6 / peak, 3)}), flush=True)
        del v, out
