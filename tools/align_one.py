"""Runs the training-step hot path (log-prior -> MAS -> expand -> CFM loss) a few times (profiling aid):
python tools/align_one.py [B T_x T_y iters]"""
import os, sys
from types import SimpleNamespace
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200 import alignment as al
B, T_x, T_y = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (32, 128, 800)))
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
mu_x, y = torch.randn(B, 80, T_x, device=dev, generator=g), torch.randn(B, 80, T_y, device=dev, generator=g)
xl = (T_x * (0.7 + 0.3 * torch.rand(B, device=dev, generator=g))).int().clamp(min=1)
yl = torch.maximum((T_y * (0.7 + 0.3 * torch.rand(B, device=dev, generator=g))).int(), xl)
for i in range(iters):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    e[0].record(); lp = al.log_prior(mu_x, y, xl, yl)
    e[1].record(); path, dur = al.maximum_path_with_durations(lp, xl, yl)
    e[2].record(); r = al.expand(dur, mu_x, xl, yl, y=y, want_attn=False, want_logw=True, want_prior=True)
    e[3].record(); torch.cuda.synchronize()
    t = [e[j].elapsed_time(e[j + 1]) * 1e3 for j in range(3)]
    nb = 4 * B * T_x * T_y
    print(f"log_prior {t[0]:.1f} us ({nb / t[0] / 1e3:.0f} GB/s written)  mas {t[1]:.1f} us  expand {t[2]:.1f} us  "
          f"prior {float(r['prior_sums'][0] / (r['prior_sums'][1] * 80)):.4f}")
