"""Times the CFM Euler solve (profiling aid): python tools/decoder_one.py [B T n_steps precision iters]"""
import os, sys, time
from types import SimpleNamespace
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200.flow_matching import CFM
B, T, n = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (32, 800, 10)))
prec = sys.argv[4] if len(sys.argv) > 4 else "fp32"
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4),
          dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4), precision=prec)
cfm.estimator.initialize_weights(seed=0)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
mu, z = torch.randn(B, 80, T, device=dev, generator=g), torch.randn(B, 80, T, device=dev, generator=g)
mask = torch.ones(B, 1, T, device=dev)
t_span = torch.linspace(0, 1, n + 1, device=dev)
for i in range(iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = cfm.solve_ode_euler(z, t_span, mu, mask, None, None); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    flops = 2 * (6017024 * T + 1536 * T * T) * B * n
    print(f"ms {ms:.3f}  mel-frames/s {B*T/ms*1e3:.0f}  TFLOP/s {flops/ms/1e9:.1f}  launches {cfm.estimator.last_launches}  finite {bool(torch.isfinite(out).all())}")
