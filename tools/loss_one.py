"""Times CFM.compute_loss (one decoder eval with per-sample t) and a solve: python tools/loss_one.py [B T]"""
import os, sys
from types import SimpleNamespace
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import torch
from matcha_b200.flow_matching import CFM
B, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (32, 800)
cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4),
          dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4), precision="fp16")
cfm.estimator.initialize_weights(seed=0)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
x1, mu = torch.randn(B, 80, T, device=dev, generator=g), torch.randn(B, 80, T, device=dev, generator=g)
mask = torch.ones(B, 1, T, device=dev)
t, z = torch.rand(B, 1, 1, device=dev, generator=g), torch.randn(B, 80, T, device=dev, generator=g)
for _ in range(4):
    loss, _ = cfm.compute_loss(x1, mask, mu, t=t, z=z)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(3):
    e0.record()
    for _ in range(20):
        loss, _ = cfm.compute_loss(x1, mask, mu, t=t, z=z)
    e1.record(); torch.cuda.synchronize()
    print(f"compute_loss B={B} T={T}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us  loss {float(loss):.6f}", flush=True)
