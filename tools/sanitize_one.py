"""One small plain-launch call of every path (no CUDA graphs), for compute-sanitizer:
   compute-sanitizer --tool memcheck|racecheck python tools/sanitize_one.py [mas|align|decoder|encoder ...]"""
import ctypes, os, sys
from types import SimpleNamespace
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import numpy as np, torch
which = sys.argv[1:] or ["mas", "align", "decoder", "encoder"]
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)

if "mas" in which:
    from matcha_b200 import monotonic_align as ma
    from matcha_b200._lib import load_library
    lib = load_library()
    for B, T_x, T_y in ((4, 70, 333), (3, 130, 404), (2, 400, 2000), (1, 1400, 1500)):      # cp.async, TMA, R = 13, tall fallback
        v = torch.randn(B, T_x, T_y, generator=g).to(dev)
        tx = torch.tensor([T_x] + [max(1, T_x * 3 // 4)] * (B - 1), dtype=torch.int32, device=dev)
        ty = torch.tensor([T_y] + [max(T_x, T_y * 3 // 4)] * (B - 1), dtype=torch.int32, device=dev)
        m = ((torch.arange(T_x, device=dev)[None, :, None] < tx[:, None, None]) & (torch.arange(T_y, device=dev)[None, None, :] < ty[:, None, None])).float()
        for mode in ("prefix", "full"):
            p = ma.maximum_path(v, m, mask_mode=mode)
        torch.cuda.synchronize()
        print("mas", B, T_x, T_y, "frames on path", int(p.sum()), "=", int(ty.sum()), flush=True)
    B, T_x, T_y = 4, 70, 333
    v = torch.randn(B, T_x, T_y, generator=g).to(dev)
    tx = torch.full((B,), T_x, dtype=torch.int32, device=dev); ty = torch.full((B,), T_y, dtype=torch.int32, device=dev)
    dur = torch.empty(B, T_x, dtype=torch.int32, device=dev)
    rc = lib.matcha_b200_mas_durations_static(dur.data_ptr(), v.data_ptr(), None, tx.data_ptr(), ty.data_ptr(), B, T_x, T_y,
                                              ctypes.c_float(-1e9), None, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    print("mas durations_static rc", rc, "sum", int(dur.sum()), flush=True)

if "align" in which:
    from matcha_b200 import alignment as al
    B, T_x, T_y = 3, 40, 90
    mu_x, y = torch.randn(B, 80, T_x, generator=g).to(dev), torch.randn(B, 80, T_y, generator=g).to(dev)
    tx, ty = torch.tensor([40, 33, 17]), torch.tensor([90, 80, 41])
    r = al.align(mu_x, y, tx, ty)
    torch.cuda.synchronize()
    print("align keys", sorted(r.keys()), flush=True)

if "decoder" in which:
    from matcha_b200.flow_matching import CFM
    for prec in ("fp16", "fp32"):
        cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4),
                  dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4), precision=prec)
        cfm.estimator.initialize_weights(seed=0)
        cfm.estimator.use_graph = False
        B, T = 2, 72
        z, mu = torch.randn(B, 80, T, generator=g).to(dev), torch.randn(B, 80, T, generator=g).to(dev)
        mask = torch.ones(B, 1, T, device=dev); mask[1, :, 50:] = 0
        out = cfm.estimator.solve_euler(z, torch.linspace(0, 1, 3), mu, mask, use_graph=False)
        loss, _ = cfm.compute_loss(z, mask, mu)
        torch.cuda.synchronize()
        print("decoder", prec, "finite", bool(torch.isfinite(out).all()), "loss", float(loss), flush=True)

if "encoder" in which:
    from matcha_b200.text_encoder import TextEncoder
    for prec in ("fp16", "fp32"):
        enc = TextEncoder(n_vocab=178, precision=prec).initialize_weights(0)
        enc.use_graph = False
        ids = torch.randint(0, 178, (2, 40), generator=g).to(dev)
        out = enc(ids, torch.tensor([40, 23]))
        torch.cuda.synchronize()
        print("encoder", prec, "finite", all(bool(torch.isfinite(o).all()) for o in out), flush=True)
print("SANITIZE_ONE DONE")
