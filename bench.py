#!/usr/bin/env python
"""bench.py -- MAS alignments/s on B200 (BASELINE.json configs[1]) with roofline, e2e and the
reference's Cython core timed on the host cores beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--batch B --tx T_x --ty T_y]

A "step" is one monotonic-alignment pass over one batch of B utterances (per GPU; weak scaling:
every rank owns its own batch slice, no data-path collective).  `value` is timed with inputs
resident in HBM (C ABI, device pointers); `e2e` goes through the host-buffer entry that
mirrors core.pyx's compute_batch_alignments(paths, values, t_xs, t_ys) with pinned numpy
buffers, H2D and D2H inside the timed region.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "mas_alignments_per_sec"
UNIT = "alignments/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="utterances per GPU")
    ap.add_argument("--tx", type=int, default=400)
    ap.add_argument("--ty", type=int, default=2000)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel-only", action="store_true", help="profiling aid: skip the e2e and CPU-baseline legs")
    ap.add_argument("--no-decoder", action="store_true", help="skip the CFM-decoder leg (BASELINE configs[2])")
    ap.add_argument("--no-pipeline-legs", action="store_true", help="skip the training-step / synthesis legs (configs[3], [4])")
    ap.add_argument("--synth-utterances", type=int, default=512, help="utterances per GPU in the synthesis leg")
    ap.add_argument("--dec-batch", type=int, default=32)
    ap.add_argument("--dec-T", type=int, default=800)
    ap.add_argument("--dec-steps", type=int, default=10, help="Euler steps per solve")
    ap.add_argument("--dec-precision", default="fp16", choices=["fp32", "fp16", "bf16"])
    return ap.parse_args()


def ragged_lengths(rng, B, T_x, T_y):
    """test_monotonic_align_speed.py:17-26: t = int(T * U[0.7, 1)), clamped to the DP's domain."""
    t_x = np.maximum(1, (T_x * (0.7 + 0.3 * rng.random(B))).astype(np.int32))
    t_y = np.maximum(t_x, (T_y * (0.7 + 0.3 * rng.random(B))).astype(np.int32)).astype(np.int32)
    return t_x.astype(np.int32), t_y


def workload_name(a):
    return f"maximum_path B={a.batch}/GPU T_text={a.tx} T_mel={a.ty} ragged 0.7-1.0 (BASELINE configs[1])"


def mas_traffic(a):
    """DRAM bytes per launch from the committed ncu --set full capture of THIS workload (profiles/r1_traffic.json), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
            t = json.load(f)["mas_kernel"]
        if t["workload"] == f"B={a.batch} T_text={a.tx} T_mel={a.ty} ragged 0.7-1.0":
            return float(t["bytes"])
    except Exception:
        pass
    return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_rate(a, seconds, max_items=256):
    """The reference's compiled core.pyx (oracle/_ref; else the C port) on the host cores."""
    from oracle import mas as omas
    use_ref = omas.ref_available()
    rng = np.random.default_rng(123)
    n = min(max_items, a.batch)
    v = rng.standard_normal((n, a.tx, a.ty), dtype=np.float32)
    t_x, t_y = ragged_lengths(rng, n, a.tx, a.ty)
    cores = min(os.cpu_count() or 1, n)
    done, t0 = 0, time.perf_counter()
    while True:
        vals = v.copy()
        paths = np.zeros(v.shape, np.int32)
        t1 = time.perf_counter()
        if use_ref:
            omas.ref_compute_batch_alignments(paths, vals, t_x, t_y)
        else:
            omas.oracle_compute_batch_alignments(paths, vals, t_x, t_y)
        dt = time.perf_counter() - t1
        done += 1
        best = dt if done == 1 else min(best, dt)  # noqa: F821
        if time.perf_counter() - t0 > seconds or done >= 50:
            break
    return {"value": n / best, "unit": UNIT, "cores": cores, "kind": "reference" if use_ref else "port",
            "sample": f"{n} utterances of the same ragged {a.tx}x{a.ty} workload, core only "
                      f"(compute_batch_alignments on numpy buffers), best of {done} passes, OpenMP prange over the batch"}


def decoder_leg(a, dev, world, rank, dist, barrier):
    """CFM decoder mel-frames/s (BASELINE configs[2]: 80-bin mel, T=800, Euler n=10, batch 32 per GPU,
    random-init Matcha-size decoder).  Returns the dict that goes under "cfm_decoder" on rank 0."""
    import torch
    from types import SimpleNamespace
    from matcha_b200.flow_matching import CFM
    B, T, n = a.dec_batch, a.dec_T, a.dec_steps
    params = dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)
    cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), params, precision=a.dec_precision)
    cfm.estimator.initialize_weights(seed=0)                   # Decoder.initialize_weights distribution (decoder.py:255)
    weights = cfm.estimator.state_dict()
    g = torch.Generator(device=dev); g.manual_seed(2000 + rank)
    mu = torch.randn(B, 80, T, device=dev, generator=g)
    z = torch.randn(B, 80, T, device=dev, generator=g)
    mask = torch.ones(B, 1, T, device=dev)
    t_span = torch.linspace(0, 1, n + 1, device=dev)
    for _ in range(3):
        out = cfm.solve_ode_euler(z, t_span, mu, mask, None, None)
    barrier()
    K = 5
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(K):
        out = cfm.solve_ode_euler(z, t_span, mu, mask, None, None)
    ev[1].record()
    barrier()
    ms = torch.tensor([ev[0].elapsed_time(ev[1]) / K], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    launches = int(cfm.estimator.last_launches)
    # e2e: pinned host tensors in, host tensor out, through the reference-shaped CFM API
    mu_h, z_h, mask_h = mu.cpu().pin_memory(), z.cpu().pin_memory(), mask.cpu().pin_memory()
    out_h = torch.empty(B, 80, T).pin_memory()
    def e2e_once():
        o = cfm.solve_ode_euler(z_h.to(dev, non_blocking=True), t_span, mu_h.to(dev, non_blocking=True),
                                mask_h.to(dev, non_blocking=True), None, None)
        out_h.copy_(o, non_blocking=True)
        torch.cuda.synchronize()
    e2e_once(); barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        e2e_once()
    barrier()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / 3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    flops = 2.0 * (6017024.0 * T + 1536.0 * T * T) * B * n
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak, src = float(json.load(f)["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    except Exception:
        peak, src = 1400.0, "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)"
    res = {
        "metric": "cfm_decoder_mel_frames_per_sec", "value": world * B * T / (ms / 1e3), "unit": "mel-frames/s",
        "ms_per_solve": ms, "n_gpus": world, "scaling": "weak", "dtype": a.dec_precision, "data": "synthetic",
        "config": {"workload": f"CFM Euler solve B={B}/GPU T_mel={T} n_timesteps={n} 80-bin mel, random-init "
                               f"Matcha-size decoder (BASELINE configs[2])", "precision": a.dec_precision},
        "roofline": {"bound": "tensor", "achieved": flops / (ms / 1e3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                     "frac": flops / (ms / 1e3) / 1e12 / peak, "traffic": None, "peak_source": src,
                     "algorithmic_flops_per_solve": flops, "kernel": "tcgen05 kernels: gemm_bf16_kernel, gemm_abuild_kernel, ffn_fused_kernel, attention_tc_kernel"},
        "e2e": {"value": world * B * T / float(e2e_s.item()), "unit": "mel-frames/s",
                "h2d_bytes_per_step": int(4 * (2 * B * 80 * T + B * T)), "d2h_bytes_per_step": int(4 * B * 80 * T),
                "api": "matcha_b200.flow_matching.CFM.solve_ode_euler on pinned host tensors"},
        "gpu_launches_per_solve": launches,
    }
    if a.dec_precision != "fp32":          # accuracy of the 16-bit mode against our own fp32 mode, same inputs
        ref = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), params, precision="fp32")
        ref.load_state_dict(weights)
        o32 = ref.solve_ode_euler(z, t_span, mu, mask, None, None)
        res["rel_l2_vs_fp32_mode"] = float(torch.linalg.norm((out - o32).double()) / torch.linalg.norm(o32.double()))
    if world == 1 and not a.no_cpu_baseline:
        # the reference's arithmetic (torch CPU fp32, oracle/decoder.py) on a bounded sample
        from oracle import decoder as odec                     # the checker, timed as the CPU baseline only
        torch.set_num_threads(os.cpu_count() or 1)
        w = weights
        bs, ns = min(4, B), 2
        xs, mus, ms_ = z[:bs].cpu(), mu[:bs].cpu(), mask[:bs].cpu()
        tsp = torch.linspace(0, 1, ns + 1)
        odec.solve_euler(w, xs, torch.linspace(0, 1, 2), mus, ms_)      # warm-up, one eval
        t1 = time.perf_counter()
        odec.solve_euler(w, xs, tsp, mus, ms_)
        dt = time.perf_counter() - t1
        res["cpu_baseline"] = {"value": bs * T / (dt * n / ns), "unit": "mel-frames/s", "cores": torch.get_num_threads(),
                               "kind": "port", "sample": f"torch CPU fp32 restatement of the reference decoder, B={bs}, T={T}, "
                                                         f"{ns} Euler steps timed and scaled to n={n}"}
    return res


def _dec_params():
    return dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)


def training_leg(a, dev, world, rank, dist, barrier):
    """BASELINE configs[3]: the training-step hot path after the text encoder (out of scope):
    mu_x -> log-prior -> MAS -> logw_ / mu_y / prior loss -> CFM flow-matching loss forward, synthetic
    LJSpeech-shaped batch 32 per GPU (T_text 128, T_mel 800, ragged 0.7-1.0)."""
    import torch
    from types import SimpleNamespace
    from matcha_b200 import alignment as al
    from matcha_b200.flow_matching import CFM
    B, T_x, T_y = a.dec_batch, 128, a.dec_T
    cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), _dec_params(), precision=a.dec_precision)
    cfm.estimator.initialize_weights(seed=0)
    rng = np.random.default_rng(3000 + rank)
    xl_h, yl_h = ragged_lengths(rng, B, T_x, T_y)
    g = torch.Generator(device=dev); g.manual_seed(3000 + rank)
    mu_x = torch.randn(B, 80, T_x, device=dev, generator=g)
    y = torch.randn(B, 80, T_y, device=dev, generator=g)
    logw = torch.randn(B, 1, T_x, device=dev, generator=g)
    xl, yl = torch.from_numpy(xl_h).to(dev), torch.from_numpy(yl_h).to(dev)
    y_mask = (torch.arange(T_y, device=dev)[None, :] < yl[:, None]).float().unsqueeze(1)
    x_mask = (torch.arange(T_x, device=dev)[None, :] < xl[:, None]).float().unsqueeze(1)

    def step(mu_x, y, logw):
        r = al.align(mu_x, y, xl, yl)                                             # matcha_tts.py:277-287, 314-321
        dur_loss = torch.sum((logw * x_mask - r["logw_"]) ** 2) / torch.sum(xl)    # utils/model.py:117-135
        diff_loss, _ = cfm.compute_loss(y, y_mask, r["mu_y"])                     # flow_matching.py:106-151
        return torch.stack([dur_loss, r["prior_loss"], diff_loss])

    for _ in range(3):
        losses = step(mu_x, y, logw)
    barrier()
    K = 10
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(K):
        losses = step(mu_x, y, logw)
    ev[1].record()
    barrier()
    ms = torch.tensor([ev[0].elapsed_time(ev[1]) / K], device=dev, dtype=torch.float64)
    # stage split (device time of each stage alone, same inputs)
    def timed(fn, n=10):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    lp = al.log_prior(mu_x, y, xl, yl)
    _, dur = al.maximum_path_with_durations(lp, xl, yl)
    split = {"log_prior_ms": timed(lambda: al.log_prior(mu_x, y, xl, yl)),
             "mas_ms": timed(lambda: al.maximum_path_with_durations(lp, xl, yl)),
             "expand_ms": timed(lambda: al.expand(dur, mu_x, xl, yl, y=y, want_logw=True, want_prior=True)),
             "cfm_loss_ms": timed(lambda: cfm.compute_loss(y, y_mask, y))}
    # e2e: pinned host tensors in, the three losses out
    mu_h, y_h, lw_h = mu_x.cpu().pin_memory(), y.cpu().pin_memory(), logw.cpu().pin_memory()
    def e2e_once():
        return step(mu_h.to(dev, non_blocking=True), y_h.to(dev, non_blocking=True), lw_h.to(dev, non_blocking=True)).cpu()
    e2e_once(); barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        e2e_once()
    barrier()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / 5], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    ms = float(ms.item())
    res = {"metric": "training_hot_path_utterances_per_sec", "value": world * B / (ms / 1e3), "unit": "utterances/s",
           "ms_per_step": ms, "n_gpus": world, "scaling": "weak", "dtype": a.dec_precision, "data": "synthetic",
           "config": {"workload": f"mu_x -> log-prior -> MAS -> mu_y/logw_/prior loss -> CFM loss forward, B={B}/GPU "
                                  f"T_text={T_x} T_mel={T_y} ragged 0.7-1.0 (BASELINE configs[3]); text encoder and "
                                  f"backward pass out of scope"},
           "stage_ms": split, "losses": [float(v) for v in losses.cpu()],
           "e2e": {"value": world * B / float(e2e_s.item()), "unit": "utterances/s",
                   "h2d_bytes_per_step": int(4 * (mu_x.numel() + y.numel() + logw.numel())), "d2h_bytes_per_step": 12,
                   "api": "matcha_b200.alignment.align + matcha_b200.flow_matching.CFM.compute_loss on pinned host tensors"}}
    if not a.no_cpu_baseline:
        from oracle import alignment as oal, decoder as odec       # the checker, timed as the CPU baseline only
        torch.set_num_threads(os.cpu_count() or 1)
        mc, yc = mu_x.cpu(), y.cpu()
        t1 = time.perf_counter()
        r = oal.align(mc, yc, xl.cpu(), yl.cpu(), use_ref=oal.omas.ref_available())
        t_align = time.perf_counter() - t1
        bs = min(4, B)
        w = cfm.estimator.state_dict()
        tt, zz = torch.rand(bs, 1, 1), torch.randn(bs, 80, T_y)
        t1 = time.perf_counter()
        odec.compute_loss(w, yc[:bs], y_mask[:bs].cpu(), r["mu_y"][:bs], tt, zz)
        t_loss = (time.perf_counter() - t1) * B / bs
        res["cpu_baseline"] = {"value": B / (t_align + t_loss), "unit": "utterances/s", "cores": torch.get_num_threads(),
                               "kind": "reference" if oal.omas.ref_available() else "port",
                               "sample": f"torch CPU restatement of matcha_tts.py:277-321 with the compiled core.pyx on the "
                                         f"full batch ({t_align * 1e3:.0f} ms) + decoder loss forward on {bs} items scaled to {B}"}
    return res


def synthesis_leg(a, dev, world, rank, dist, barrier):
    """BASELINE configs[4]: batched synthesis throughput, 512 synthetic utterances per GPU (4096 over 8 GPUs) in
    batches of 32, MAS-free inference: durations -> generate_path / mu_y (matcha_tts.py:218-229) -> CFM Euler
    solve with n_timesteps 2 / 4 / 10, then one all-gather of the mels over NCCL when N > 1."""
    import torch
    from types import SimpleNamespace
    from matcha_b200 import alignment as al
    from matcha_b200.flow_matching import CFM
    n_utt, bs, T_x, T_max = a.synth_utterances, a.dec_batch, 128, a.dec_T
    cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), _dec_params(), precision=a.dec_precision)
    cfm.estimator.initialize_weights(seed=0)
    rng = np.random.default_rng(4000 + rank)
    g = torch.Generator(device=dev); g.manual_seed(4000 + rank)
    batches = []
    for _ in range(n_utt // bs):
        xl_h, yl_h = ragged_lengths(rng, bs, T_x, T_max)
        yl_h[0], xl_h[0] = T_max, T_x                       # every batch pads to T_mel = 800, as LJSpeech-length text does
        d = np.zeros((bs, T_x), np.float32)
        for i in range(bs):                                 # integer durations >= 1 summing to the utterance's mel length
            cuts = np.sort(rng.choice(np.arange(1, yl_h[i]), xl_h[i] - 1, replace=False))
            d[i, : xl_h[i]] = np.diff(np.concatenate([[0], cuts, [yl_h[i]]]))
        logw = torch.from_numpy(np.log(np.maximum(d, 1.0) - 0.5)).unsqueeze(1).to(dev)     # ceil(exp(logw)) == d
        x_mask = (torch.arange(T_x, device=dev)[None, :] < torch.from_numpy(xl_h).to(dev)[:, None]).float().unsqueeze(1)
        batches.append((torch.randn(bs, 80, T_x, device=dev, generator=g), logw, x_mask, int(yl_h.sum())))
    frames = sum(b[3] for b in batches)
    mels = torch.empty(n_utt, 80, T_max, device=dev)
    gathered = torch.empty(world * n_utt, 80, T_max, device=dev) if world > 1 else None

    def run_all(n):
        for i, (mu_x, logw, x_mask, _) in enumerate(batches):
            pre = al.synthesis_prestep(mu_x, logw, x_mask, 1.0)                   # matcha_tts.py:218-229
            out = cfm(pre["mu_y"], pre["y_mask"], n, 1.0)                         # :234 (randn + Euler solve)
            mels[i * bs:(i + 1) * bs, :, : out.shape[-1]] = out
        if world > 1:
            dist.all_gather_into_tensor(gathered, mels)

    per_n = {}
    for n in (2, 4, 10):
        run_all(n); barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        run_all(n)
        ev[1].record()
        barrier()
        ms = torch.tensor([ev[0].elapsed_time(ev[1])], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        per_n[f"n_timesteps_{n}"] = {"ms_per_pass": ms, "utterances_per_sec": world * n_utt / (ms / 1e3),
                                    "mel_frames_per_sec": world * frames / (ms / 1e3)}
    if rank != 0:
        return None
    return {"metric": "synthesis_utterances_per_sec", "value": per_n["n_timesteps_10"]["utterances_per_sec"],
            "unit": "utterances/s", "n_gpus": world, "scaling": "weak", "dtype": a.dec_precision, "data": "synthetic",
            "config": {"workload": f"{n_utt} utterances/GPU in batches of {bs}, T_text={T_x}, T_mel<={T_max} (valid frames "
                                   f"counted), durations -> generate_path/mu_y -> CFM solve, all-gather of mels at N>1 "
                                   f"(BASELINE configs[4]); text encoder and vocoder out of scope"},
            "by_n_timesteps": per_n}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    best = None
    from oracle import mas as omas
    use_ref = omas.ref_available()
    rng = np.random.default_rng(123)
    n = min(128, a.batch)
    v = rng.standard_normal((n, a.tx, a.ty), dtype=np.float32)
    t_x, t_y = ragged_lengths(rng, n, a.tx, a.ty)
    times = []
    for i in range(a.warmup + a.steps):
        vals, paths = v.copy(), np.zeros(v.shape, np.int32)
        t1 = time.perf_counter()
        if use_ref:
            omas.ref_compute_batch_alignments(paths, vals, t_x, t_y)
        else:
            omas.oracle_compute_batch_alignments(paths, vals, t_x, t_y)
        if i >= a.warmup:
            times.append(time.perf_counter() - t1)
    ms = 1e3 * float(np.mean(times))
    val = n / (ms / 1e3)
    cores = min(os.cpu_count() or 1, n)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample_per_step": f"{n} utterances on the host"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "reference" if use_ref else "port",
                         "sample": f"{n} utterances per step, core.pyx compute_batch_alignments, OpenMP over the batch"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    _emit(line)


def run_ours(a):
    import torch
    import torch.distributed as dist
    from matcha_b200 import monotonic_align as ma
    from matcha_b200._lib import load_library
    import ctypes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, T_x, T_y = a.batch, a.tx, a.ty
    rng = np.random.default_rng(1000 + rank)
    t_x_h, t_y_h = ragged_lengths(rng, B, T_x, T_y)
    g = torch.Generator(device=dev); g.manual_seed(1000 + rank)
    value = torch.randn(B, T_x, T_y, device=dev, generator=g)
    t_x, t_y = torch.from_numpy(t_x_h).to(dev), torch.from_numpy(t_y_h).to(dev)
    path = torch.empty(B, T_x, T_y, dtype=torch.float32, device=dev)

    def step():
        ma.maximum_path_from_lengths(value, t_x, t_y, out=path)

    for _ in range(a.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    ev[0].record()
    for i in range(a.steps):
        step()
        ev[i + 1].record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.steps)]
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / a.steps
    value_rate = world * B / (ms_per_step / 1e3)

    if a.kernel_only:
        if rank == 0:
            _emit({"metric": METRIC, "value": value_rate, "unit": UNIT, "ms_per_step": ms_per_step, "kernel_only": True})
        if world > 1:
            dist.destroy_process_group()
        return

    # e2e: pinned numpy buffers through the core.pyx-shaped host entry (H2D + kernel + D2H per call)
    vh = torch.empty(B, T_x, T_y, dtype=torch.float32).pin_memory()
    vh.copy_(value)
    ph = torch.zeros(B, T_x, T_y, dtype=torch.int32).pin_memory()   # pre-zeroed by the caller, as the reference's np.zeros (__init__.py:49)
    vn, pn = vh.numpy(), ph.numpy()
    for _ in range(1):
        ma.compute_batch_alignments(pn, vn, t_x_h, t_y_h)
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.e2e_steps):
        ma.compute_batch_alignments(pn, vn, t_x_h, t_y_h)
    barrier()
    e2e_s = (time.perf_counter() - t0) / a.e2e_steps
    te = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_rate = world * B / float(te.item())
    same = bool(torch.equal(ph.to(dev).float(), path))

    if rank == 0:
        peak, peak_src = measured_peaks()
        alg_bytes = 4.0 * float((t_x_h.astype(np.int64) * t_y_h.astype(np.int64)).sum()) + 4.0 * B * T_x * T_y
        k_ms = float(np.mean(kern_ms))
        achieved = alg_bytes / (k_ms / 1e3) / 1e9
        out = (ctypes.c_int * 8)()
        load_library().matcha_b200_mas_query_launch(B, T_x, T_y, 0, 1, ctypes.byref(out))
        line = {
            "metric": METRIC, "value": value_rate, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(a), "batch_per_gpu": B, "T_text": T_x, "T_mel": T_y,
                       "l2": "inputs+outputs 6.5 GB per step >> 126 MB L2, no flush needed",
                       "launch": {"grid": out[0], "block": out[1], "smem": out[2], "rows_per_lane": out[3],
                                  "stages": out[4], "tma": out[5], "ctas_per_sm": out[6]},
                       "parallelism": f"batch-sharded x{world}, no collective in the timed region"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": mas_traffic(a), "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "kernel_ms": k_ms, "kernel": "mas_kernel<R,HAS_MASK=false>"},
            "e2e": {"value": e2e_rate, "unit": UNIT,
                    "h2d_bytes_per_step": int(4 * int((t_x_h.astype(np.int64) * t_y_h).sum()) + 8 * B),
                    "d2h_bytes_per_step": int(4 * B * T_x + 4), "steps": a.e2e_steps,
                    "note": "valid [t_x][t_y] value blocks in, per-token durations out (expanded into the caller's pre-zeroed int32 paths on the host), sliced H2D / kernel / D2H pipeline",
                    "api": "matcha_b200.monotonic_align.compute_batch_alignments(numpy pinned) -> matcha_b200_mas_maximum_path_host",
                    "matches_device_result": same},
            "gpu_launches": a.steps, "clocks": clocks,
        }
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_rate(a, a.cpu_seconds)
    dec = None
    if not a.no_decoder:
        del value, path
        torch.cuda.empty_cache()
        dec = decoder_leg(a, dev, world, rank, dist, barrier)
    train = synth = None
    if not a.no_decoder and not a.no_pipeline_legs:
        train = training_leg(a, dev, world, rank, dist, barrier)
        synth = synthesis_leg(a, dev, world, rank, dist, barrier)
    if rank == 0:
        if dec is not None:
            line["cfm_decoder"] = dec
            line["gpu_launches"] = a.steps + 5 * dec["gpu_launches_per_solve"]
        if train is not None:
            line["training_step"] = train
            line["gpu_launches"] += 10 * (3 + 2 + 185)      # 10 timed steps: align kernels, loss kernels, one graph-replayed eval
        if synth is not None:
            line["synthesis"] = synth
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


def _emit(obj):
    """The one JSON line goes to the process's ORIGINAL stdout; everything else any library prints to fd 1
    during the run (NCCL prints its version banner there) has been redirected to stderr."""
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


if __name__ == "__main__":
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
