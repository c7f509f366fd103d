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
    print(json.dumps(line), flush=True)


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
            print(json.dumps({"metric": METRIC, "value": value_rate, "unit": UNIT, "ms_per_step": ms_per_step,
                              "kernel_only": True}), flush=True)
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
                         "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "kernel_ms": k_ms, "kernel": "mas_kernel<R,HAS_MASK=false>"},
            "e2e": {"value": e2e_rate, "unit": UNIT,
                    "h2d_bytes_per_step": int(4 * int((t_x_h.astype(np.int64) * t_y_h).sum()) + 8 * B),
                    "d2h_bytes_per_step": int(4 * int((t_x_h.astype(np.int64) * t_y_h).sum()) + 4), "steps": a.e2e_steps,
                    "note": "valid [t_x][t_y] blocks only, sliced H2D / kernel / D2H pipeline; paths pre-zeroed by the caller as in the reference",
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
    if rank == 0:
        if dec is not None:
            line["cfm_decoder"] = dec
            line["gpu_launches"] = a.steps + 5 * dec["gpu_launches_per_solve"]
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
