#!/usr/bin/env python
"""bench.py -- MAS alignments/s on B200 (BASELINE.json configs[1]) with roofline, e2e and the
reference's Cython core timed on the host cores beside it; the CFM decoder (configs[2]), the training-step
hot path (configs[3]) and batched synthesis (configs[4]) ride along as extra keys of the same JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--batch B --tx T_x --ty T_y]

A "step" is one monotonic-alignment pass over one batch of B utterances (per GPU; weak scaling: every rank
owns its own batch slice).  At N > 1 the step also all-gathers the alignment over NCCL in its compact form
(the kernel's per-token durations), on a high-priority side stream under the next step's kernel; the dense forms of
the gather are timed beside it under "mas_gather".  `value` is timed with inputs resident in HBM (C ABI, device pointers);
`e2e` goes through the host-buffer entry that mirrors core.pyx's compute_batch_alignments(paths, values,
t_xs, t_ys) with pinned numpy buffers, H2D and D2H inside the timed region.  Prints ONE JSON line on rank 0.
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel-only", action="store_true", help="profiling aid: skip the e2e and CPU-baseline legs")
    ap.add_argument("--no-decoder", action="store_true", help="skip the CFM-decoder leg (BASELINE configs[2])")
    ap.add_argument("--no-pipeline-legs", action="store_true", help="skip the training-step / synthesis legs (configs[3], [4])")
    ap.add_argument("--no-mas-extra", action="store_true", help="skip the config-1 / sweep / gather-variant MAS legs")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the PyTorch-eager decoder on the same GPU")
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


def bench_config(a):
    """The `config` object of the JSON line: the same dict in both arms (per-arm remarks live under `notes`)."""
    gb = 2 * 4.0 * a.batch * a.tx * a.ty / 1e9
    return {"workload": workload_name(a), "batch_per_gpu": a.batch, "T_text": a.tx, "T_mel": a.ty,
            "l2": f"inputs+outputs {gb:.1f} GB per step >> 126 MB L2, no flush needed"}


def bench_summary(line):
    """Compact digest (< 600 characters) of every leg, emitted as the LAST key of the JSON line."""
    def r(x, n=3):
        return None if x is None else float(f"{x:.{n}g}")
    sm = {"n": line["n_gpus"], "mas_Malign_s": r(line["value"] / 1e6), "mas_hbm_frac": r(line["roofline"]["frac"]),
          "mas_e2e_kalign_s": r(line["e2e"]["value"] / 1e3)}
    dec = line.get("cfm_decoder")
    if dec:
        bp = dec.get("by_precision", {})
        sm["dec_ms"] = {k: r(v["ms_per_solve"]) for k, v in bp.items()}
        sm["dec_frac"] = r(dec["roofline"]["frac"])
        sm["dec_relL2"] = {k: r(v, 2) for k, v in (dec.get("rel_l2_vs_oracle") or {}).items()}
        gb = dec.get("gpu_baseline", {}).get("speedup_ours_over_eager")
        if gb:
            sm["dec_vs_eager"] = {"fp32": r(gb["fp32_mode_vs_eager_fp32"]), "fp16_vs_bf16ac": r(gb["fp16_mode_vs_eager_bf16_autocast"])}
    te = line.get("text_encoder")
    if te:
        sm["te_ms"] = {k: r(v["ms_per_forward"]) for k, v in te["modes"].items()}
    tr = line.get("training_step")
    if tr:
        sm["train_utt_s"] = r(tr["value"])
        sm["train_e2e_utt_s"] = r((tr.get("e2e") or {}).get("value"))
    sy = line.get("synthesis")
    if sy:
        sm["synth_utt_s"] = {k.replace("n_timesteps_", "n"): r(v["utterances_per_sec"]) for k, v in sy["by_n_timesteps"].items()}
    return sm


def traffic_record(kernel, workload):
    """DRAM bytes per launch (or per solve) from the committed ncu --set full capture of THIS workload, else None.
    profiles/r2_traffic.json is preferred; profiles/r1_traffic.json is the round-1 capture (older kernel revision)."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                t = json.load(f)[kernel]
            if t["workload"] == workload:
                return float(t["bytes"]), name
        except Exception:
            continue
    return None, None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """SM clock and throttle reasons DURING a timed region (B200_PROFILING.md recipe).  NVML polled every ~2 ms from a
    thread (the timed regions are tens of ms; `nvidia-smi -lms 100` would see two samples); falls back to nvidia-smi."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.index, self.sm, self.mx, self.reasons = index, [], [], set()
        self._stop, self._thread, self._nvml, self._proc = threading.Event(), None, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            phys = self.index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    phys = int(vis.split(",")[self.index])
                except Exception:
                    phys = self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml

            def pump():
                while not self._stop.is_set():
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        self.mx.append(mx)
                        r = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)) if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                            else int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        for name, bit in self.REASONS:
                            if r & bit:
                                self.reasons.add(name)
                    except Exception:
                        pass
                    time.sleep(0.002)
            self._thread = threading.Thread(target=pump, daemon=True)
            self._thread.start()
            return
        except Exception:
            self._nvml = None
        try:
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            self._proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                           "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

            def pump_smi():
                for line in self._proc.stdout:
                    r = [c.strip() for c in line.split(",")]
                    try:
                        self.sm.append(float(r[0])); self.mx.append(float(r[1]))
                    except Exception:
                        continue
                    for (name, _), v in zip(self.REASONS, r[2:6]):
                        if v.lower().startswith("active"):
                            self.reasons.add(name)
            self._thread = threading.Thread(target=pump_smi, daemon=True)
            self._thread.start()
        except Exception:
            self._proc = None

    def stop(self):
        self._stop.set()
        if self._proc is not None:
            time.sleep(0.05)
            self._proc.terminate()
        if self._thread is not None:
            self._thread.join(timeout=1.0)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": max(self.mx), "reasons": sorted(self.reasons),
                "samples": len(self.sm), "source": "nvml" if self._nvml is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------------ CPU reference (MAS)
def mas_cpu_rate(B, T_x, T_y, threads, warmup, steps, wrapper=False, full_length=False, seed=123):
    """The reference's own MAS on the host cores: compiled core.pyx from oracle/_ref (else the C restatement), either the
    core alone on pre-built numpy buffers (compute_batch_alignments, core.pyx:101-128) or through the restated wrapper
    maximum_path on torch CPU tensors (__init__.py:40-55).  One protocol everywhere: `warmup` untimed calls, then the
    MEAN of `steps` timed calls.  Returns (alignments/s, ms per call, kind, threads in force)."""
    from oracle import mas as omas
    use_ref = omas.ref_available()
    eff = omas.set_omp_threads(threads)
    rng = np.random.default_rng(seed)
    nb = min(B, 128)                 # N(0,1) scores; batches above 128 utterances tile a 128-utterance block (the lengths below stay per item)
    v = rng.standard_normal((nb, T_x, T_y), dtype=np.float32)
    if B > nb:
        v = np.concatenate([v] * ((B + nb - 1) // nb), 0)[:B]
    if full_length:
        t_x, t_y = np.full(B, T_x, np.int32), np.full(B, T_y, np.int32)
    else:
        t_x, t_y = ragged_lengths(rng, B, T_x, T_y)
    times = []
    if wrapper and use_ref:
        import torch
        torch.set_num_threads(max(1, int(threads)))
        vt = torch.from_numpy(v)
        mask = torch.zeros(B, T_x, T_y)
        for b in range(B):
            mask[b, : t_x[b], : t_y[b]] = 1
        for i in range(warmup + steps):
            t1 = time.perf_counter()
            omas.maximum_path_ref_torch(vt, mask)
            if i >= warmup:
                times.append(time.perf_counter() - t1)
    else:
        for i in range(warmup + steps):
            vals, paths = v.copy(), np.zeros(v.shape, np.int32)     # core.pyx clobbers values and needs zeroed paths
            t1 = time.perf_counter()
            if use_ref:
                omas.ref_compute_batch_alignments(paths, vals, t_x, t_y)
            else:
                omas.oracle_compute_batch_alignments(paths, vals, t_x, t_y, n_threads=threads)
            if i >= warmup:
                times.append(time.perf_counter() - t1)
    ms = 1e3 * float(np.mean(times))
    return B / (ms / 1e3), ms, ("reference" if use_ref else "port"), eff


def mas_cpu_baseline_block(B, T_x, T_y, full=True):
    """core-only and wrapper, each at all cores and at 1 thread (bounded samples)."""
    cores = host_cores()
    out = {"host_cores": cores}
    nb = min(B, 128)
    v, ms, kind, eff = mas_cpu_rate(min(B, 1024), T_x, T_y, cores, 1, 3)
    out["core_all_cores"] = {"value": v, "ms_per_call": ms, "threads": eff, "utterances": min(B, 1024)}
    if full:
        v, ms, kind, eff = mas_cpu_rate(nb, T_x, T_y, 1, 1, 2)
        out["core_1_thread"] = {"value": v, "ms_per_call": ms, "threads": eff, "utterances": nb}
        v, ms, kind, eff = mas_cpu_rate(nb, T_x, T_y, cores, 1, 2, wrapper=True)
        out["wrapper_all_cores"] = {"value": v, "ms_per_call": ms, "threads": eff, "utterances": nb}
        v, ms, kind, eff = mas_cpu_rate(nb, T_x, T_y, 1, 1, 2, wrapper=True)
        out["wrapper_1_thread"] = {"value": v, "ms_per_call": ms, "threads": eff, "utterances": nb}
    out["kind"] = kind
    out["unit"] = UNIT
    out["protocol"] = "1 warm-up call, mean of 2-3 timed calls; core = compute_batch_alignments on numpy buffers, wrapper = maximum_path(value, mask) on torch CPU tensors"
    return out


# ------------------------------------------------------------------------------------------------ MAS extra legs (N = 1)
def mas_config1_leg(dev, with_cpu):
    """BASELINE configs[0]: 16 x 128 x 800, ragged and full-length, 3 warm-ups + 20 runs as
    test_monotonic_align_speed.py:73-84 does, Cython on the host cores beside it."""
    import torch
    from matcha_b200 import monotonic_align as ma
    B, T_x, T_y = 16, 128, 800
    res = {"workload": "maximum_path 16 x 128 x 800 (BASELINE configs[0], test_monotonic_align_speed.py:17-26,73-84), 3 warm-ups + 20 runs"}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, full in (("ragged", False), ("full_length", True)):
        rng = np.random.default_rng(42)
        v = torch.from_numpy(rng.standard_normal((B, T_x, T_y), dtype=np.float32)).to(dev)
        if full:
            tx_h, ty_h = np.full(B, T_x, np.int32), np.full(B, T_y, np.int32)
        else:
            tx_h, ty_h = ragged_lengths(rng, B, T_x, T_y)
        mask = torch.zeros(B, T_x, T_y, device=dev)
        for b in range(B):
            mask[b, : tx_h[b], : ty_h[b]] = 1
        t_x, t_y = torch.from_numpy(tx_h).to(dev), torch.from_numpy(ty_h).to(dev)
        out = torch.empty_like(v)
        for _ in range(3):
            ma.maximum_path_from_lengths(v, t_x, t_y, out=out)
        torch.cuda.synchronize()
        us = []
        for _ in range(20):
            flush.zero_()                                            # L2 flush: the 13 MB working set would otherwise stay resident
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ma.maximum_path_from_lengths(v, t_x, t_y, out=out); e1.record()
            torch.cuda.synchronize()
            us.append(1e3 * e0.elapsed_time(e1))
        for _ in range(3):
            ma.maximum_path(v, mask)
        torch.cuda.synchronize()
        wall = []
        for _ in range(20):                                          # the reference's own measurement shape: wall time of maximum_path(value, mask)
            t0 = time.perf_counter(); ma.maximum_path(v, mask); torch.cuda.synchronize(); wall.append(time.perf_counter() - t0)
        r = {"kernel_us": float(np.mean(us)), "kernel_alignments_per_sec": B / (float(np.mean(us)) * 1e-6),
             "maximum_path_call_us": 1e6 * float(np.mean(wall)), "maximum_path_alignments_per_sec": B / float(np.mean(wall)),
             "l2": "256 MB write between timed launches"}
        if with_cpu:
            cores = host_cores()
            c_all, ms_all, kind, eff = mas_cpu_rate(B, T_x, T_y, cores, 3, 20, full_length=full, seed=42)
            c_one, ms_one, _, _ = mas_cpu_rate(B, T_x, T_y, 1, 3, 20, full_length=full, seed=42)
            w_all, wms_all, _, _ = mas_cpu_rate(B, T_x, T_y, cores, 3, 20, wrapper=True, full_length=full, seed=42)
            w_one, wms_one, _, _ = mas_cpu_rate(B, T_x, T_y, 1, 3, 20, wrapper=True, full_length=full, seed=42)
            r["cpu_baseline"] = {"kind": kind, "cores": min(cores, B), "unit": UNIT,
                                 "core_all_cores": c_all, "core_1_thread": c_one, "wrapper_all_cores": w_all, "wrapper_1_thread": w_one,
                                 "note": "prange uses at most B = 16 threads"}
            r["speedup_kernel_vs_core_all_cores"] = r["kernel_alignments_per_sec"] / c_all
            r["speedup_call_vs_wrapper_all_cores"] = r["maximum_path_alignments_per_sec"] / w_all
        res[name] = r
    return res


def mas_sweep_leg(dev, world, rank, dist, barrier, with_cpu):
    """BASELINE configs[1]: global batch 64-1024 x three (T_text, T_mel) shapes with ragged masks, batch-sharded
    contiguous-by-rank over the N GPUs (strong scaling for this leg: the global batch is fixed), compact all-gather of
    the alignment at N > 1; the Cython core on the host cores beside each row at N = 1."""
    import torch
    from matcha_b200 import alignment as al, sharding
    rows = []
    peak, _ = measured_peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (T_x, T_y) in ((128, 800), (200, 1000), (400, 2000)):
        for B in (64, 128, 256, 512, 1024):
            lo, hi = sharding.batch_slice(B, rank, world)
            rng = np.random.default_rng(B * 7 + T_x)
            tx_h, ty_h = ragged_lengths(rng, B, T_x, T_y)
            g = torch.Generator(device=dev); g.manual_seed(B + T_x + rank)
            v = torch.randn(hi - lo, T_x, T_y, device=dev, generator=g)
            t_x, t_y = torch.from_numpy(tx_h[lo:hi]).to(dev), torch.from_numpy(ty_h[lo:hi]).to(dev)
            gathered = torch.empty(B, T_x, dtype=torch.int32, device=dev) if world > 1 else None

            def step():
                _, dur = al.maximum_path_with_durations(v, t_x, t_y)
                if world > 1:
                    sharding.all_gather_durations(dur, B, out=gathered)
            for _ in range(3):
                step()
            barrier()
            ms = []
            for _ in range(5):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); step(); e1.record()
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
            t = torch.tensor([float(np.mean(ms))], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            m = float(t.item())
            alg = 4.0 * float((tx_h.astype(np.int64) * ty_h).sum()) + 4.0 * B * T_x * T_y
            row = {"B": B, "T_text": T_x, "T_mel": T_y, "ms": m, "alignments_per_sec": B / (m / 1e3),
                   "hbm_frac_of_measured_peak": alg / (m / 1e3) / 1e9 / (peak * world)}
            if with_cpu and world == 1:
                nb = min(B, 256)
                c, cms, kind, eff = mas_cpu_rate(nb, T_x, T_y, host_cores(), 1, 2, seed=B * 7 + T_x)
                row["cpu_core_all_cores"] = c
                row["speedup"] = row["alignments_per_sec"] / c
            rows.append(row)
            del v, gathered
    return {"workload": "maximum_path sweep, global B in {64..1024} x {(128,800),(200,1000),(400,2000)}, ragged 0.7-1.0, "
                        f"batch-sharded x{world} (BASELINE configs[1]); 3 warm-ups, mean of 5 launches, 256 MB L2 flush between",
            "cpu": "compiled core.pyx, all host cores, min(B,256) utterances, 1 warm-up + mean of 2" if (with_cpu and world == 1) else None,
            "rows": rows}


def mas_gather_leg(a, dev, world, rank, dist, barrier, value, t_x, t_y):
    """SURVEY §8e at N > 1: the three ways every rank can end up with the whole batch's alignment, each timed as
    kernel + collective (+ expansion) per step, max over ranks.  B is per GPU (weak scaling, as the headline)."""
    import torch
    from matcha_b200 import alignment as al, sharding
    B, T_x, T_y = value.shape
    Bg = min(B, 128)                              # the dense forms at the full batch would need world x 3.3 GB per rank per buffer
    total = world * Bg
    v, tx, ty = value[:Bg], t_x[:Bg], t_y[:Bg]
    dur_all = torch.empty(total, T_x, dtype=torch.int32, device=dev)
    dense_all = torch.empty(total, T_x, T_y, device=dev)

    def compact():
        _, dur = al.maximum_path_with_durations(v, tx, ty)
        sharding.all_gather_durations(dur, total, out=dur_all)

    def compact_expand():
        _, dur = al.maximum_path_with_durations(v, tx, ty)
        sharding.all_gather_durations(dur, total, out=dur_all)
        sharding.path_from_durations(dur_all, T_y)

    def dense():
        path, _ = al.maximum_path_with_durations(v, tx, ty)
        dist.all_gather_into_tensor(dense_all, path)

    def kernel_only():
        al.maximum_path_with_durations(v, tx, ty)

    res = {"utterances_per_gpu": Bg, "shape": [T_x, T_y]}
    for name, fn in (("kernel_only", kernel_only), ("compact_durations", compact), ("compact_then_expand_on_every_rank", compact_expand),
                     ("dense_fp32_paths", dense)):
        for _ in range(2):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / 5], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = {"ms_per_step": float(t.item()), "alignments_per_sec": total / (float(t.item()) / 1e3)}
    res["bytes_per_rank_received"] = {"compact": int(4 * (world - 1) * Bg * T_x), "dense": int(4 * (world - 1) * Bg * T_x * T_y)}
    res["limiters"] = {"compact_durations": "NCCL launch latency (1.6 KB per utterance on the wire)",
                       "compact_then_expand_on_every_rank": "expand_kernel's HBM write of world x B x T_x x T_y fp32 on every rank",
                       "dense_fp32_paths": "NVLink: (world-1)/world x 3.2 MB per utterance into every rank"}
    res["headline_uses"] = "compact_durations (lossless for a monotonic alignment; matcha_tts.py:287,314 reduce the alignment to exactly this)"
    return res


# ------------------------------------------------------------------------------------------------ CFM decoder
def _dec_params():
    return dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)


def _time_solves(fn, K, barrier, world, dist, dev):
    """K solves between barriers, twice; per rank the better of the two repeats, then the max over ranks.  (One repeat of 5
    solves is ~45 ms: a single host hiccup on one of 8 ranks doubled the N=8 figure of one run, profiles/r2b_bench_n8.json.)"""
    import torch
    best = None
    for _ in range(2):
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(K):
            out = fn()
        ev[1].record()
        barrier()
        t = ev[0].elapsed_time(ev[1]) / K
        best = t if best is None else min(best, t)
    ms = torch.tensor([best], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), out


def decoder_leg(a, dev, world, rank, dist, barrier):
    """CFM decoder mel-frames/s (BASELINE configs[2]: 80-bin mel, T=800, Euler n=10, batch 32 per GPU,
    random-init Matcha-size decoder), "fp32 then bf16": every precision mode of the library is timed; the headline of
    the sub-dict is --dec-precision.  Beside it on the same GPU: the reference's arithmetic through PyTorch eager
    (oracle/decoder.py, the restatement pinned to the reference modules) in fp32 (TF32 off), TF32 and bf16 autocast."""
    import torch
    from types import SimpleNamespace
    from matcha_b200.flow_matching import CFM
    B, T, n = a.dec_batch, a.dec_T, a.dec_steps
    flops = 2.0 * (6017024.0 * T + 1536.0 * T * T) * B * n
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak, src = float(json.load(f)["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    except Exception:
        peak, src = 1400.0, "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)"
    g = torch.Generator(device=dev); g.manual_seed(2000 + rank)
    mu = torch.randn(B, 80, T, device=dev, generator=g)
    z = torch.randn(B, 80, T, device=dev, generator=g)
    mask = torch.ones(B, 1, T, device=dev)
    t_span = torch.linspace(0, 1, n + 1, device=dev).cpu()      # host copy, read back once: what CFM.forward hands solve_ode_euler
    modes, outs, weights, launches = {}, {}, None, 0
    order = [a.dec_precision] + [p for p in ("fp32", "bf16", "fp16") if p != a.dec_precision]
    sampler = ClockSampler(dev.index or 0)
    for prec in order:
        cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), _dec_params(), precision=prec)
        cfm.estimator.initialize_weights(seed=0)                   # Decoder.initialize_weights distribution (decoder.py:255)
        weights = cfm.estimator.state_dict()
        for _ in range(3):
            cfm.solve_ode_euler(z, t_span, mu, mask, None, None)
        K = 2 if prec == "fp32" else 5
        if prec == a.dec_precision and rank == 0:
            sampler.start()
        ms, out = _time_solves(lambda: cfm.solve_ode_euler(z, t_span, mu, mask, None, None), K, barrier, world, dist, dev)
        clocks = sampler.stop() if (prec == a.dec_precision and rank == 0) else None
        outs[prec] = out
        modes[prec] = {"ms_per_solve": ms, "mel_frames_per_sec": world * B * T / (ms / 1e3), "tflops": flops / (ms / 1e3) / 1e12,
                       "frac_of_bf16_peak": flops / (ms / 1e3) / 1e12 / peak, "timed_solves": K,
                       "gpu_launches_per_solve": int(cfm.estimator.last_launches)}
        if clocks is not None:
            modes[prec]["clocks"] = clocks
        if prec == a.dec_precision:
            launches = int(cfm.estimator.last_launches)
            head_cfm = cfm
        else:
            del cfm
            torch.cuda.empty_cache()
    ms = modes[a.dec_precision]["ms_per_solve"]
    # e2e: pinned host tensors in, host tensor out, through the reference-shaped CFM API
    mu_h, z_h, mask_h = mu.cpu().pin_memory(), z.cpu().pin_memory(), mask.cpu().pin_memory()
    out_h = torch.empty(B, 80, T).pin_memory()

    def e2e_once():
        o = head_cfm.solve_ode_euler(z_h.to(dev, non_blocking=True), t_span, mu_h.to(dev, non_blocking=True),
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
    # parity at the benchmarked configuration: samples of the first, a middle and the last sub-batch against the fp64
    # oracle (oracle/decoder.py, the checker; run on the GPU in fp64 on those samples alone -- batch items never interact)
    rel = {}
    if not a.no_cpu_baseline and rank == 0:
        from oracle import decoder as odec
        idx = sorted(set([0, B // 2, B - 1]))
        w64 = odec.cast_weights({k: v.to(dev) for k, v in weights.items()}, torch.float64)
        want = odec.solve_euler(w64, z[idx].double(), t_span, mu[idx].double(), mask[idx].double())
        for prec, o in outs.items():
            rel[prec] = float(torch.linalg.norm((o[idx].double() - want)) / torch.linalg.norm(want))
        del w64, want
    if rank != 0:
        return None
    traffic, traffic_src = traffic_record("cfm_decoder", f"B={B} T_mel={T} n_timesteps={n} {a.dec_precision}")
    res = {
        "metric": "cfm_decoder_mel_frames_per_sec", "value": world * B * T / (ms / 1e3), "unit": "mel-frames/s",
        "ms_per_solve": ms, "n_gpus": world, "scaling": "weak", "dtype": a.dec_precision, "data": "synthetic",
        "config": {"workload": f"CFM Euler solve B={B}/GPU T_mel={T} n_timesteps={n} 80-bin mel, random-init "
                               f"Matcha-size decoder (BASELINE configs[2])", "precision": a.dec_precision},
        "roofline": {"bound": "tensor", "achieved": flops / (ms / 1e3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                     "frac": flops / (ms / 1e3) / 1e12 / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": src,
                     "algorithmic_flops_per_solve": flops, "kernel": "tcgen05 kernels: gemm_bf16_kernel, gemm_abuild_kernel, ffn_fused_kernel, attention_tc_kernel"},
        "e2e": {"value": world * B * T / float(e2e_s.item()), "unit": "mel-frames/s",
                "h2d_bytes_per_step": int(4 * (2 * B * 80 * T + B * T)), "d2h_bytes_per_step": int(4 * B * 80 * T),
                "api": "matcha_b200.flow_matching.CFM.solve_ode_euler on pinned host tensors"},
        "gpu_launches_per_solve": launches,
        "by_precision": modes,
        "rel_l2_vs_oracle": rel, "rel_l2_note": "fp64 oracle on samples [0, B/2, B-1]; bars: fp32 <= 1e-5, 16-bit tensor-core mode <= 2e-3 (north_star)",
    }
    if world == 1 and not a.no_gpu_baseline and not a.no_cpu_baseline:
        res["gpu_baseline"] = decoder_gpu_baseline(weights, z, t_span, mu, mask, flops, modes)
    if world == 1 and not a.no_cpu_baseline:
        # the reference's arithmetic (torch CPU fp32, oracle/decoder.py) on a bounded sample
        from oracle import decoder as odec                     # the checker, timed as the CPU baseline only
        torch.set_num_threads(host_cores())
        bs, ns = min(4, B), 2
        xs, mus, ms_ = z[:bs].cpu(), mu[:bs].cpu(), mask[:bs].cpu()
        tsp = torch.linspace(0, 1, ns + 1)
        odec.solve_euler(weights, xs, torch.linspace(0, 1, 2), mus, ms_)      # warm-up, one eval
        t1 = time.perf_counter()
        odec.solve_euler(weights, xs, tsp, mus, ms_)
        dt = time.perf_counter() - t1
        res["cpu_baseline"] = {"value": bs * T / (dt * n / ns), "unit": "mel-frames/s", "cores": torch.get_num_threads(),
                               "kind": "port", "sample": f"torch CPU fp32 restatement of the reference decoder, B={bs}, T={T}, "
                                                         f"{ns} Euler steps timed and scaled to n={n}"}
    return res


def decoder_gpu_baseline(weights, z, t_span, mu, mask, flops, modes):
    """The "kernel to beat" (BASELINE.md §2 row 4): the reference decoder's arithmetic through PyTorch
    eager on this same B200 -- cuBLAS / cuDNN / SDPA kernels -- in fp32 with TF32 off, with TF32 on, and under bf16
    autocast.  /root/reference does not exist on the GPU box, so oracle/decoder.py (pinned to the reference modules to
    1.1e-6) stands in, with its attention through F.scaled_dot_product_attention as diffusers' AttnProcessor2_0 calls it:
    kind "port"."""
    import torch
    from oracle import decoder as odec
    dev = z.device
    w = {k: v.to(dev) for k, v in weights.items()}
    B, _, T = z.shape
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, odec.USE_SDPA)
    odec.USE_SDPA = True
    res = {"kind": "port", "what": "oracle/decoder.py (torch restatement pinned to the reference modules) on cuda through PyTorch eager, "
                                   "attention via F.scaled_dot_product_attention(attn_mask=float)",
           "timing": "same protocol as the library arm: 3 warm-up solves, better of two repeats of K solves (K = 2 for fp32, 5 otherwise), CUDA events"}

    def timed(fn, K):
        # the library arm's protocol (_time_solves): 3 warm-up solves, then the better of two repeats of K solves, CUDA events
        with torch.inference_mode():
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            ms = None
            for _ in range(2):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(K):
                    fn()
                e1.record(); torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / K
                ms = t if ms is None else min(ms, t)
        return {"ms_per_solve": ms, "mel_frames_per_sec": B * T / (ms / 1e3), "tflops": flops / (ms / 1e3) / 1e12, "timed_solves": K}

    try:
        torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False
        res["eager_fp32_tf32_off"] = timed(lambda: odec.solve_euler(w, z, t_span, mu, mask), modes["fp32"].get("timed_solves", 2))
        torch.backends.cuda.matmul.allow_tf32 = True; torch.backends.cudnn.allow_tf32 = True
        res["eager_tf32"] = timed(lambda: odec.solve_euler(w, z, t_span, mu, mask), 5)

        def bf16():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return odec.solve_euler(w, z, t_span, mu, mask)
        res["eager_bf16_autocast"] = timed(bf16, 5)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, odec.USE_SDPA = old
    res["speedup_ours_over_eager"] = {
        "fp32_mode_vs_eager_fp32": res["eager_fp32_tf32_off"]["ms_per_solve"] / modes["fp32"]["ms_per_solve"],
        "fp16_mode_vs_eager_bf16_autocast": res["eager_bf16_autocast"]["ms_per_solve"] / modes["fp16"]["ms_per_solve"],
        "bf16_mode_vs_eager_bf16_autocast": res["eager_bf16_autocast"]["ms_per_solve"] / modes["bf16"]["ms_per_solve"],
        "fp16_mode_vs_eager_tf32": res["eager_tf32"]["ms_per_solve"] / modes["fp16"]["ms_per_solve"]}
    return res


def text_encoder_leg(a, dev, world, rank, dist, barrier):
    """SURVEY §8f-4 / BASELINE configs[3] "text-encoder mu_x": TextEncoder.forward (text_encoder.py:377-402) on a synthetic
    LJSpeech-shaped batch (32 x T_text 128 per GPU, ragged 0.7-1.0), in both precisions, with the torch restatement of the
    reference module timed on the same GPU (eager fp32 with TF32 off / on, bf16 autocast) and on the host cores."""
    import torch
    from matcha_b200.text_encoder import TextEncoder
    from oracle import text_encoder as oenc            # the checker: rel-L2 of the kernels' outputs, and the baselines
    B, T = a.dec_batch, 128
    rng = np.random.default_rng(5000 + rank)
    xl_h, _ = ragged_lengths(rng, B, T, 800)
    g = torch.Generator(device=dev); g.manual_seed(5000 + rank)
    ids = torch.randint(0, 178, (B, T), device=dev, generator=g)
    xl = torch.from_numpy(xl_h).to(dev)
    encs = {p: TextEncoder(n_vocab=178, precision=p).initialize_weights(seed=1) for p in ("fp32", "fp16")}
    w32 = {k: v.to(dev) for k, v in encs["fp32"].state_dict().items()}
    with torch.no_grad():
        ref = oenc.text_encoder_forward(oenc.cast_weights(w32, torch.float64), ids, xl.long())

    def rel(x, r):
        return float(torch.linalg.norm(x.double() - r) / torch.linalg.norm(r))

    def timed(fn, n):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    modes = {}
    for prec, enc in encs.items():
        ms = timed(lambda: enc(ids, xl), 20)
        mu, logw, _ = enc(ids, xl)
        modes[prec] = {"ms_per_forward": ms, "tokens_per_sec": world * B * T / (ms / 1e3), "launches": int(enc.last_launches),
                       "rel_l2_vs_oracle_fp64": {"mu": rel(mu, ref[0]), "logw": rel(logw, ref[1])}}
    gpu_base = None
    if not a.no_gpu_baseline and world == 1:
        gpu_base = {"kind": "port", "what": "oracle/text_encoder.py (torch restatement of the reference module, pinned to it bit for bit) "
                                           "in PyTorch eager on the same GPU"}
        old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        with torch.no_grad():
            for name, tf32, ac in (("fp32", False, False), ("tf32", True, False), ("bf16_autocast", True, True)):
                torch.backends.cudnn.allow_tf32 = tf32; torch.backends.cuda.matmul.allow_tf32 = tf32
                if ac:
                    def fn():
                        with torch.autocast("cuda", dtype=torch.bfloat16):
                            return oenc.text_encoder_forward(w32, ids, xl.long())
                else:
                    fn = lambda: oenc.text_encoder_forward(w32, ids, xl.long())
                ms = timed(fn, 5)
                out = fn()
                gpu_base[name] = {"ms_per_forward": ms, "tokens_per_sec": B * T / (ms / 1e3), "rel_l2_mu_vs_oracle_fp64": rel(out[0].float(), ref[0])}
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    if rank != 0:
        return None
    best = modes["fp16"]
    res = {"metric": "text_encoder_tokens_per_sec", "value": best["tokens_per_sec"], "unit": "tokens/s", "n_gpus": world, "scaling": "weak",
           "dtype": "fp16", "data": "synthetic",
           "config": {"workload": f"TextEncoder.forward B={B}/GPU T_text={T} ragged 0.7-1.0, random-init Matcha-size encoder (7.2 M parameters)"},
           "modes": modes}
    if gpu_base is not None:
        gpu_base["speedup_fp16_vs_eager_bf16"] = gpu_base["bf16_autocast"]["ms_per_forward"] / modes["fp16"]["ms_per_forward"]
        gpu_base["speedup_fp32_vs_eager_fp32"] = gpu_base["fp32"]["ms_per_forward"] / modes["fp32"]["ms_per_forward"]
        res["gpu_baseline"] = gpu_base
    if not a.no_cpu_baseline and world == 1:
        torch.set_num_threads(host_cores())
        w_cpu = encs["fp32"].state_dict()
        with torch.no_grad():
            oenc.text_encoder_forward(w_cpu, ids[:4].cpu(), xl[:4].cpu().long())
            t1 = time.perf_counter()
            oenc.text_encoder_forward(w_cpu, ids.cpu(), xl.cpu().long())
            dt = time.perf_counter() - t1
        res["cpu_baseline"] = {"value": B * T / dt, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port",
                               "sample": f"one forward of the same batch through the torch restatement on the host ({dt * 1e3:.0f} ms)"}
    return res


def training_leg(a, dev, world, rank, dist, barrier):
    """BASELINE configs[3]: the training-step hot path, forward values: token ids -> text encoder (mu_x, logw) ->
    log-prior -> MAS -> logw_ / mu_y / prior loss -> CFM flow-matching loss forward, synthetic LJSpeech-shaped batch 32
    per GPU (T_text 128, T_mel 800, ragged 0.7-1.0)."""
    import torch
    from types import SimpleNamespace
    from matcha_b200 import alignment as al
    from matcha_b200.flow_matching import CFM
    from matcha_b200.text_encoder import TextEncoder
    B, T_x, T_y = a.dec_batch, 128, a.dec_T
    cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), _dec_params(), precision=a.dec_precision)
    cfm.estimator.initialize_weights(seed=0)
    enc = TextEncoder(n_vocab=178, precision="fp32" if a.dec_precision == "fp32" else "fp16").initialize_weights(seed=1)
    rng = np.random.default_rng(3000 + rank)
    xl_h, yl_h = ragged_lengths(rng, B, T_x, T_y)
    g = torch.Generator(device=dev); g.manual_seed(3000 + rank)
    ids = torch.randint(0, 178, (B, T_x), device=dev, generator=g)
    y = torch.randn(B, 80, T_y, device=dev, generator=g)
    xl, yl = torch.from_numpy(xl_h).to(dev), torch.from_numpy(yl_h).to(dev)
    y_mask = (torch.arange(T_y, device=dev)[None, :] < yl[:, None]).float().unsqueeze(1)

    def step(ids, y):
        mu_x, logw, x_mask = enc(ids, xl)                                         # matcha_tts.py:271 (text_encoder.py:377-402)
        r = al.align(mu_x, y, xl, yl)                                             # matcha_tts.py:277-287, 314-321
        dur_loss = torch.sum((logw - r["logw_"]) ** 2) / torch.sum(xl)            # utils/model.py:117-135
        diff_loss, _ = cfm.compute_loss(y, y_mask, r["mu_y"])                     # flow_matching.py:106-151
        return torch.stack([dur_loss, r["prior_loss"], diff_loss])

    for _ in range(3):
        losses = step(ids, y)
    barrier()
    K = 10
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(K):
        losses = step(ids, y)
    ev[1].record()
    barrier()
    ms = torch.tensor([ev[0].elapsed_time(ev[1]) / K], device=dev, dtype=torch.float64)
    mu_x, logw, _ = enc(ids, xl)
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
    split = {"text_encoder_ms": timed(lambda: enc(ids, xl)),
             "log_prior_ms": timed(lambda: al.log_prior(mu_x, y, xl, yl)),
             "mas_ms": timed(lambda: al.maximum_path_with_durations(lp, xl, yl)),
             "expand_ms": timed(lambda: al.expand(dur, mu_x, xl, yl, y=y, want_logw=True, want_prior=True)),
             "cfm_loss_ms": timed(lambda: cfm.compute_loss(y, y_mask, y))}
    # e2e: pinned host tensors in (token ids, mel), the three losses out
    ids_h, y_h = ids.cpu().pin_memory(), y.cpu().pin_memory()

    def e2e_once():
        return step(ids_h.to(dev, non_blocking=True), y_h.to(dev, non_blocking=True)).cpu()
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
           "config": {"workload": f"token ids -> text encoder ({enc.precision}) -> log-prior -> MAS -> mu_y/logw_/prior loss -> CFM loss "
                                  f"forward, B={B}/GPU T_text={T_x} T_mel={T_y} ragged 0.7-1.0 (BASELINE configs[3]); forward values, "
                                  f"backward pass out of scope"},
           "stage_ms": split, "losses": [float(v) for v in losses.cpu()],
           "e2e": {"value": world * B / float(e2e_s.item()), "unit": "utterances/s",
                   "h2d_bytes_per_step": int(8 * ids.numel() + 4 * y.numel()), "d2h_bytes_per_step": 12,
                   "api": "matcha_b200.text_encoder.TextEncoder + alignment.align + flow_matching.CFM.compute_loss on pinned host tensors"}}
    if not a.no_cpu_baseline and world == 1:
        from oracle import alignment as oal, decoder as odec       # the checker, timed as the CPU baseline only
        from oracle import mas as omas
        from oracle import text_encoder as oenc
        torch.set_num_threads(host_cores())
        omas.set_omp_threads(host_cores())
        yc = y.cpu()
        t1 = time.perf_counter()
        with torch.no_grad():
            mc = oenc.text_encoder_forward(enc.state_dict(), ids.cpu(), xl.cpu().long())[0]
        t_enc = time.perf_counter() - t1
        t1 = time.perf_counter()
        r = oal.align(mc, yc, xl.cpu(), yl.cpu(), use_ref=oal.omas.ref_available())
        t_align = time.perf_counter() - t1 + t_enc
        bs = min(4, B)
        w = cfm.estimator.state_dict()
        tt, zz = torch.rand(bs, 1, 1), torch.randn(bs, 80, T_y)
        t1 = time.perf_counter()
        odec.compute_loss(w, yc[:bs], y_mask[:bs].cpu(), r["mu_y"][:bs], tt, zz)
        t_loss = (time.perf_counter() - t1) * B / bs
        res["cpu_baseline"] = {"value": B / (t_align + t_loss), "unit": "utterances/s", "cores": torch.get_num_threads(),
                               "kind": "reference" if oal.omas.ref_available() else "port",
                               "sample": f"torch CPU restatement of matcha_tts.py:271-321: text encoder ({t_enc * 1e3:.0f} ms) + alignment with "
                                         f"the compiled core.pyx on the full batch ({t_align * 1e3:.0f} ms together) + decoder loss forward "
                                         f"on {bs} items scaled to {B}"}
    return res


def synthesis_leg(a, dev, world, rank, dist, barrier):
    """BASELINE configs[4]: batched synthesis throughput, 512 synthetic utterances per GPU (4096 over 8 GPUs) in
    batches of 32, MAS-free inference: token ids -> text encoder -> durations -> generate_path / mu_y
    (matcha_tts.py:212-229) -> CFM Euler solve with n_timesteps 2 / 4 / 10; at N > 1 each batch's mels are all-gathered
    over NCCL on a side stream while the next batch's solve runs (the gathers are inside the timed region).
    The encoder's mu_x is used as it comes; its predicted log-durations are computed and then REPLACED by synthetic
    LJSpeech-shaped ones (a random-init duration predictor says ~1.5 frames per token, i.e. T_mel ~ 200, not the 800 the
    config names)."""
    import torch
    from types import SimpleNamespace
    from matcha_b200 import alignment as al
    from matcha_b200.flow_matching import CFM
    from matcha_b200.text_encoder import TextEncoder
    n_utt, bs, T_x, T_max = a.synth_utterances, a.dec_batch, 128, a.dec_T
    cfm = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), _dec_params(), precision=a.dec_precision)
    cfm.estimator.initialize_weights(seed=0)
    enc = TextEncoder(n_vocab=178, precision="fp32" if a.dec_precision == "fp32" else "fp16").initialize_weights(seed=1)
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
        ids = torch.randint(0, 178, (bs, T_x), device=dev, generator=g)
        batches.append((ids, torch.from_numpy(xl_h).to(dev), logw, int(yl_h.sum())))
    frames = sum(b[3] for b in batches)
    nb = len(batches)
    mels = torch.zeros(nb, bs, 80, T_max, device=dev)
    gathered = torch.empty(nb, world * bs, 80, T_max, device=dev) if world > 1 else None     # [batch][rank * bs + i]
    comm = torch.cuda.Stream(device=dev) if world > 1 else None

    def run_all(n):
        main = torch.cuda.current_stream(dev)
        for i, (ids, xl, logw, _) in enumerate(batches):
            mu_x, _logw_pred, x_mask = enc(ids, xl)                               # matcha_tts.py:212
            pre = al.synthesis_prestep(mu_x, logw, x_mask, 1.0)                   # matcha_tts.py:218-229 (synthetic durations)
            out = cfm(pre["mu_y"], pre["y_mask"], n, 1.0)                         # :234 (randn + Euler solve)
            mels[i, :, :, : out.shape[-1]] = out
            if world > 1:
                comm.wait_stream(main)
                with torch.cuda.stream(comm):                                     # overlaps the next batch's solve
                    dist.all_gather_into_tensor(gathered[i], mels[i])
        if world > 1:
            main.wait_stream(comm)

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
                                   f"counted), token ids -> text encoder ({enc.precision}) -> synthetic LJSpeech-shaped durations -> "
                                   f"generate_path/mu_y -> CFM solve, per-batch all-gather of mels on a side stream at N>1 "
                                   f"(BASELINE configs[4]); vocoder out of scope"},
            "by_n_timesteps": per_n}


# ------------------------------------------------------------------------------------------------ --impl reference
def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path (compiled core.pyx from oracle/_ref) on
    the box's host cores, ALL of them: the OpenMP team size is set explicitly (torch.distributed.run exports
    OMP_NUM_THREADS=1 to every rank, which made the round-1 N >= 2 numbers single-threaded).  Same config as our arm:
    every step is the whole B-utterance batch.  Rank 0 alone runs; the other ranks exit 0."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    cores = host_cores()
    val, ms, kind, eff = mas_cpu_rate(a.batch, a.tx, a.ty, cores, a.warmup, a.steps)
    variants = {}
    nb = min(a.batch, 128)
    for name, thr, wrap in (("core_1_thread", 1, False), ("wrapper_all_cores", cores, True), ("wrapper_1_thread", 1, True)):
        v, m, _, e = mas_cpu_rate(nb, a.tx, a.ty, thr, 1, 2, wrapper=wrap)
        variants[name] = {"value": v, "ms_per_call": m, "threads": e, "utterances": nb}
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": bench_config(a),
        "notes": {"sample_per_step": f"{a.batch} utterances on the host (the whole batch of our arm's step; a CPU has no ranks, "
                                     f"so the value does not grow with --gpus)",
                  "inputs": "one block of min(B,128) N(0,1) utterances tiled to B (the DP's cost does not depend on the data); our arm draws B distinct ones"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": min(cores, a.batch), "kind": kind, "omp_threads_in_force": eff,
                         "sample": f"{a.batch} utterances per step, core.pyx compute_batch_alignments on numpy buffers, OpenMP prange over "
                                   f"the batch, {a.warmup} warm-up + mean of {a.steps} timed steps",
                         "variants": variants},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    _emit(line)


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(a):
    import torch
    import torch.distributed as dist
    from matcha_b200 import alignment as al, monotonic_align as ma, sharding
    from matcha_b200._lib import load_library
    import ctypes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))

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
    lib = load_library()
    if world > 1:
        # N > 1: the kernel also emits the per-token durations (matcha_b200_mas_align) and every step all-gathers them over
        # NCCL, inside the timed region.
        # MB200_BENCH_GATHER=overlap (default): the gather of step i runs on a HIGH-PRIORITY side stream while the kernel of
        # step i+1 runs (durations double-buffered); the persistent MAS kernel fills every SM, so the NCCL kernel's CTAs get in as the
        # first CTAs of the running kernel retire -- with the higher priority they are placed before the next kernel's CTAs.  The
        # collective's latency and the wait for the slowest rank then sit beside the kernel instead of behind it; the timed region
        # ends only when the last gather has landed.  =inorder: kernel, then gather, on one stream.
        mode = os.environ.get("MB200_BENCH_GATHER", "overlap")
        dur = [torch.empty(B, T_x, dtype=torch.int32, device=dev) for _ in range(2)]
        dur_all = [torch.empty(world * B, T_x, dtype=torch.int32, device=dev) for _ in range(2)]
        stream_ptr = lambda: ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        comm = torch.cuda.Stream(device=dev, priority=-1) if mode == "overlap" else None
        gathered = [None, None]          # event of the last gather that read dur[k]
        state = {"i": 0}

        def step():
            k = state["i"] & 1
            state["i"] += 1
            if comm is not None and gathered[k] is not None:
                torch.cuda.current_stream(dev).wait_event(gathered[k])          # buffer k is free: gather i-2 has finished
            rc = lib.matcha_b200_mas_align(path.data_ptr(), 1, dur[k].data_ptr(), value.data_ptr(), None, t_x.data_ptr(), t_y.data_ptr(),
                                           B, T_x, T_y, ctypes.c_float(-1e9), None, stream_ptr())
            assert rc == 0, rc
            if comm is None:
                dist.all_gather_into_tensor(dur_all[k], dur[k])
            else:
                comm.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(comm):
                    dist.all_gather_into_tensor(dur_all[k], dur[k])
                    gathered[k] = torch.cuda.Event()
                    gathered[k].record(comm)

        def finish():
            if comm is not None:
                torch.cuda.current_stream(dev).wait_stream(comm)
    else:
        def step():
            ma.maximum_path_from_lengths(value, t_x, t_y, out=path)

        def finish():
            pass

    for _ in range(a.warmup):
        step()
    finish()
    # the clock sampler (rank 0: NVML start-up, a few ms of host time) starts BEFORE the barrier: a rank that enters the timed
    # region late makes every other rank wait in the first step's collective, inside their timed regions
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    ev[0].record()
    for i in range(a.steps):
        step()
        if i == a.steps - 1:
            finish()
        ev[i + 1].record()
    barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.steps)]
    # the timed region is ~20 ms: keep the same kernel running for ~0.25 s more so that the clock record has enough samples
    # of the GPU under THIS load (not part of `value`); the same fixed count on every rank (the step holds a collective at N > 1)
    if not a.kernel_only:
        for _ in range(250):
            step()
        finish()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    barrier()
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

    # M1 itself: maximum_path(value, mask) with a dense mask (reads a second B x T_x x T_y tensor; lengths reduced on the device)
    m1 = None
    if not a.no_mas_extra:
        mask = torch.zeros(B, T_x, T_y, device=dev)
        ar_x, ar_y = torch.arange(T_x, device=dev), torch.arange(T_y, device=dev)
        mask = ((ar_x[None, :, None] < t_x[:, None, None]) & (ar_y[None, None, :] < t_y[:, None, None])).float()
        m1 = {}
        for mode in ("prefix", "full"):
            for _ in range(2):
                p1 = ma.maximum_path(value, mask, mask_mode=mode)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                p1 = ma.maximum_path(value, mask, mask_mode=mode)
            e1.record()
            barrier()
            tt = torch.tensor([e0.elapsed_time(e1) / 5], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            m1[mode] = {"ms_per_call": float(tt.item()), "alignments_per_sec": world * B / (float(tt.item()) / 1e3),
                        "equals_headline_path": bool(torch.equal(p1, path))}
        del mask, p1
        torch.cuda.empty_cache()

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
    del vh, ph, vn, pn

    gather = None
    if world > 1 and not a.no_mas_extra:
        gather = mas_gather_leg(a, dev, world, rank, dist, barrier, value, t_x, t_y)

    if rank == 0:
        peak, peak_src = measured_peaks()
        alg_bytes = 4.0 * float((t_x_h.astype(np.int64) * t_y_h.astype(np.int64)).sum()) + 4.0 * B * T_x * T_y
        k_ms = float(np.mean(kern_ms))
        achieved = alg_bytes / (k_ms / 1e3) / 1e9
        out = (ctypes.c_int * 8)()
        lib.matcha_b200_mas_query_launch(B, T_x, T_y, 0, 1, ctypes.byref(out))
        traffic, traffic_src = traffic_record("mas_kernel", f"B={a.batch} T_text={a.tx} T_mel={a.ty} ragged 0.7-1.0")
        h2d = int(4 * int((t_x_h.astype(np.int64) * t_y_h).sum()) + 8 * B)
        line = {
            "metric": METRIC, "value": value_rate, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": bench_config(a),
            "notes": {"launch": {"grid": out[0], "block": out[1], "smem": out[2], "rows_per_lane": out[3],
                                  "stages": out[4], "tma": out[5], "ctas_per_sm": out[6]},
                       "parallelism": (f"batch-sharded x{world}; every step all-gathers the alignment (compact: int32 durations) over NCCL, "
                                       + ("on a high-priority side stream under the next step's kernel" if os.environ.get("MB200_BENCH_GATHER", "overlap") == "overlap"
                                          else "right behind the kernel on the same stream") + ", inside the timed region") if world > 1 else "single GPU, no collective"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                         "kernel_ms": k_ms, "kernel": "mas_kernel<R,HAS_MASK=false>"},
            "e2e": {"value": e2e_rate, "unit": UNIT,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(4 * B * T_x + 4), "steps": a.e2e_steps,
                    "pcie_h2d_gb_per_s": h2d / float(te.item()) / 1e9,
                    "note": "valid [t_x][t_y] value blocks in (one cudaMemcpy2DAsync per utterance), per-token durations out (expanded into the caller's "
                            "pre-zeroed int32 paths on the host), sliced H2D / kernel / D2H pipeline; streams, events, device slabs and the pinned "
                            "durations buffer are cached in the library's host handle across calls",
                    "api": "matcha_b200.monotonic_align.compute_batch_alignments(numpy pinned) -> matcha_b200_mas_maximum_path_host",
                    "matches_device_result": same},
            "gpu_launches": a.steps, "clocks": clocks,
        }
        if m1 is not None:
            line["maximum_path_with_mask"] = m1
        if gather is not None:
            line["mas_gather"] = gather
    del path
    torch.cuda.empty_cache()
    with_cpu = (world == 1 and not a.no_cpu_baseline)
    if rank == 0 and with_cpu:
        blk = mas_cpu_baseline_block(B, T_x, T_y)
        c = blk["core_all_cores"]
        line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": min(blk["host_cores"], B), "kind": blk["kind"],
                                "sample": f"{c['utterances']} utterances of the same ragged {T_x}x{T_y} workload per call, core only "
                                          f"(compute_batch_alignments on numpy buffers), OpenMP prange over the batch, {c['threads']} threads; "
                                          + blk["protocol"],
                                "variants": {k: blk[k] for k in ("core_1_thread", "wrapper_all_cores", "wrapper_1_thread") if k in blk}}
    cfg1 = sweep = None
    if not a.no_mas_extra:
        del value
        torch.cuda.empty_cache()
        if world == 1:
            cfg1 = mas_config1_leg(dev, with_cpu)
        sweep = mas_sweep_leg(dev, world, rank, dist, barrier, with_cpu)
    dec = None
    if not a.no_decoder:
        torch.cuda.empty_cache()
        dec = decoder_leg(a, dev, world, rank, dist, barrier)
    train = synth = tenc = None
    if not a.no_decoder and not a.no_pipeline_legs:
        tenc = text_encoder_leg(a, dev, world, rank, dist, barrier)
        train = training_leg(a, dev, world, rank, dist, barrier)
        synth = synthesis_leg(a, dev, world, rank, dist, barrier)
    if rank == 0:
        if cfg1 is not None:
            line["mas_config1"] = cfg1
        if sweep is not None:
            line["mas_sweep"] = sweep
        if dec is not None:
            line["cfm_decoder"] = dec
            line["gpu_launches"] = a.steps + 5 * dec["gpu_launches_per_solve"]
        if tenc is not None:
            line["text_encoder"] = tenc
            line["gpu_launches"] += 20 * 2 * 38             # 20 timed forwards in each precision, 38 kernels each
        if train is not None:
            line["training_step"] = train
            line["gpu_launches"] += 10 * (38 + 3 + 2 + 185)   # 10 timed steps: encoder, align kernels, loss kernels, one graph-replayed eval
        if synth is not None:
            line["synthesis"] = synth
        line["summary"] = bench_summary(line)        # last key: a driver that keeps only the line's tail still sees every leg
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
