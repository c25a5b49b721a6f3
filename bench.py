#!/usr/bin/env python3
"""bench.py -- reduce_dimension(t) input GB/s of the B200 reducer path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R]

Workload (config.workload): BASELINE.json configs[1], the Sentinel-2-shaped median composite
``reduce_dimension(median,'t')`` on a (t=100, bands=4, y, x=10980) float32 cube with 30 % NaN.  The
full cube (192.9 GB) does not fit one B200's 180 GB, so a rank holds a y-strip of it resident in HBM
(default 5490 of 10980 rows = 96.4 GB) and ONE STEP is one reduction of that resident strip; with N
ranks every rank owns its own strip of the same height (weak scaling, no data-path collective: a t
reduction of a y-strip needs no exchange).  Throughput is size-independent at these sizes.

value  = input bytes of all ranks / max-over-ranks device time of the K timed steps (CUDA events)
e2e    = the same metric through the public process API with HOST (pinned) buffers: the strip
         pipeline's H2D copies, kernels and D2H copies are all inside the timed region
roofline: the median kernel's algorithmic bytes (4*R+4 per pixel) / its mean launch duration,
         against the measured HBM copy bandwidth of MEASURED_PEAKS.json
cpu_baseline / --impl reference: the reference's numpy call for this path (np.nanmedian, the oracle
         port -- the reference itself cannot be installed here: xarray/dask are not available)
         chunk-parallel on all host cores, on a bounded sample of the same workload.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T, B, X, Y_FULL = 100, 4, 10980, 10980
NAN_PROB = 0.30
SEED = 43  # 42 + config index
METRIC = "reduce_dimension(t) input GB/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=0, help="resident strip height per rank (0 = auto)")
    ap.add_argument("--e2e-rows", type=int, default=128)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ---- CPU reference arm (the oracle port of the reference's numpy path) ------------------------------------
def cpu_reference(seconds: float, reducer: str = "median"):
    """np.nanmedian over t on chunks of (T, B, rows, X), one chunk per task, all host cores.

    This is the per-chunk work dask's threaded scheduler runs for the reference (numpy releases the
    GIL).  Returns (GB/s, cores, sample description)."""
    import warnings
    from concurrent.futures import ThreadPoolExecutor

    from oracle import reducers as R
    from oracle import synth

    warnings.simplefilter("ignore", RuntimeWarning)
    cores = os.cpu_count() or 1
    rows = 2
    chunk = synth.synth_window(SEED, (T, B, Y_FULL, X), (slice(0, T), slice(0, B), slice(0, rows), slice(0, X)),
                               np.float32, NAN_PROB)
    fn = R.BY_NAME[reducer]
    t0 = time.perf_counter()
    fn(chunk, axis=0)
    per_chunk = time.perf_counter() - t0
    reps = max(1, int(round(seconds / per_chunk)))
    chunks = [chunk.copy() for _ in range(min(cores, 16))]  # distinct buffers, revisited round-robin

    def work(k):
        return fn(chunks[k % len(chunks)], axis=0).shape

    n_tasks = cores * reps
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(work, range(n_tasks)))
    dt = time.perf_counter() - t0
    gbs = n_tasks * chunk.nbytes / dt / 1e9
    sample = (f"np.nan{reducer} over t on {n_tasks} chunks of ({T},{B},{rows},{X}) fp32, 30% NaN "
              f"({n_tasks * chunk.nbytes / 1e9:.2f} GB) on {cores} threads in {dt:.1f} s")
    return gbs, cores, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(args.warmup and 1):
        cpu_reference(min(2.0, args.cpu_seconds))
    sample = ""
    cores = os.cpu_count() or 1
    t_total = 0.0
    for _ in range(max(1, args.steps)):
        t0 = time.perf_counter()
        gbs, cores, sample = cpu_reference(args.cpu_seconds / max(1, args.steps) + 1.0)
        t_total += time.perf_counter() - t0
        vals.append(gbs)
    v = float(np.median(vals))
    line = {
        "metric": METRIC, "value": v, "unit": "GB/s", "impl": "reference", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.rows or 5490, args.gpus),
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(rows, n):
    return {
        "workload": "BASELINE configs[1]: reduce_dimension(median,'t'), t=100,bands=4,x=10980 fp32, 30% NaN; "
                    f"resident y-strip of {rows} rows per GPU (full cube 10980 rows = 192.9 GB > 180 GB HBM)",
        "reducer": "median", "layout": "(t,bands,y,x) C-order", "t": T, "bands": B, "rows_per_gpu": rows, "x": X,
        "nan_fraction": NAN_PROB, "seed": SEED, "parallelism": f"y-strips x{n}, no collective",
        "cache": "inputs larger than L2 (strip >> 126 MB), no flush needed",
    }


# ---- clocks ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def known_traffic(reducer):
    """dram bytes per launch from the committed ncu --set full capture, if one exists for this config."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(reducer)
    except Exception:
        return None


# ---- GPU arm ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback for the reducer kernels)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    # resident strip: default half the cube, clipped to what fits beside the output and a safety margin
    free, _total = torch.cuda.mem_get_info()
    row_bytes = T * B * X * 4
    rows = args.rows or min(5490, int(free * 0.80) // row_bytes)
    y_full = max(Y_FULL, rows * world)
    strip = torch.empty((T, B, rows, X), dtype=torch.float32, device=dev)
    engine.synth_fill(strip.view(T * B, rows * X), T * B, y_full * X, rank * rows * X, rows * X, SEED, NAN_PROB)
    torch.cuda.synchronize()
    in_bytes = strip.numel() * 4
    pixels = B * rows * X

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        engine.launch_count(reset=True)
        t0 = time.perf_counter()
        for a, b in ev:
            a.record()
            fn()
            b.record()
        barrier()
        wall = time.perf_counter() - t0
        per = np.array([a.elapsed_time(b) for a, b in ev])  # ms, device time of each step
        span = ev[0][0].elapsed_time(ev[-1][1])             # ms, first start -> last end on the device
        return per, span, wall, engine.launch_count()

    with ClockSampler(local) as clocks:
        per, span, wall, launches = timed(lambda: P.median(strip, axis=0), args.steps, args.warmup)
    t = torch.tensor([span], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    span_max = float(t.item())
    value = world * in_bytes * args.steps / (span_max * 1e-3) / 1e9

    # the dominant kernel on its own: algorithmic bytes / mean launch duration (events bracket one launch)
    peak, peak_src = measured_peak()
    alg_bytes = pixels * (4 * T + 4)
    achieved = alg_bytes / (float(per.mean()) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": known_traffic("median"), "peak_source": peak_src, "kernel": "k_select_regs<float,100>",
                "algorithmic_bytes_per_launch": alg_bytes}

    # secondary reducers on the same resident strip (not part of the headline timed region)
    extra = {}
    for name, fn, out_b in (("mean", lambda: P.mean(strip, axis=0), 4), ("sd", lambda: P.sd(strip, axis=0), 4),
                            ("max", lambda: P._max(strip, axis=0), 4),
                            ("quantiles_0.1_0.5_0.9", lambda: P.quantiles(strip, probabilities=[0.1, 0.5, 0.9], axis=0), 24)):
        p2, s2, _, _ = timed(fn, max(3, args.steps // 2), 2)
        t2 = torch.tensor([s2 / len(p2)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        ms = float(t2.item())
        extra[name] = {"input_gbs": world * in_bytes / (ms * 1e-3) / 1e9, "ms_per_step": ms,
                       "hbm_frac": (pixels * (4 * T + out_b)) / (float(p2.mean()) * 1e-3) / 1e9 / peak}

    # end to end through the process API with pinned HOST buffers (strip pipeline inside the call)
    e2e = None
    if not args.no_e2e:
        er = min(args.e2e_rows, rows)
        host = torch.empty((T, B, er, X), dtype=torch.float32, pin_memory=True)
        host.copy_(strip[:, :, :er, :])
        torch.cuda.synchronize()
        harr = host.numpy()
        for _ in range(2):
            P.median(harr, axis=0)
        barrier()
        k = max(3, args.steps // 2)
        t0 = time.perf_counter()
        for _ in range(k):
            out = P.median(harr, axis=0)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * harr.nbytes * k / float(dt.item()) / 1e9, "unit": "GB/s",
               "h2d_bytes_per_step": int(harr.nbytes), "d2h_bytes_per_step": int(out.nbytes),
               "sample": f"median over t of a pinned host strip ({T},{B},{er},{X}) fp32 per rank, numpy in -> numpy out"}
        del host

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, cores, sample = cpu_reference(args.cpu_seconds)
        cpu = {"value": v, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": span_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(rows, world), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks.summary(),
            "pct_hbm_peak": 100.0 * value / world / peak,
            "step_ms": {"median": float(np.median(per)), "mean": float(per.mean()), "min": float(per.min()),
                        "max": float(per.max())},
            "reducers": extra, "wall_s_timed_region": wall,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
