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
import functools
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
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary workloads")
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
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed region runs."""

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
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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


def known_traffic(reducer, alg_bytes_per_launch, t=T):
    """dram bytes per launch for this bench's launch size, scaled from the committed `ncu --set full` capture
    of the same kernel on a smaller strip of the same cube (profiles/traffic.json, written by
    tools/ncu_summary.py): traffic = captured dram bytes / captured algorithmic bytes * this launch's
    algorithmic bytes.  None when no capture is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            cap = json.load(f)[reducer]
        rows = int(cap["plain_run"].split("rows=")[1].split()[0])
        out_b = {"median": 4, "mean": 4, "sd": 4, "quantiles": 24, "quant_sd": 28}.get(reducer, 4)
        cap_alg = rows * X * B * (4 * t + out_b)
        return (cap["dram_read_bytes"] + cap["dram_write_bytes"]) / cap_alg * alg_bytes_per_launch
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
        # rank 0's stdout carries ONE JSON line: keep NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION) off it
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
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
                "traffic": known_traffic("median", alg_bytes), "traffic_source": "profiles/traffic.json (ncu --set full "
                "capture of the same kernel on a 512-row strip of the same cube, scaled by algorithmic bytes)",
                "peak_source": peak_src, "kernel": "k_select_tile<100,median,256 columns> (TMA-staged, lockstep CTA)",
                "algorithmic_bytes_per_launch": alg_bytes}

    # secondary workloads (not part of the headline timed region): the other reducers on the same resident
    # strip, and BASELINE configs[2..4] on strips of their own cubes
    extra = {}

    def side(name, fn, in_b, alg_b, note=None):
        p2, s2, _, _ = timed(fn, max(3, args.steps // 2), 2)
        t2 = torch.tensor([s2 / len(p2)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        ms = float(t2.item())
        extra[name] = {"input_gbs": world * in_b / (ms * 1e-3) / 1e9, "ms_per_step": ms,
                       "hbm_frac": alg_b / (float(p2.mean()) * 1e-3) / 1e9 / peak}
        if note:
            extra[name]["note"] = note

    if not args.no_extras:
        from openeo_processes_dask_b200.process_implementations.math import reduce_many

        q3 = [0.1, 0.5, 0.9]
        ref = P.ParameterReference(from_parameter="data")
        side("mean", lambda: P.mean(strip, axis=0), in_bytes, pixels * (4 * T + 4))
        side("sd", lambda: P.sd(strip, axis=0), in_bytes, pixels * (4 * T + 4), "TMA tile kernel, single HBM pass")
        side("max", lambda: P._max(strip, axis=0), in_bytes, pixels * (4 * T + 4))
        side("quantiles_0.1_0.5_0.9", lambda: P.quantiles(strip, probabilities=q3, axis=0), in_bytes,
             pixels * (4 * T + 24), "float64 quantiles like numpy")
        side("config2_quantiles+sd_fused", lambda: reduce_many(strip, [("quantiles", {"probabilities": q3}), "sd"], axis=0),
             in_bytes, pixels * (4 * T + 28), "BASELINE configs[2] per GPU: one pass (eoc_select_fused)")
        # SURVEY 8f rows on the same strip: cloud mask -> NaN -> median in one pass, cumulative sum, sort along t
        try:
            r8 = min(1024, rows)
            sub = strip[:, :, :r8, :].contiguous()   # a cube of its own (a row slice of the strip is not one view)
            gen = torch.Generator(device=dev).manual_seed(5 + rank)
            cloud = torch.rand((T, r8, X), device=dev, generator=gen) < 0.25
            cube8 = P.DataCube(sub, ("t", "bands", "y", "x"), {"t": list(range(T))})
            mcube = P.DataCube(cloud, ("t", "y", "x"), {"t": list(range(T))})
            px8 = B * r8 * X
            side("next_mask_then_median_fused", lambda: P.median(P.mask(data=cube8, mask=mcube, fuse=True).data, axis=0),
                 sub.numel() * 4, px8 * (4 * T + 4) + T * r8 * X,
                 "(t,y,x) uint8 cloud mask + (t,bands,y,x) cube -> median over t, one pass (eoc_masked_select)")
            side("next_cumsum_t", lambda: P.cumsum(sub, axis=0), sub.numel() * 4, px8 * 8 * T,
                 "apply_dimension(cumsum): read + write of the cube (eoc_scan)")
            side("next_sort_t", lambda: P.sort(sub, nodata=True, axis=0), sub.numel() * 4, px8 * 8 * T,
                 "apply_dimension(sort): read + write of the cube (eoc_sort)")
            side("next_first_t", lambda: P.first(sub, axis=0), sub.numel() * 4, px8 * (4 * T + 4))
            # merge_cubes overlap resolver x + y on two cubes of T/2 time steps, and the
            # (y,x)-plane-over-cube broadcast of its example 4: input GB/s counts both operands
            half = T // 2
            c_a = P.DataCube(sub[:half], ("t", "bands", "y", "x"), {"t": list(range(half))})
            c_b = P.DataCube(sub[half:], ("t", "bands", "y", "x"), {"t": list(range(half))})
            plane = P.DataCube(sub[0, 0], ("y", "x"))
            refx, refy = P.ParameterReference(from_parameter="x"), P.ParameterReference(from_parameter="y")
            add_xy = functools.partial(P.process(P.add), x=refx, y=refy)
            side("next_merge_cubes_add", lambda: P.merge_cubes(c_a, c_b, overlap_resolver=add_xy).data, sub.numel() * 4,
                 sub.numel() * 4 + half * px8 * 4, "merge_cubes(overlap_resolver=add): x + y, one launch, no stacked copy (eoc_binary)")
            side("next_merge_cubes_plane", lambda: P.merge_cubes(c_a, plane, overlap_resolver=add_xy).data, half * px8 * 4,
                 half * px8 * 8 + r8 * X * 4, "merge_cubes example 4: (y,x) plane broadcast over the cube by stride 0")
            del c_a, c_b, plane
            del cloud, cube8, mcube, sub
        except Exception as e:  # pragma: no cover
            extra["next_rows"] = {"error": repr(e)[:200]}
        # configs[3]: ndvi(nir, red) fused with reduce_dimension(max, 't') on a uint16 cube, t=200, 12 bands
        try:
            t4, b4, r4 = 200, 12, min(1024, rows)
            u16 = torch.empty((t4, b4, r4, X), dtype=torch.uint16, device=dev)
            engine.synth_fill(u16.view(t4 * b4, r4 * X), t4 * b4, Y_FULL * X, rank * r4 * X, r4 * X, SEED + 2, 0.0)
            cube4 = P.DataCube(u16, ("t", "bands", "y", "x"), {"bands": [f"B{k:02d}" for k in range(1, 13)]})
            red_max = functools.partial(P.process(P._max), data=ref)

            def ndvi_max():
                nd = P.ndvi(cube4, nir="B08", red="B04", fuse=True)
                return P.reduce_dimension(data=nd, reducer=red_max, dimension="t")

            px4 = r4 * X
            side("config3_ndvi_max_uint16", ndvi_max, 2 * t4 * px4 * 2, px4 * (2 * 2 * t4 + 4),
                 "two of twelve uint16 band planes read, NDVI never materialised (eoc_nd_reduce)")
            del u16, cube4
        except Exception as e:  # pragma: no cover
            extra["config3_ndvi_max_uint16"] = {"error": repr(e)[:200]}
        # configs[4]: aggregate_temporal, 12 monthly intervals, daily t=365, bands=4, x=5490, 40 % NaN
        try:
            import pandas as pd

            t5, b5, x5, r5 = 365, 4, 5490, min(1024, rows)
            c5 = torch.empty((t5, b5, r5, x5), dtype=torch.float32, device=dev)
            engine.synth_fill(c5.view(t5 * b5, r5 * x5), t5 * b5, x5 * x5, rank * r5 * x5, r5 * x5, SEED + 3, 0.40)
            days = pd.date_range("2021-01-01T12:00:00", periods=t5, freq="D").values
            months = pd.date_range("2021-01-01", "2022-01-01", freq="MS").strftime("%Y-%m-%dT%H:%M:%S")
            intervals = [[a, b] for a, b in zip(months[:-1], months[1:])]
            cube5 = P.DataCube(c5, ("t", "bands", "y", "x"), {"t": days})
            for nm, fn_, ob in (("mean", P.mean, 4), ("count", P.count, 8)):
                red = functools.partial(P.process(fn_), data=ref)
                side(f"config4_aggregate_temporal_{nm}",
                     lambda red=red: P.aggregate_temporal(data=cube5, intervals=intervals, reducer=red),
                     c5.numel() * 4, b5 * r5 * x5 * (4 * t5 + 12 * ob), "segmented reduce (eoc_segmented_reduce)")
            del c5, cube5
        except Exception as e:  # pragma: no cover
            extra["config4_aggregate_temporal"] = {"error": repr(e)[:200]}
        torch.cuda.empty_cache()

    # end to end through the process API with pinned HOST buffers (strip pipeline inside the call)
    e2e = None
    if not args.no_e2e:
        er = min(args.e2e_rows, rows)
        # one process per GPU: allocate (first-touch) this rank's host strip on the GPU's own memory node
        from openeo_processes_dask_b200.sharding import bind_rank_to_gpu_numa

        numa = bind_rank_to_gpu_numa(local) if world > 1 else {"bound": False}
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
               "sample": f"median over t of a pinned host strip ({T},{B},{er},{X}) fp32 per rank, numpy in -> numpy out",
               "numa": numa}
        if world == 1:  # informational: the same call on a PAGEABLE copy of the cube (an ordinary numpy chunk)
            parr = np.array(harr)
            P.median(parr, axis=0)
            t0 = time.perf_counter()
            for _ in range(3):
                P.median(parr, axis=0)
            e2e["pageable_value"] = parr.nbytes * 3 / (time.perf_counter() - t0) / 1e9
            del parr
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
    # stdout carries exactly ONE JSON line.  Libraries write banners to file descriptor 1 behind python's back
    # (NCCL prints "NCCL version ..." at every NCCL_DEBUG level from VERSION up, WARN included), so fd 1 points at
    # stderr while the benchmark runs and sys.stdout keeps a private duplicate of the real stdout.
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
