#!/usr/bin/env python3
"""bench.py -- reduce_dimension(t) input GB/s of the B200 reducer path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R]

Headline workload (config.workload): BASELINE.json configs[1], the Sentinel-2-shaped median composite
``reduce_dimension(median,'t')`` on a (t=100, bands=4, y, x=10980) float32 cube with 30 % NaN.  The full cube
(192.9 GB) does not fit one B200's 180 GB, so a rank holds a y-strip of it resident in HBM (default 5490 of 10980
rows = 96.4 GB) and ONE STEP is one reduction of that resident strip; with N ranks every rank owns its own strip of
the same height (weak scaling, no data-path collective: a t reduction of a y-strip needs no exchange).

value     = input bytes of all ranks / max-over-ranks device time of the K timed steps (CUDA events)
e2e       = the same metric through the public process API with HOST (pinned) buffers: the strip pipeline's H2D
            copies, kernels and D2H copies are all inside the timed region; next to it the raw pinned H2D rate
            measured in the same run on the same ranks (the ceiling any host-fed path has on this box)
roofline  = the median kernel's algorithmic bytes (4*R+4 per pixel) / its mean launch duration, against the measured
            HBM copy bandwidth of MEASURED_PEAKS.json
configs   = EVERY BASELINE.json config run AS CONFIGURED: the full cube is cut into y-strips with
            views.strip_bounds(rows, N, rank) (strong scaling; when a rank's share does not fit its HBM the share
            is processed in equal passes and the entry says so), each verified in the same run: >= 64 random
            128-pixel windows are regenerated from the hash on the host (oracle/synth.py), reduced by the oracle and
            compared with what the GPU wrote (`parity_checked`); a mismatch makes the bench exit non-zero
cpu_baseline / --impl reference: the reference's numpy calls for this path (oracle port -- the reference package
            itself is probed for and cannot be imported here: xarray/dask are not available) chunk-parallel on all
            host cores over (T, 1, 512, 512)-pixel chunks, on a bounded sample of the same workload.
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
Q3 = [0.1, 0.5, 0.9]
RTOL = 1e-6  # BASELINE.json north_star: mean / sd / interpolated quantiles within 1e-6 relative


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
    ap.add_argument("--windows", type=int, default=64, help="parity windows per verified entry")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary reducers and the SURVEY 8f rows")
    ap.add_argument("--no-configs", action="store_true", help="skip the as-configured runs of BASELINE configs[0..4]")
    return ap.parse_args()


# ---- CPU reference arm (the oracle port of the reference's numpy path) ------------------------------------
def probe_real_reference():
    """probe, don't assume.  Returns (usable, reason)."""
    missing = []
    for mod in ("dask", "xarray", "openeo_pg_parser_networkx", "openeo_processes_dask"):
        try:
            if mod == "dask" and getattr(sys.modules.get("dask.array"), "__eoc_shim__", False):
                raise ImportError("test shim, not dask")
            __import__(mod)
        except Exception as e:  # ImportError, or the package's own import-time failures
            missing.append(f"{mod}: {type(e).__name__}: {str(e)[:80]}")
    return (not missing), "; ".join(missing)


CPU_CALLS = {  # reducer -> (oracle name, reference call site)
    "median": ("median", "np.nanmedian (math.py:186)"), "mean": ("mean", "np.nanmean (math.py:193)"),
    "sd": ("sd", "np.nanstd ddof=1 (math.py:200)"), "max": ("max", "np.nanmax (math.py:179)"),
    "quantiles": ("quantiles", "np.nanquantile linear (math.py:377-379)"),
}


def cpu_reference(seconds: float, reducer: str = "median", t: int = T, nan_prob: float = NAN_PROB, seed: int = SEED,
                  full_shape=None):
    """The reference's numpy call over t on spatial chunks of (t, 1, 512, 512) pixels, one chunk per task, all host
    cores: the per-chunk work dask's threaded scheduler runs for the reference (numpy releases the GIL).  The chunk
    edge is what BASELINE.md section 3 names; one band per chunk keeps a chunk at 105 MB (np.nanmedian works on
    copies: four bands per chunk on 32 threads would need > 40 GB of host memory).  When one round of full chunks
    would exceed the time budget the chunk height is cut (throughput does not depend on it past the caches).
    Returns (GB/s, cores, sample description)."""
    import warnings
    from concurrent.futures import ThreadPoolExecutor

    from oracle import reducers as R
    from oracle import synth

    warnings.simplefilter("ignore", RuntimeWarning)
    cores = os.cpu_count() or 1
    full_shape = full_shape or (t, B, Y_FULL, X)
    name = CPU_CALLS[reducer][0]
    fn = R.BY_NAME[name] if name != "quantiles" else functools.partial(R.quantiles, probabilities=Q3)
    probe = synth.synth_window(seed, full_shape, (slice(0, t), slice(0, 1), slice(0, 8), slice(0, 512)),
                               np.float32, nan_prob)
    t0 = time.perf_counter()
    fn(probe, axis=0)
    per_row = (time.perf_counter() - t0) / 8  # seconds per 512-pixel row of one band on one core
    rows = int(min(512, max(8, seconds / max(per_row, 1e-9))))
    rows -= rows % 8
    chunk = synth.synth_window(seed, full_shape, (slice(0, t), slice(0, 1), slice(0, rows), slice(0, 512)),
                               np.float32, nan_prob)
    reps = max(1, int(round(seconds / (per_row * rows))))
    chunks = [chunk.copy() for _ in range(min(cores, 16))]  # distinct buffers, revisited round-robin

    def work(k):
        return np.shape(fn(chunks[k % len(chunks)], axis=0))

    n_tasks = cores * reps
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(work, range(n_tasks)))
    dt = time.perf_counter() - t0
    gbs = n_tasks * chunk.nbytes / dt / 1e9
    sample = (f"{CPU_CALLS[reducer][1]} over t on {n_tasks} chunks of ({t},1,{rows},512) fp32, {int(nan_prob * 100)}% NaN "
              f"({n_tasks * chunk.nbytes / 1e9:.2f} GB) on {cores} threads in {dt:.1f} s")
    return gbs, cores, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    real, why = probe_real_reference()
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
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample,
                         "real_reference": real, "real_reference_probe": why},
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


# ---- parity in the same run ---------------------------------------------------------------------------
class Parity:
    """>= `n` random 128-pixel windows of a result, regenerated from the hash on the host, reduced by the oracle and
    compared with what the GPU wrote.  exact: bit-equal values and NaN masks; else rtol (north_star tolerance)."""

    def __init__(self, n, rank):
        self.n = n
        self.rng = np.random.default_rng(20261105 + rank)
        self.failures = []

    def check(self, name, seed, full_shape, dtype, nan_prob, row0, rows, bands, oracle_fn, fetch, exact, width=128):
        """oracle_fn(window (t, nb, 1, w)) -> array whose LAST axis is the window's x; fetch(y, x0, w) -> the GPU's
        values for the same pixels (numpy, same shape)."""
        from oracle import synth

        t_len, x_len = full_shape[0], full_shape[3]
        max_rel, same_mask, checked = 0.0, True, 0
        for _ in range(self.n):
            y = int(self.rng.integers(0, rows))
            x0 = int(self.rng.integers(0, max(1, x_len - width)))
            w = min(width, x_len - x0)
            parts = [synth.synth_window(seed, full_shape, (slice(0, t_len), slice(b, b + 1), slice(row0 + y, row0 + y + 1),
                                                           slice(x0, x0 + w)), dtype, nan_prob) for b in bands]
            win = np.concatenate(parts, axis=1)
            with np.errstate(all="ignore"):
                want = np.asarray(oracle_fn(win))
            got = np.asarray(fetch(y, x0, w))
            if got.shape != want.shape:
                self.failures.append(f"{name}: shape {got.shape} vs oracle {want.shape}")
                return {"windows": checked, "ok": False}
            wn, gn = np.isnan(want.astype(np.float64)), np.isnan(got.astype(np.float64))
            same_mask &= bool(np.array_equal(wn, gn))
            ok = ~wn & ~gn
            if exact:
                good = np.array_equal(want[ok], got[ok])
            else:
                wv, gv = want[ok].astype(np.float64), got[ok].astype(np.float64)
                finite = np.isfinite(wv)
                rel = np.abs(gv - wv)[finite] / np.maximum(np.abs(wv[finite]), 1e-300)
                max_rel = max(max_rel, float(rel.max()) if rel.size else 0.0)
                good = (not rel.size or rel.max() <= RTOL) and np.array_equal(wv[~finite], gv[~finite])
            if not (good and same_mask):
                self.failures.append(f"{name}: window y={row0 + y} x0={x0} differs from the oracle")
                return {"windows": checked + 1, "ok": False, "nan_mask_equal": same_mask}
            checked += 1
        out = {"windows": checked, "ok": True, "nan_mask_equal": same_mask, "bar": "bit-exact" if exact else f"rtol {RTOL}"}
        if not exact:
            out["max_rel"] = max_rel
        return out


# ---- GPU arm ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.views import strip_bounds
    from oracle import cubes as OC
    from oracle import reducers as R

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

    def over_ranks(x, op):
        t_ = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_, op=op)
        return float(t_.item())

    def max_over_ranks(x):
        return over_ranks(x, dist.ReduceOp.MAX)

    def sum_over_ranks(x):
        return over_ranks(x, dist.ReduceOp.SUM)

    parity = Parity(args.windows, rank)
    ref = P.ParameterReference(from_parameter="data")
    peak, peak_src = measured_peak()

    def host(a):
        a = getattr(a, "tensor", a)
        return a.cpu().numpy() if hasattr(a, "cpu") else np.asarray(a)

    def timed(fn, steps, warmup):
        out = None
        for _ in range(warmup):
            out = None
            out = fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        engine.launch_count(reset=True)
        t0 = time.perf_counter()
        for a, b in ev:
            out = None  # release the previous result first: two live outputs of 9-18 GB push the allocator to re-malloc
            a.record()
            out = fn()
            b.record()
        barrier()
        wall = time.perf_counter() - t0
        per = np.array([a.elapsed_time(b) for a, b in ev])  # ms, device time of each step
        span = ev[0][0].elapsed_time(ev[-1][1])             # ms, first start -> last end on the device
        return per, span, wall, engine.launch_count(), out

    # ---- headline: resident strip, default half the cube, clipped to what fits beside the output -----------------
    free, _total = torch.cuda.mem_get_info()
    row_bytes = T * B * X * 4
    rows = args.rows or min(5490, int(free * 0.80) // row_bytes)
    y_full = max(Y_FULL, rows * world)
    row0 = rank * rows
    strip = torch.empty((T, B, rows, X), dtype=torch.float32, device=dev)
    engine.synth_fill(strip.view(T * B, rows * X), T * B, y_full * X, row0 * X, rows * X, SEED, NAN_PROB)
    torch.cuda.synchronize()
    in_bytes = strip.numel() * 4
    pixels = B * rows * X
    full2 = (T, B, y_full, X)

    with ClockSampler(local) as clocks:
        per, span, wall, launches, med = timed(lambda: P.median(strip, axis=0), args.steps, args.warmup)
    span_max = max_over_ranks(span)
    value = world * in_bytes * args.steps / (span_max * 1e-3) / 1e9
    head_parity = parity.check("median", SEED, full2, np.float32, NAN_PROB, row0, rows, range(B),
                               lambda w: R.median(w, axis=0)[:, 0], lambda y, x0, w: host(med[:, y, x0:x0 + w]), True)

    # the dominant kernel on its own: algorithmic bytes / mean launch duration (events bracket one launch)
    alg_bytes = pixels * (4 * T + 4)
    achieved = alg_bytes / (float(per.mean()) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": known_traffic("median", alg_bytes), "traffic_source": "profiles/traffic.json (ncu --set full "
                "capture of the same kernel on a 512-row strip of the same cube, scaled by algorithmic bytes)",
                "peak_source": peak_src, "kernel": "k_select_quad<25,128 columns> (TMA-staged, four raw-sorted runs of 25 + PTX rank elimination)",
                "algorithmic_bytes_per_launch": alg_bytes}

    extra = {}

    def side(name, fn, in_b, alg_b, note=None, check=None):
        p2, s2, _, _, out = timed(fn, max(3, args.steps // 2), 2)
        ms = max_over_ranks(s2 / len(p2))
        extra[name] = {"input_gbs": sum_over_ranks(in_b) / (ms * 1e-3) / 1e9, "ms_per_step": ms,
                       "hbm_frac": alg_b / (float(p2.mean()) * 1e-3) / 1e9 / peak}
        if note:
            extra[name]["note"] = note
        if check is not None:
            extra[name]["parity_checked"] = check(out)
        return out

    def strip_check(name, oracle_fn, exact, pick=lambda o: o):
        def run(out):
            o = pick(out)
            lead = (slice(None),) * (o.ndim - 3)
            return parity.check(name, SEED, full2, np.float32, NAN_PROB, row0, rows, range(B),
                                lambda w: np.asarray(oracle_fn(w))[..., 0, :],
                                lambda y, x0, w: host(o[lead + (slice(None), y, slice(x0, x0 + w))]), exact)
        return run

    if not args.no_extras:
        from openeo_processes_dask_b200.process_implementations.math import reduce_many

        side("mean", lambda: P.mean(strip, axis=0), in_bytes, pixels * (4 * T + 4),
             check=strip_check("mean", lambda w: R.mean(w, axis=0), True))
        side("sd", lambda: P.sd(strip, axis=0), in_bytes, pixels * (4 * T + 4), "TMA tile kernel, single HBM pass",
             check=strip_check("sd", lambda w: R.sd(w, axis=0), True))
        side("max", lambda: P._max(strip, axis=0), in_bytes, pixels * (4 * T + 4),
             check=strip_check("max", lambda w: R._max(w, axis=0), True))
        side("quantiles_0.1_0.5_0.9", lambda: P.quantiles(strip, probabilities=Q3, axis=0), in_bytes,
             pixels * (4 * T + 24), "float64 quantiles like numpy (k_select_quad: four sorted runs + rank cursor)",
             check=strip_check("quantiles", lambda w: R.quantiles(w, probabilities=Q3, axis=0), False))
        side("quantiles+sd_one_pass_4band_strip",
             lambda: reduce_many(strip, [("quantiles", {"probabilities": Q3}), "sd"], axis=0),
             in_bytes, pixels * (4 * T + 28), "eoc_select_fused on the headline strip (configs[2] proper is under `configs`)",
             check=strip_check("quantiles+sd(sd)", lambda w: R.sd(w, axis=0), True, pick=lambda o: o[1]))
        # SURVEY 8f rows on the same strip: cloud mask -> NaN -> median in one pass, cumulative sum, sort along t
        try:
            r8 = min(1024, rows)
            sub = strip[:, :, :r8, :].contiguous()   # a cube of its own (a row slice of the strip is not one view)
            gen = torch.Generator(device=dev).manual_seed(5 + rank)
            cloud = torch.rand((T, r8, X), device=dev, generator=gen) < 0.25
            cube8 = P.DataCube(sub, ("t", "bands", "y", "x"), {"t": list(range(T))})
            mcube = P.DataCube(cloud, ("t", "y", "x"), {"t": list(range(T))})
            px8 = B * r8 * X
            side("next_mask_then_median_fused", lambda: P.median(P.mask(data=cube8, mask=mcube).data, axis=0),
                 sub.numel() * 4, px8 * (4 * T + 4) + T * r8 * X,
                 "(t,y,x) uint8 cloud mask + (t,bands,y,x) cube -> median over t, one pass (eoc_masked_select)")
            side("next_cumsum_t", lambda: P.cumsum(sub, axis=0), sub.numel() * 4, px8 * 8 * T,
                 "apply_dimension(cumsum): read + write of the cube (eoc_scan)")
            side("next_sort_t", lambda: P.sort(sub, nodata=True, axis=0), sub.numel() * 4, px8 * 8 * T,
                 "apply_dimension(sort): read + write of the cube (eoc_sort)")
            side("next_first_t", lambda: P.first(sub, axis=0), sub.numel() * 4, px8 * (4 * T + 4))
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
            del c_a, c_b, plane, cloud, cube8, mcube, sub
        except Exception as e:  # pragma: no cover
            extra["next_rows"] = {"error": repr(e)[:200]}
        # long columns (a year of daily scenes: north_star "selection over T <= 512") and gap filling along t, each on a
        # synthetic cube of its own, verified on windows like everything else
        try:
            for tl in (200, 365):
                rl = max(64, min(512, int(8e9 // (tl * B * X * 4))))
                cl = torch.empty((tl, B, rl, X), dtype=torch.float32, device=dev)
                engine.synth_fill(cl.view(tl * B, rl * X), tl * B, y_full * X, row0 * X, rl * X, SEED + tl, NAN_PROB)
                side(f"median_t{tl}_long_columns", lambda: P.median(cl, axis=0), cl.numel() * 4, B * rl * X * (4 * tl + 4),
                     "k_select_multi: eight sorted runs with NaN sentinels, several threads per column",
                     check=lambda o, tl=tl, rl=rl: parity.check(
                         f"median t={tl}", SEED + tl, (tl, B, y_full, X), np.float32, NAN_PROB, row0, rl, range(B),
                         lambda w: R.median(w, axis=0)[:, 0], lambda y, x0, w: host(o[:, y, x0:x0 + w]), True))
                del cl
            # raw digital numbers (Sentinel-2 L2A is uint16) and float64 cubes on the four-run kernels
            for dt, tdt, nanp in ((np.uint16, torch.uint16, 0.0), (np.float64, torch.float64, NAN_PROB)):
                rd = min(256, rows)
                cd = torch.empty((T, B, rd, X), dtype=tdt, device=dev)
                sd_ = SEED + (16 if dt is np.uint16 else 64)
                engine.synth_fill(cd.view(T * B, rd * X), T * B, y_full * X, row0 * X, rd * X, sd_, nanp)
                es = np.dtype(dt).itemsize
                side(f"median_{np.dtype(dt).name}_t{T}", lambda: P.median(cd, axis=0), cd.numel() * es,
                     B * rd * X * (es * T + 8), "k_select_quad<uint16> / k_select_quad64: float64 results like numpy's",
                     check=lambda o, dt=dt, sd_=sd_, nanp=nanp, rd=rd: parity.check(
                         f"median {np.dtype(dt).name}", sd_, (T, B, y_full, X), dt, nanp, row0, rd, range(B),
                         lambda w: R.median(w, axis=0)[:, 0], lambda y, x0, w: host(o[:, y, x0:x0 + w]), True))
                del cd
            ri = min(512, rows)
            ci = strip[:, :, :ri, :].contiguous()
            side("next_interpolate_linear_t", lambda: P.array_interpolate_linear(ci, axis=0), ci.numel() * 4,
                 2 * ci.numel() * 4, "apply_dimension(array_interpolate_linear): read + write of the cube (k_interp_tma)",
                 check=lambda o: parity.check(
                     "interpolate", SEED, full2, np.float32, NAN_PROB, row0, ri, range(B),
                     lambda w: R.array_interpolate_linear(w, axis=0)[:, :, 0], lambda y, x0, w: host(o[:, :, y, x0:x0 + w]), True))
            del ci
            torch.cuda.empty_cache()
        except Exception as e:  # pragma: no cover
            extra["long_columns_interp"] = {"error": repr(e)[:300]}
        # aggregate_spatial(mean) of one polygon spanning the strips of ALL ranks -- per-rank
        # rasterise + zonal partials, then the only collective on a measured path (all_reduce of (sum, n) per plane)
        try:
            from openeo_processes_dask_b200 import sharding

            xs_pix = np.arange(X, dtype=np.float64) + 0.5
            ys_pix = float(y_full) - (np.arange(row0, row0 + rows, dtype=np.float64) + 0.5)
            cx, cy, rx, ry = X / 2.0, y_full / 2.0, 0.45 * X, 0.45 * y_full
            ang = np.linspace(0.0, 2.0 * np.pi, 48, endpoint=False)
            rings = [np.stack([cx + rx * np.cos(ang), cy + ry * np.sin(ang)], axis=1)]
            inside_rows = int(np.sum(np.abs(ys_pix - cy) < ry))
            side("next_aggregate_spatial_mean", lambda: sharding.zonal_reduce("mean", strip, ys_pix, xs_pix, rings),
                 T * B * inside_rows * int(0.9 * X * 0.785) * 4, T * B * inside_rows * int(0.9 * X) * 4,
                 "48-gon over 90 % of the cube's extent; bytes ~ samples under the polygon; "
                 + ("NCCL all_reduce of the per-plane partials" if world > 1 else "single rank: no exchange"))
            zm = sharding.zonal_reduce("mean", strip, ys_pix, xs_pix, rings)
            zc = sharding.zonal_reduce("count", strip, ys_pix, xs_pix, rings)
            extra["next_aggregate_spatial_mean"]["planes"] = int(zm.numel())
            extra["next_aggregate_spatial_mean"]["pixels_under_polygon_all_ranks"] = int(zc.max().item())
            # value check: DN are uniform on 0..9999, so every plane's mean is 4999.5 +- a few
            ok = bool(((zm - 4999.5).abs() < 5.0).all().item())
            extra["next_aggregate_spatial_mean"]["parity_checked"] = {"ok": ok, "bar": "|mean - 4999.5| < 5 on every plane "
                                                                      "(uniform DN; exact values are tested in tests/)"}
            if not ok:
                parity.failures.append("aggregate_spatial mean off")
        except Exception as e:  # pragma: no cover
            extra["next_aggregate_spatial_mean"] = {"error": repr(e)[:200]}

    # ---- end to end through the process API with pinned HOST buffers (strip pipeline inside the call) ------------
    e2e = None
    if not args.no_e2e:
        er = min(args.e2e_rows, rows)
        from openeo_processes_dask_b200.sharding import bind_rank_to_gpu_numa

        numa = bind_rank_to_gpu_numa(local) if world > 1 else {"bound": False}
        hbuf = torch.empty((T, B, er, X), dtype=torch.float32, pin_memory=True)
        hbuf.copy_(strip[:, :, :er, :])
        torch.cuda.synchronize()
        harr = hbuf.numpy()
        for _ in range(2):
            P.median(harr, axis=0)
        barrier()
        k = max(3, args.steps // 2)
        t0 = time.perf_counter()
        for _ in range(k):
            out = P.median(harr, axis=0)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": world * harr.nbytes * k / dt / 1e9, "unit": "GB/s",
               "h2d_bytes_per_step": int(harr.nbytes), "d2h_bytes_per_step": int(out.nbytes),
               "sample": f"median over t of a pinned host strip ({T},{B},{er},{X}) fp32 per rank, numpy in -> numpy out",
               "numa": numa}
        want = R.median(harr[:, :, :2, :4096], axis=0)
        e2e["parity_checked"] = {"ok": bool(np.array_equal(out[:, :2, :4096], want, equal_nan=True)), "bar": "bit-exact",
                                 "pixels": int(want.size)}
        if not e2e["parity_checked"]["ok"]:
            parity.failures.append("e2e median differs from the oracle")
        # the ceiling of ANY host-fed path on this box: raw pinned H2D of the same buffer, all ranks at once
        dbuf = torch.empty((T, B, er, X), dtype=torch.float32, device=dev)
        dbuf.copy_(hbuf, non_blocking=True)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            dbuf.copy_(hbuf, non_blocking=True)
        barrier()
        ceil_dt = max_over_ranks(time.perf_counter() - t0)
        e2e["h2d_ceiling_gbs"] = world * harr.nbytes * k / ceil_dt / 1e9
        e2e["frac_of_h2d_ceiling"] = e2e["value"] / e2e["h2d_ceiling_gbs"]
        e2e["h2d_ceiling_how"] = (f"{k} x cudaMemcpyAsync of the same pinned strip on every rank concurrently, wall clock, "
                                  "max over ranks")
        del dbuf
        if world == 1:  # informational: the same call on a PAGEABLE copy of the cube (an ordinary numpy chunk)
            parr = np.array(harr)
            P.median(parr, axis=0)
            t0 = time.perf_counter()
            for _ in range(3):
                P.median(parr, axis=0)
            e2e["pageable_value"] = parr.nbytes * 3 / (time.perf_counter() - t0) / 1e9
            del parr
        del hbuf

    del strip, med
    torch.cuda.empty_cache()

    # ---- every BASELINE config as configured: full cube cut by strip_bounds over the ranks (strong scaling) --------
    configs = {}

    def plan(y_rows, bytes_per_row):
        """(row0, rows, passes): this rank's share of the cube, cut into equal passes until one pass fits in HBM."""
        free_b, _ = torch.cuda.mem_get_info()
        y0, y1 = strip_bounds(y_rows, world, rank)
        passes = 1
        while (-(-(y1 - y0) // passes)) * bytes_per_row > free_b * 0.90:
            passes *= 2
        passes = int(max_over_ranks(passes))
        return y0, -(-(y1 - y0) // passes), passes

    def record(name, shape, fn, in_b, alg_b, passes, note, checks, steps=None):
        p2, s2, _, nl, out = timed(fn, steps or max(3, args.steps // 2), 2)
        ms = max_over_ranks(s2 / len(p2))
        entry = {"shape": list(shape), "rows_this_rank_per_pass": int(shape[2]), "passes": passes,
                 "ms_per_pass": ms, "full_cube_ms": ms * passes,
                 "input_gbs": sum_over_ranks(in_b) / (ms * 1e-3) / 1e9,
                 "hbm_frac": alg_b / (float(p2.mean()) * 1e-3) / 1e9 / peak, "launches_per_step": nl // len(p2),
                 "note": note, "parity_checked": {k_: c(out) for k_, c in checks.items()}}
        configs[name] = entry
        return out

    if not args.no_configs:
        try:
            # configs[0]: reduce_dimension(mean, 't') on (32, 4, 1024, 1024) fp32 -- the reference's CPU-runnable case
            t1, b1, y1_, x1 = 32, 4, 1024, 1024
            ya, yb = strip_bounds(y1_, world, rank)
            c1 = torch.empty((t1, b1, yb - ya, x1), dtype=torch.float32, device=dev)
            engine.synth_fill(c1.view(t1 * b1, (yb - ya) * x1), t1 * b1, y1_ * x1, ya * x1, (yb - ya) * x1, 42, 0.0)
            cube1 = P.DataCube(c1, ("t", "bands", "y", "x"))
            red = functools.partial(P.process(P.mean), data=ref)
            full1 = (t1, b1, y1_, x1)
            record("C1_mean_t32_b4_1024x1024", c1.shape,
                   lambda: P.reduce_dimension(data=cube1, reducer=red, dimension="t").data,
                   c1.numel() * 4, b1 * (yb - ya) * x1 * (4 * t1 + 4), 1,
                   "whole cube resident; a launch is ~0.1 ms: launch-latency bound",
                   {"mean": lambda o: parity.check("C1 mean", 42, full1, np.float32, 0.0, ya, yb - ya, range(b1),
                                                   lambda w: R.mean(w, axis=0)[:, 0],
                                                   lambda y, x0, w: host(o[:, y, x0:x0 + w]), True)},
                   steps=max(10, args.steps))
            if not args.no_e2e:
                h1 = np.ascontiguousarray(host(c1))
                P.mean(h1, axis=0)
                barrier()
                t0 = time.perf_counter()
                for _ in range(5):
                    P.mean(h1, axis=0)
                dt1 = max_over_ranks(time.perf_counter() - t0)
                configs["C1_mean_t32_b4_1024x1024"]["e2e_pageable_host_gbs"] = sum_over_ranks(h1.nbytes) * 5 / dt1 / 1e9
            del c1, cube1
        except Exception as e:  # pragma: no cover
            configs["C1_mean_t32_b4_1024x1024"] = {"error": repr(e)[:300]}
        try:
            # configs[1] as a whole cube: median, (100, 4, 10980, 10980), split over the ranks
            y0, r2, passes = plan(Y_FULL, T * B * X * 4 + B * X * 4)
            c2 = torch.empty((T, B, r2, X), dtype=torch.float32, device=dev)
            engine.synth_fill(c2.view(T * B, r2 * X), T * B, Y_FULL * X, y0 * X, r2 * X, SEED, NAN_PROB)
            cube2 = P.DataCube(c2, ("t", "bands", "y", "x"))
            red = functools.partial(P.process(P.median), data=ref)
            record("C2_median_t100_b4_10980x10980", c2.shape,
                   lambda: P.reduce_dimension(data=cube2, reducer=red, dimension="t").data,
                   c2.numel() * 4, B * r2 * X * (4 * T + 4), passes, "through reduce_dimension on a DataCube",
                   {"median": lambda o: parity.check("C2 median", SEED, (T, B, Y_FULL, X), np.float32, NAN_PROB, y0, r2,
                                                     range(B), lambda w: R.median(w, axis=0)[:, 0],
                                                     lambda y, x0, w: host(o[:, y, x0:x0 + w]), True)})
            del c2, cube2
            torch.cuda.empty_cache()
        except Exception as e:  # pragma: no cover
            configs["C2_median_t100_b4_10980x10980"] = {"error": repr(e)[:300]}
        try:
            # configs[2]: quantiles([0.1, 0.5, 0.9]) + sd over t on (100, 12, 10980, 10980) fp32, y-strip sharded
            from openeo_processes_dask_b200.process_implementations.math import reduce_many

            b3 = 12
            y0, r3, passes = plan(Y_FULL, T * b3 * X * 4 + b3 * X * (3 * 8 + 4))
            c3 = torch.empty((T, b3, r3, X), dtype=torch.float32, device=dev)
            engine.synth_fill(c3.view(T * b3, r3 * X), T * b3, Y_FULL * X, y0 * X, r3 * X, SEED + 1, NAN_PROB)
            full3 = (T, b3, Y_FULL, X)
            bsel = [0, 5, 11]
            chk = {
                "quantiles": lambda o: parity.check("C3 quantiles", SEED + 1, full3, np.float32, NAN_PROB, y0, r3, bsel,
                                                    lambda w: R.quantiles(w, probabilities=Q3, axis=0)[:, :, 0],
                                                    lambda y, x0, w: host(o[0][:, bsel, y, x0:x0 + w]), False),
                "sd": lambda o: parity.check("C3 sd", SEED + 1, full3, np.float32, NAN_PROB, y0, r3, bsel,
                                             lambda w: R.sd(w, axis=0)[:, 0],
                                             lambda y, x0, w: host(o[1][bsel, y, x0:x0 + w]), True),
            }
            record("C3_quantiles+sd_t100_b12_10980x10980", c3.shape,
                   lambda: reduce_many(c3, [("quantiles", {"probabilities": Q3}), "sd"], axis=0),
                   c3.numel() * 4, b3 * r3 * X * (4 * T + 28), passes,
                   "one pass (eoc_select_fused): quantiles float64 + sd float32 out of the same resident column", chk)
            record("C3_two_calls_quantiles_then_sd", c3.shape,
                   lambda: (P.quantiles(c3, probabilities=Q3, axis=0), P.sd(c3, axis=0)),
                   c3.numel() * 4, b3 * r3 * X * (8 * T + 28), passes,
                   "the same two processes called one after the other (two reads of the cube)", chk)
            del c3
            torch.cuda.empty_cache()
        except Exception as e:  # pragma: no cover
            configs["C3_quantiles+sd_t100_b12_10980x10980"] = {"error": repr(e)[:300]}
        try:
            # configs[3]: ndvi(nir, red) -> reduce_dimension(max, 't') on (200, 12, 10980, 10980) uint16 -> fp32
            t4, b4 = 200, 12
            y0, r4, passes = plan(Y_FULL, t4 * b4 * X * 2 + X * 4)
            u16 = torch.empty((t4, b4, r4, X), dtype=torch.uint16, device=dev)
            engine.synth_fill(u16.view(t4 * b4, r4 * X), t4 * b4, Y_FULL * X, y0 * X, r4 * X, SEED + 2, 0.0)
            names = [f"B{k:02d}" for k in range(1, 13)]
            cube4 = P.DataCube(u16, ("t", "bands", "y", "x"), {"bands": names}).astype("float32")
            red_max = functools.partial(P.process(P._max), data=ref)

            def ndvi_max():
                return P.reduce_dimension(data=P.ndvi(cube4, nir="B08", red="B04"), reducer=red_max, dimension="t").data

            def oracle4(w):
                f = w.astype(np.float32)
                return R._max(R.normalized_difference(f[:, 0], f[:, 1]), axis=0)[0]

            record("C4_ndvi_max_t200_b12_uint16_10980x10980", u16.shape, ndvi_max, 2 * t4 * r4 * X * 2,
                   r4 * X * (2 * 2 * t4 + 4), passes,
                   "cube.astype(float32) -> ndvi -> reduce_dimension(max): two of twelve uint16 band planes read, "
                   "converted in the load stage, NDVI never materialised (eoc_nd_reduce)",
                   {"ndvi_max": lambda o: parity.check("C4 ndvi max", SEED + 2, (t4, b4, Y_FULL, X), np.uint16, 0.0, y0, r4,
                                                       [7, 3], oracle4, lambda y, x0, w: host(o[y, x0:x0 + w]), True)})
            del u16, cube4
            torch.cuda.empty_cache()
        except Exception as e:  # pragma: no cover
            configs["C4_ndvi_max_t200_b12_uint16_10980x10980"] = {"error": repr(e)[:300]}
        try:
            # configs[4]: aggregate_temporal, 12 monthly intervals, daily t=365, bands=4, 5490x5490, 40 % NaN
            import pandas as pd

            t5, b5, x5, y5 = 365, 4, 5490, 5490
            y0, r5, passes = plan(y5, t5 * b5 * x5 * 4 + 12 * b5 * x5 * 12)
            c5 = torch.empty((t5, b5, r5, x5), dtype=torch.float32, device=dev)
            engine.synth_fill(c5.view(t5 * b5, r5 * x5), t5 * b5, y5 * x5, y0 * x5, r5 * x5, SEED + 3, 0.40)
            days = pd.date_range("2021-01-01T12:00:00", periods=t5, freq="D").values
            months = pd.date_range("2021-01-01", "2022-01-01", freq="MS").strftime("%Y-%m-%dT%H:%M:%S")
            intervals = [[a, b] for a, b in zip(months[:-1], months[1:])]
            cube5 = P.DataCube(c5, ("t", "bands", "y", "x"), {"t": days})
            full5 = (t5, b5, y5, x5)
            for nm, fn_, ofn, ob in (("mean", P.mean, R.mean, 4), ("count", P.count, R.count, 8)):
                red = functools.partial(P.process(fn_), data=ref)

                def oracle5(w, ofn=ofn):
                    out, _, _ = OC.aggregate_temporal(w, ("t", "bands", "y", "x"), {"t": days}, intervals, ofn)
                    return out[:, :, 0]

                record(f"C5_aggregate_temporal_{nm}_t365_b4_5490x5490", c5.shape,
                       lambda red=red: P.aggregate_temporal(data=cube5, intervals=intervals, reducer=red).data,
                       c5.numel() * 4, b5 * r5 * x5 * (4 * t5 + 12 * ob), passes,
                       "12 monthly bins, one segmented launch (eoc_segmented_reduce)",
                       {nm: lambda o, oracle5=oracle5, nm=nm: parity.check(
                           f"C5 {nm}", SEED + 3, full5, np.float32, 0.40, y0, r5, range(b5), oracle5,
                           lambda y, x0, w: host(o[:, :, y, x0:x0 + w]), True)})
            if not args.no_e2e:
                # end to end for C5: aggregate_temporal(mean) of a PINNED numpy chunk through the host strip pipeline
                e5 = min(64, r5)
                h5 = torch.empty((t5, b5, e5, x5), dtype=torch.float32, pin_memory=True)
                h5.copy_(c5[:, :, :e5, :])
                torch.cuda.synchronize()
                hcube = P.DataCube(h5.numpy(), ("t", "bands", "y", "x"), {"t": days})
                red = functools.partial(P.process(P.mean), data=ref)
                for _ in range(2):
                    P.aggregate_temporal(data=hcube, intervals=intervals, reducer=red)
                barrier()
                t0 = time.perf_counter()
                for _ in range(3):
                    oh = P.aggregate_temporal(data=hcube, intervals=intervals, reducer=red)
                barrier()
                dt5 = max_over_ranks(time.perf_counter() - t0)
                want5, _, _ = OC.aggregate_temporal(h5.numpy()[:, :, :1, :2048], ("t", "bands", "y", "x"), {"t": days},
                                                    intervals, R.mean)
                ok5 = bool(np.array_equal(np.asarray(oh.data)[:, :, :1, :2048], want5, equal_nan=True))
                if not ok5:
                    parity.failures.append("C5 e2e aggregate_temporal differs from the oracle")
                configs["C5_aggregate_temporal_mean_t365_b4_5490x5490"]["e2e_pinned_host"] = {
                    "value": sum_over_ranks(h5.numel() * 4) * 3 / dt5 / 1e9, "unit": "GB/s",
                    "h2d_bytes_per_step": int(h5.numel() * 4), "d2h_bytes_per_step": int(np.asarray(oh.data).nbytes),
                    "parity_checked": {"ok": ok5, "bar": "bit-exact"},
                    "sample": f"aggregate_temporal(mean) of a pinned numpy chunk ({t5},{b5},{e5},{x5}) per rank"}
                del h5, hcube
            del c5, cube5
            torch.cuda.empty_cache()
        except Exception as e:  # pragma: no cover
            configs["C5_aggregate_temporal"] = {"error": repr(e)[:300]}

    # ---- the reference's CPU path on this box's host cores (rank 0; the other ranks wait at the barrier) ----------
    cpu, cpu_other = None, {}
    if rank == 0 and not args.no_cpu:
        real, why = probe_real_reference()
        v, cores, sample = cpu_reference(args.cpu_seconds)
        cpu = {"value": v, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample, "real_reference": real,
               "real_reference_probe": why}
        for red_ in ("mean", "sd", "max", "quantiles"):
            v2, _, s2 = cpu_reference(max(2.0, args.cpu_seconds / 4), red_)
            cpu_other[red_] = {"value": v2, "unit": "GB/s", "cores": cores, "kind": "port", "sample": s2}
        v1, _, s1 = cpu_reference(2.0, "mean", t=32, nan_prob=0.0, seed=42, full_shape=(32, 4, 1024, 1024))
        cpu_other["C1_mean"] = {"value": v1, "unit": "GB/s", "cores": cores, "kind": "port", "sample": s1}
    barrier()
    for key, name in (("mean", "mean"), ("sd", "sd"), ("max", "max"), ("quantiles", "quantiles_0.1_0.5_0.9")):
        if key in cpu_other and name in extra:
            extra[name]["cpu_baseline"] = cpu_other[key]
    if "C1_mean" in cpu_other and "C1_mean_t32_b4_1024x1024" in configs:
        configs["C1_mean_t32_b4_1024x1024"]["cpu_baseline"] = cpu_other["C1_mean"]

    n_fail = int(sum_over_ranks(len(parity.failures)))
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
            "parity_checked": head_parity, "parity_failures": n_fail, "parity_failure_notes": parity.failures[:8],
            "reducers": extra, "configs": configs, "wall_s_timed_region": wall,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if n_fail:
        sys.stdout.flush()
        raise SystemExit(f"bench.py: {n_fail} parity check(s) failed: {parity.failures[:3]}")


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
