#!/usr/bin/env python3
"""A/B timing of the selection kernels on a resident C2-shaped strip (one JSON line; env knobs pick the variant).

    EOC_SELECT_SPLIT=0 python tools/ab_select.py --rows 1024         # round-1 kernels
    EOC_SELECT_SPLIT=1 EOC_SPLIT_W=256 python tools/ab_select.py     # split kernel everywhere, 256-column tiles

Input GB/s = strip bytes / mean device time of `--iters` launches (CUDA events around each call, after warm-up).
"""

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--t", type=int, default=100)
    ap.add_argument("--only", default="")
    ap.add_argument("--ballast-gb", type=float, default=0.0, help="resident dummy allocation (the bench holds a 96 GB strip)")
    ap.add_argument("--slice-of", type=int, default=0, help="make the strip a .contiguous() row slice of a taller one")
    args = ap.parse_args()

    import torch

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.process_implementations.math import reduce_many

    T, B, X, Y = args.t, 4, 10980, 10980
    ballast = torch.empty(int(args.ballast_gb * 2**30), dtype=torch.uint8, device="cuda") if args.ballast_gb else None
    if args.slice_of:
        tall = torch.empty((T, B, args.slice_of, X), dtype=torch.float32, device="cuda")
        engine.synth_fill(tall.view(T * B, args.slice_of * X), T * B, Y * X, 0, args.slice_of * X, 43, 0.30)
        strip = tall[:, :, :args.rows, :].contiguous()
    else:
        strip = torch.empty((T, B, args.rows, X), dtype=torch.float32, device="cuda")
        engine.synth_fill(strip.view(T * B, args.rows * X), T * B, Y * X, 0, args.rows * X, 43, 0.30)
    g = torch.Generator(device="cuda").manual_seed(5)
    m = torch.rand((T, args.rows, X), device="cuda", generator=g) < 0.25
    cube = P.DataCube(strip, ("t", "bands", "y", "x"), {"t": list(range(T))})
    mcube = P.DataCube(m, ("t", "y", "x"), {"t": list(range(T))})
    q3 = [0.1, 0.5, 0.9]
    fns = {
        "median": lambda: P.median(strip, axis=0),
        "quantiles": lambda: P.quantiles(strip, probabilities=q3, axis=0),
        "quant_sd": lambda: reduce_many(strip, [("quantiles", {"probabilities": q3}), "sd"], axis=0),
        "median_mean": lambda: reduce_many(strip, ["median", "mean"], axis=0),
        "masked_median": lambda: P.median(P.mask(data=cube, mask=mcube).data, axis=0),
        "quantiles8": lambda: P.quantiles(strip, probabilities=[0.05, 0.1, 0.25, 0.4, 0.5, 0.6, 0.75, 0.9], axis=0),
    }
    only = [s for s in args.only.split(",") if s]
    out = {"env": {k: v for k, v in os.environ.items() if k.startswith("EOC_")}, "rows": args.rows, "t": T}
    for name, fn in fns.items():
        if only and name not in only:
            continue
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.iters)]
        for a, b in ev:
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
        out[name] = round(strip.numel() * 4 / (ms * 1e-3) / 1e9, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
