#!/usr/bin/env python3
"""Run ONE reducer a few times on a resident C2-shaped strip: the short command ncu wraps.

    python tools/profile_one.py median --rows 256 --iters 3
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("reducer", choices=["median", "mean", "sd", "max", "quantiles", "quant_sd", "median_mean", "ndvi_max", "first", "last", "any", "all",
                                        "count", "masked_median", "masked_median_2pass", "masked_mean", "masked_quantiles", "cumsum", "cummax", "sort", "interp", "lsr", "add", "add_plane", "gt", "sort_only"])
    ap.add_argument("--rows", type=int, default=256)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--t", type=int, default=100)
    args = ap.parse_args()

    import torch

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.process_implementations.math import reduce_many

    T, B, X, Y = args.t, 4, 10980, 10980
    strip = torch.empty((T, B, args.rows, X), dtype=torch.float32, device="cuda")
    engine.synth_fill(strip.view(T * B, args.rows * X), T * B, Y * X, 0, args.rows * X, 43, 0.30)
    fns = {
        "median": lambda: P.median(strip, axis=0),
        "mean": lambda: P.mean(strip, axis=0),
        "sd": lambda: P.sd(strip, axis=0),
        "max": lambda: P._max(strip, axis=0),
        "quantiles": lambda: P.quantiles(strip, probabilities=[0.1, 0.5, 0.9], axis=0),
        "quant_sd": lambda: reduce_many(strip, [("quantiles", {"probabilities": [0.1, 0.5, 0.9]}), "sd"], axis=0),
        "median_mean": lambda: reduce_many(strip, ["median", "mean"], axis=0),
        "first": lambda: P.first(strip, axis=0),
        "last": lambda: P.last(strip, axis=0),
        "any": lambda: P._any(strip, axis=0),
        "all": lambda: P._all(strip, axis=0),
        "count": lambda: P.count(strip, axis=0),
        "cumsum": lambda: P.cumsum(strip, axis=0),
        "cummax": lambda: P.cummax(strip, axis=0),
        "sort": lambda: P.sort(strip, nodata=True, axis=0),
        "interp": lambda: P.array_interpolate_linear(strip, axis=0),
        "lsr": lambda: P.linear_scale_range(strip, 0, 8000, 0, 255),
        "add": lambda: P.add(strip[: T // 2], strip[T // 2:]),            # x + y, equal shapes (12 B per sample)
        "add_plane": lambda: P.add(strip, strip[0, 0]),                    # cube + (y,x) plane broadcast
        "gt": lambda: P.gt(strip[: T // 2], strip[T // 2:]),              # float64 result (16 B per sample)
    }
    if args.reducer.startswith("masked_"):  # (t,y,x) cloud mask -> NaN -> reduce over t
        g = torch.Generator(device="cuda").manual_seed(5)
        m = torch.rand((T, args.rows, X), device="cuda", generator=g) < 0.25
        cube = P.DataCube(strip, ("t", "bands", "y", "x"), {"t": list(range(T))})
        mcube = P.DataCube(m, ("t", "y", "x"), {"t": list(range(T))})
        fns["masked_median"] = lambda: P.median(P.mask(data=cube, mask=mcube, fuse=True).data, axis=0)
        fns["masked_median_2pass"] = lambda: P.median(P.mask(data=cube, mask=mcube).data, axis=0)
        fns["masked_mean"] = lambda: P.mean(P.mask(data=cube, mask=mcube, fuse=True).data, axis=0)
        fns["masked_quantiles"] = lambda: P.quantiles(P.mask(data=cube, mask=mcube, fuse=True).data,
                                                      probabilities=[0.1, 0.5, 0.9], axis=0)
    if args.reducer == "ndvi_max":  # BASELINE configs[3]: uint16 cube, t=200, 12 bands, two of them read
        import functools

        del strip
        t4, b4 = 200, 12
        u16 = torch.empty((t4, b4, args.rows, X), dtype=torch.uint16, device="cuda")
        engine.synth_fill(u16.view(t4 * b4, args.rows * X), t4 * b4, Y * X, 0, args.rows * X, 45, 0.0)
        cube = P.DataCube(u16, ("t", "bands", "y", "x"), {"bands": [f"B{k:02d}" for k in range(1, 13)]}).astype("float32")
        red = functools.partial(P.process(P._max), data=P.ParameterReference(from_parameter="data"))
        fns["ndvi_max"] = lambda: P.reduce_dimension(data=P.ndvi(cube, nir="B08", red="B04"), reducer=red,
                                                     dimension="t")
        strip = u16[:, :2]  # bytes actually read: two band planes
    fn = fns[args.reducer]
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ms = [a.elapsed_time(b) for a, b in ev]
    gb = strip.numel() * strip.element_size() / 1e9
    print(f"{args.reducer}: rows={args.rows} input={gb:.2f} GB ms={['%.3f' % m for m in ms]} "
          f"best={gb / (min(ms) * 1e-3):.0f} GB/s")


if __name__ == "__main__":
    main()
