#!/usr/bin/env python3
"""median / mean over t of non-float32 cubes (uint16, int16, uint8, float64): python tools/profile_dtype.py uint16 --t 100"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("dtype")
    ap.add_argument("--t", type=int, default=100)
    ap.add_argument("--rows", type=int, default=128)
    args = ap.parse_args()
    import torch

    import openeo_processes_dask_b200.process_implementations as P

    T, B, X = args.t, 4, 10980
    dt = getattr(torch, args.dtype)
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randint(0, 10000 if args.dtype != "uint8" else 255, (T, B, args.rows, X), device="cuda", generator=g,
                      dtype=torch.int32).to(dt)
    gb = x.numel() * x.element_size() / 1e9
    for name, fn in (("median", lambda: P.median(x, axis=0)), ("mean", lambda: P.mean(x, axis=0)),
                     ("quantiles", lambda: P.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0))):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(3)]
        for a, b in ev:
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        ms = min(a.elapsed_time(b) for a, b in ev)
        print(f"{args.dtype} t={T} {name}: {gb:.2f} GB in {ms:.3f} ms = {gb / ms * 1e3:.0f} GB/s")


if __name__ == "__main__":
    main()
