#!/usr/bin/env python3
"""Throughput of mean / sd / median over differently laid out cubes (dispatch sanity: no layout should fall off a cliff)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import openeo_processes_dask_b200.process_implementations as P

CASES = [
    ("(t,b,y,x) over t", (100, 4, 256, 4096), 0),
    ("(b,t,y,x) over t", (4, 100, 256, 4096), 1),
    ("(t,b,y,x) over bands", (100, 12, 256, 1024), 1),
    ("(y,x,t) over t", (1024, 2048, 100), 2),
    ("(x,y,t,b) over t", (1024, 1024, 100, 4), 2),
    ("(x,y,t,b) over bands", (1024, 1024, 100, 4), 3),
    ("(t,y,x) over y", (100, 1024, 2048), 1),
    ("(t,b,y,x) over x", (100, 4, 512, 1024), 3),
    ("(t,b,y,x) odd x over t", (100, 4, 255, 4097), 0),
    ("(t,b,y,x) x=130 over t", (100, 4, 4096, 130), 0),
    ("(b,t,y,x) small yx over t", (4000, 100, 8, 12), 1),
]
for name, shape, axis in CASES:
    x = torch.randn(shape, device="cuda")
    x[torch.rand(shape, device="cuda") < 0.3] = float("nan")
    row = []
    for fn in (P.mean, P.sd, P.median):
        for _ in range(2):
            fn(x, axis=axis)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn(x, axis=axis)
        b.record()
        torch.cuda.synchronize()
        row.append(x.numel() * 4 / a.elapsed_time(b) / 1e6)
    print(f"{name:32s} {str(shape):24s} mean {row[0]:6.0f}  sd {row[1]:6.0f}  median {row[2]:6.0f} GB/s")
