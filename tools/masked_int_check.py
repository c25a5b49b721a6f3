#!/usr/bin/env python3
"""mask -> median of a raw uint16 cube: the fused four-run kernel (EOC_QUAD_INT=1) against materialise-first (=0)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import openeo_processes_dask_b200.process_implementations as P
from openeo_processes_dask_b200 import engine

T, B, rows, X = 100, 4, int(os.environ.get("ROWS", "512")), 10980
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.randint(0, 10000, (T, B, rows, X), device="cuda", generator=g, dtype=torch.int32).to(torch.uint16)
m = torch.rand((T, rows, X), device="cuda", generator=g) < 0.25
cube = P.DataCube(x, ("t", "bands", "y", "x"), {"t": list(range(T))})
mcube = P.DataCube(m, ("t", "y", "x"), {"t": list(range(T))})
res = {}
for flag in ("1", "0"):
    os.environ["EOC_QUAD_INT"] = flag
    fn = lambda: P.median(P.mask(data=cube, mask=mcube).data, axis=0)
    for _ in range(2):
        out = fn()
    torch.cuda.synchronize()
    engine.launch_count(reset=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    res[flag] = out
    print(f"EOC_QUAD_INT={flag}: {ms:.3f} ms, {x.numel() * 2 / ms / 1e6:.0f} GB/s of uint16 input, "
          f"{engine.launch_count() // 3} launches per call, out dtype {out.dtype}")
t1, t0 = res["1"], res["0"]
t1 = t1.tensor if hasattr(t1, "tensor") else t1
t0 = t0.tensor if hasattr(t0, "tensor") else t0
print("equal:", bool(((t1 == t0) | (torch.isnan(t1) & torch.isnan(t0))).all()))
