#!/usr/bin/env python3
"""median over t of cubes whose inner extent is small: (x, y, t, bands) -- the reference's test-fixture layout -- and (y, x, t)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import openeo_processes_dask_b200.process_implementations as P

for shape, axis in (((1024, 1024, 100, 4), 2), ((2048, 1024, 100), 2), ((1024, 1024, 100, 12), 2)):
    x = torch.randn(shape, device="cuda")
    x[torch.rand(shape, device="cuda") < 0.3] = float("nan")
    for env in ({}, {"EOC_SELECT_TMA": "0"}):
        os.environ.pop("EOC_SELECT_TMA", None)
        os.environ.update(env)
        for _ in range(2):
            out = P.median(x, axis=axis)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = P.median(x, axis=axis)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        print(f"{shape} axis {axis} {env or 'default'}: {ms:.2f} ms = {x.numel() * 4 / ms / 1e6:.0f} GB/s")
