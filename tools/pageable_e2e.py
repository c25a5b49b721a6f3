#!/usr/bin/env python3
"""End-to-end median over t from a PAGEABLE numpy cube (what a dask chunk is), next to the pinned case."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import openeo_processes_dask_b200.process_implementations as P

T, B, R, X = 100, 4, 128, 10980
x = np.random.default_rng(0).random((T, B, R, X), dtype=np.float32)
x[x < 0.3] = np.nan
pinned = torch.empty(x.shape, dtype=torch.float32, pin_memory=True)
pinned.numpy()[...] = x
for name, arr in (("pageable", x), ("pinned", pinned.numpy())):
    P.median(arr, axis=0)
    t0 = time.perf_counter()
    for _ in range(3):
        P.median(arr, axis=0)
    dt = (time.perf_counter() - t0) / 3
    print(f"{name}: {x.nbytes / dt / 1e9:.1f} GB/s end to end")
