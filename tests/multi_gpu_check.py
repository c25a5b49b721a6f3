#!/usr/bin/env python3
"""Two (or more) GPUs, one process each, NCCL: run with

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/multi_gpu_check.py

Checks (a) a t-reduction of y-strips needs no exchange: the concatenated per-rank medians / sd equal the
oracle on the whole cube; (b) a spatial reduction combines per-rank kernel partials through NCCL and equals
numpy on the whole cube.  Not collected by pytest (needs a multi-GPU launch); the same host logic runs on
the CPU over gloo in tests/test_sharding_gloo.py.
"""
import os
import sys
import warnings

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from openeo_processes_dask_b200 import sharding  # noqa: E402
import openeo_processes_dask_b200.process_implementations as P  # noqa: E402
from oracle import synth  # noqa: E402

warnings.simplefilter("ignore", RuntimeWarning)


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    world, rank = dist.get_world_size(), dist.get_rank()
    T, B, Y, X = 40, 3, 61, 256
    full = synth.synth_cube(7, (T, B, Y, X), np.float32, 0.3)
    y0, y1 = sharding.my_strip(Y)
    strip = torch.from_numpy(np.ascontiguousarray(full[:, :, y0:y1])).to(dev)
    # (a) t reductions: strip in, strip out, no collective
    med = sharding.reduce_strip(P.median, strip, 0)
    sd = sharding.reduce_strip(P.sd, strip, 0)
    np.testing.assert_array_equal(med.cpu().numpy(), np.nanmedian(full, axis=0)[:, y0:y1])
    np.testing.assert_array_equal(sd.cpu().numpy(), np.nanstd(full, axis=0, ddof=1)[:, y0:y1])
    # (b) spatial reductions: kernel partials + NCCL combine; every rank holds the full answer
    flat = full.reshape(T, B, -1).astype(np.float64)
    want = {"sum": np.nansum(flat, axis=2), "mean": np.nanmean(flat, axis=2), "min": np.nanmin(flat, axis=2),
            "max": np.nanmax(flat, axis=2), "sd": np.nanstd(flat, axis=2, ddof=1),
            "count": np.sum(np.isfinite(flat), axis=2)}
    for op, w in want.items():
        got = sharding.spatial_reduce(op, strip, (2, 3)).cpu().numpy()
        np.testing.assert_allclose(got, w, rtol=1e-6, equal_nan=True, err_msg=op)
    small = strip[:4, :, :, :8].contiguous()
    got = sharding.spatial_reduce("median", small, (2, 3), shard_axis=2).cpu().numpy()
    np.testing.assert_array_equal(got, np.nanmedian(full[:4, :, :, :8].reshape(4, B, -1), axis=2))
    # (c) aggregate_spatial: one polygon (with a hole) across the strips -- per-rank rasterise + zonal partials,
    # then the same NCCL combine
    from oracle import cubes as OC

    xs, ys = np.arange(X) * 10.0 + 5.0, 1000.0 - np.arange(Y) * 10.0
    rings = [np.array([[103.0, 950.0], [2203.0, 870.0], [1800.0, 455.0], [400.0, 430.0]]),
             np.array([[900.0, 800.0], [1300.0, 800.0], [1300.0, 600.0], [900.0, 600.0]])]
    inside = OC.point_in_polygon(xs, ys, rings)
    masked = np.where(inside[None, None], full.astype(np.float64), np.nan).reshape(T, B, -1)
    wantz = {"mean": np.nanmean(masked, axis=2), "sum": np.nansum(masked, axis=2), "max": np.nanmax(masked, axis=2),
             "sd": np.nanstd(masked, axis=2, ddof=1), "count": np.sum(np.isfinite(masked), axis=2)}
    for op, w in wantz.items():
        got = sharding.zonal_reduce(op, strip, ys[y0:y1], xs, rings).cpu().numpy()
        np.testing.assert_allclose(got, w, rtol=1e-6, equal_nan=True, err_msg="zonal " + op)
    dist.barrier(device_ids=[local])
    if rank == 0:
        print(f"multi_gpu_check ok: world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
