"""N > 1 host logic on the CPU: two ranks over gloo.

The per-rank partials are produced here with numpy (the kernels need a GPU; their parity is the job of
tests/test_gpu_parity.py) -- what is under test is the product's partitioning and its combine step:
``strip_bounds`` / ``my_strip`` and ``combine_spatial_partials`` with its collectives.
"""

import socket
import warnings

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from openeo_processes_dask_b200 import sharding  # noqa: E402
from openeo_processes_dask_b200.views import strip_bounds  # noqa: E402

warnings.simplefilter("ignore", RuntimeWarning)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _cube(seed=5):
    rng = np.random.default_rng(seed)
    x = rng.normal(3000, 1000, size=(6, 3, 37, 20))  # (t, bands, y, x)
    x[rng.random(x.shape) < 0.3] = np.nan
    x[0, 0] = np.nan            # one (t, band) plane without any valid sample
    x[1, 1, :19] = np.nan       # ... and one whose valid samples all sit in the second strip
    x[2, 2, 5, 5] = np.inf
    return x


def _numpy_partials(strip):
    flat = strip.reshape(strip.shape[0], strip.shape[1], -1)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        n = np.sum(~np.isnan(flat), axis=2)
        mean = np.nanmean(flat, axis=2)
        part = {
            "sum": np.nansum(flat, axis=2), "n": n, "count": np.sum(np.isfinite(flat), axis=2),
            "min": np.nanmin(flat, axis=2), "max": np.nanmax(flat, axis=2), "mean": mean,
            "m2": np.nanvar(flat, axis=2, ddof=0) * n, "len": np.full(n.shape, flat.shape[2]),
        }
    return {k: torch.from_numpy(np.asarray(v, dtype=np.int64 if k in ("n", "count", "len") else np.float64))
            for k, v in part.items()}


def _worker(rank, world, port, ignore):
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        x = _cube()
        y0, y1 = sharding.my_strip(x.shape[2])
        assert (y0, y1) == strip_bounds(x.shape[2], world, rank)
        part = _numpy_partials(x[:, :, y0:y1])
        flat = x.reshape(x.shape[0], x.shape[1], -1)
        want = {
            "sum": np.nansum(flat, axis=2), "mean": np.nanmean(flat, axis=2), "min": np.nanmin(flat, axis=2),
            "max": np.nanmax(flat, axis=2), "variance": np.nanvar(flat, axis=2, ddof=1),
            "sd": np.nanstd(flat, axis=2, ddof=1), "count": np.sum(np.isfinite(flat), axis=2),
        }
        if not ignore:
            plain = {"sum": np.sum, "mean": np.mean, "min": np.min, "max": np.max}
            want.update({k: f(flat, axis=2) for k, f in plain.items()})
            want["variance"] = np.var(flat, axis=2, ddof=1)
            want["sd"] = np.std(flat, axis=2, ddof=1)
        for op in sharding.MERGEABLE:
            got = sharding.combine_spatial_partials(op, part, ignore_nodata=ignore, ddof=1).numpy()
            np.testing.assert_array_equal(np.isnan(got.astype(np.float64)), np.isnan(want[op].astype(np.float64)),
                                          err_msg=op)
            np.testing.assert_allclose(got, want[op], rtol=1e-12, equal_nan=True, err_msg=op)
        with pytest.raises(ValueError):
            sharding.combine_spatial_partials("median", part)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("ignore", [True, False])
def test_spatial_combine_two_ranks_gloo(ignore):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        mp.spawn(_worker, args=(2, _free_port(), ignore), nprocs=2, join=True)


def test_single_rank_needs_no_process_group():
    """world_size 1 (or no torch.distributed at all): the combine step is the identity on the partials."""
    x = _cube()
    part = _numpy_partials(x)
    flat = x.reshape(x.shape[0], x.shape[1], -1)
    got = sharding.combine_spatial_partials("sd", part).numpy()
    np.testing.assert_allclose(got, np.nanstd(flat, axis=2, ddof=1), rtol=1e-12, equal_nan=True)
    got = sharding.combine_spatial_partials("max", part).numpy()
    np.testing.assert_array_equal(got, np.nanmax(flat, axis=2))


def test_strips_tile_the_cube():
    for rows, world in ((10980, 8), (10980, 4), (10980, 2), (5490, 8), (7, 8), (37, 2)):
        bounds = [strip_bounds(rows, world, r) for r in range(world)]
        assert bounds[0][0] == 0 and bounds[-1][1] == rows
        assert all(a[1] == b[0] for a, b in zip(bounds, bounds[1:]))
        sizes = [b - a for a, b in bounds]
        assert max(sizes) - min(sizes) <= 1
    assert [b - a for a, b in (strip_bounds(10980, 8, r) for r in range(8))] == [1373] * 4 + [1372] * 4
