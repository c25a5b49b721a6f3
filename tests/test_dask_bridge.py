"""The dask bridge: every ``_is_dask`` branch of the host layer, driven through ``tests/daskshim.py``.

Reference behaviour being matched: the reducers never compute -- a dask-backed cube gives a lazy dask array back
(``tests/general_checks.py:19`` asserts ``type(out.data) is type(in.data)``, ``tests/test_logic.py:147,166``
``isinstance(output_cube.data, da.Array)``); median / quantiles need the reduced axis in ONE chunk;
numpy and dask inputs give equal results (``tests/test_aggregate.py:153-181``).

* not gpu: structure only -- result kind, chunk grid, dtype, quantile axis position -- and that NOTHING is evaluated
  before ``compute()``; on a box without CUDA ``compute()`` must fail loudly (no CPU fallback).
* gpu: ``compute()`` equals the oracle on the whole array; each block went through the host strip pipeline.
"""

import functools
import warnings

import numpy as np
import pytest

from tests import daskshim

da = daskshim.install()
SHIM = getattr(da, "__eoc_shim__", False)
warnings.simplefilter("ignore", RuntimeWarning)

import openeo_processes_dask_b200.process_implementations as P  # noqa: E402
from oracle import reducers as R  # noqa: E402


def _cube(seed=5, shape=(40, 3, 12, 20), nan=0.3, dtype=np.float32):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 10000, size=shape).astype(dtype)
    if np.dtype(dtype).kind == "f":
        x[rng.random(shape) < nan] = np.nan
    return x


def _lazy(x, chunks):
    return da.from_array(x, chunks=chunks)


def _is_lazy(a):
    return isinstance(a, da.Array)


CASES = [  # name, call(data), expected shape fn, expected dtype
    ("mean", lambda d: P.mean(d, axis=0), (3, 12, 20), np.float32),
    ("median", lambda d: P.median(d, axis=0), (3, 12, 20), np.float32),
    ("sd", lambda d: P.sd(d, axis=0), (3, 12, 20), np.float32),
    ("max_bands", lambda d: P._max(d, axis=1), (40, 12, 20), np.float32),
    ("sum_keepdims", lambda d: P._sum(d, axis=0, keepdims=True), (1, 3, 12, 20), np.float32),
    ("count", lambda d: P.count(d, axis=0), (3, 12, 20), np.int64),
    ("count_true", lambda d: P.count(d, condition=True, axis=0), (3, 12, 20), np.float32),
    ("quantiles", lambda d: P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=0), (3, 3, 12, 20), np.float64),
    ("quantiles_axis1", lambda d: P.quantiles(d, probabilities=[0.25, 0.75], axis=1), (40, 2, 12, 20), np.float64),
    ("cumsum", lambda d: P.cumsum(d, axis=0), (40, 3, 12, 20), np.float32),
    ("first", lambda d: P.first(d, axis=0), (3, 12, 20), np.float32),
    ("any", lambda d: P._any(d, axis=0), (3, 12, 20), None),
    ("interp", lambda d: P.array_interpolate_linear(d, axis=0), (40, 3, 12, 20), np.float32),
    ("add_scalar", lambda d: P.add(d, 2.5), (40, 3, 12, 20), np.float32),
    ("gt_self", lambda d: P.gt(d, d), (40, 3, 12, 20), np.float64),
    ("clip", lambda d: P.clip(d, 100, 5000), (40, 3, 12, 20), np.float32),
    ("nd", lambda d: P.normalized_difference(d, d), (40, 3, 12, 20), np.float32),
    ("extrema", lambda d: P.extrema(d, axis=0), (2, 3, 12, 20), np.float32),
]


@pytest.mark.parametrize("name,call,shape,dtype", CASES, ids=[c[0] for c in CASES])
def test_dask_in_dask_out_and_nothing_runs(name, call, shape, dtype):
    x = _cube()
    d = _lazy(x, (10, 3, 6, 7))  # the reduced axis (t) arrives in 4 chunks; y, x in ragged chunks (20 = 7+7+6)
    before = da.Array.calls if SHIM else 0
    out = call(d)
    assert _is_lazy(out), f"{name}: {type(out)} is not a dask array"
    assert tuple(out.shape) == shape
    if dtype is not None:
        assert np.dtype(out.dtype) == np.dtype(dtype)
    if SHIM:
        assert da.Array.calls == before, "a block was evaluated before compute()"


def test_reduced_axis_is_rechunked_and_quantile_axis_lands_at_axis():
    d = _lazy(_cube(), (10, 3, 6, 7))
    med = P.median(d, axis=0)
    assert med.chunks == ((3,), (6, 6), (7, 7, 6))  # spatial chunks kept, t gone
    q = P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=0)
    assert q.chunks == ((3,), (3,), (6, 6), (7, 7, 6))
    q1 = P.quantiles(_lazy(_cube(), (40, 1, 6, 20)), probabilities=4, axis=1)  # 3 probabilities at axis 1
    assert q1.chunks == ((40,), (3,), (6, 6), (20,))
    cs = P.cumsum(d, axis=0)
    assert cs.chunks == ((40,), (3,), (6, 6), (7, 7, 6))


def test_reduce_dimension_on_a_dask_cube_stays_lazy():
    """tests/general_checks.py:19 of the reference: the output cube holds the same array type as the input."""
    x = _cube()
    cube = P.DataCube(_lazy(x, (10, 3, 6, 7)), ("t", "bands", "y", "x"), {"bands": ["a", "b", "c"]})
    ref = P.ParameterReference(from_parameter="data")
    for fn in (P.mean, P.median, P._max, P.sd, P.count):
        out = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(fn), data=ref), dimension="t")
        assert type(out.data) is type(cube.data) and out.dims == ("bands", "y", "x")


def _has_cuda():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_cuda(), reason="only meaningful on a box without CUDA")
def test_compute_fails_loudly_without_cuda():
    """No CPU fallback: evaluating a block without a device raises the library's error."""
    from openeo_processes_dask_b200 import _lib

    out = P.mean(_lazy(_cube(shape=(4, 2, 3, 5)), (4, 2, 3, 5)), axis=0)
    with pytest.raises(_lib.EocError):
        out.compute()


# ---- evaluation on the GPU ---------------------------------------------------------------------------------------
ORACLE = {
    "mean": lambda x: R.mean(x, axis=0), "median": lambda x: R.median(x, axis=0), "sd": lambda x: R.sd(x, axis=0),
    "max_bands": lambda x: R._max(x, axis=1), "sum_keepdims": lambda x: R._sum(x, axis=0, keepdims=True),
    "count": lambda x: R.count(x, axis=0), "count_true": lambda x: R.count(x, condition=True, axis=0),
    "quantiles": lambda x: R.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0),
    "quantiles_axis1": lambda x: R.quantiles(x, probabilities=[0.25, 0.75], axis=1),
    "cumsum": lambda x: R.cumsum(x, axis=0), "first": lambda x: R.first(x, axis=0),
    "any": lambda x: R._any(x, axis=0), "interp": lambda x: R.array_interpolate_linear(x, axis=0),
    "add_scalar": lambda x: x + np.float32(2.5), "gt_self": lambda x: R.gt(x, x), "clip": lambda x: np.clip(x, 100, 5000),
    "nd": lambda x: R.normalized_difference(x, x), "extrema": lambda x: np.stack([R._min(x, axis=0), R._max(x, axis=0)]),
}


@pytest.mark.gpu
@pytest.mark.parametrize("name,call,shape,dtype", CASES, ids=[c[0] for c in CASES])
def test_gpu_dask_compute_equals_oracle(name, call, shape, dtype):
    x = _cube()
    del daskshim.BLOCK_LOG[:]
    out = call(_lazy(x, (10, 3, 6, 7)))
    got = out.compute()
    want = np.asarray(ORACLE[name](x))
    assert got.shape == want.shape
    if name.startswith("quantiles"):
        np.testing.assert_allclose(got, want, rtol=1e-6, equal_nan=True)
    else:
        np.testing.assert_array_equal(got, want.astype(got.dtype) if name in ("any",) else want)
    if SHIM and name in ("mean", "median", "quantiles", "count"):
        # 2 x 3 spatial blocks, each with the whole t axis
        assert sorted(daskshim.BLOCK_LOG) == sorted([(40, 3, 6, 7)] * 4 + [(40, 3, 6, 6)] * 2)


@pytest.mark.gpu
def test_gpu_numpy_equals_dask_for_integer_and_float_cubes():
    """tests/test_aggregate.py:153-181 of the reference: numpy == dask for int32 / int64 / float32 / float64."""
    for dt in (np.int32, np.int64, np.float32, np.float64):
        x = _cube(seed=9, shape=(24, 2, 8, 10), dtype=dt)
        d = _lazy(x, (6, 2, 4, 5))
        for fn in (P.mean, P._sum, P._min, P.median):
            a = np.asarray(fn(x, axis=0))
            b = fn(d, axis=0).compute()
            np.testing.assert_array_equal(a, b)
            assert a.dtype == b.dtype, (dt, fn.__name__, a.dtype, b.dtype)


@pytest.mark.gpu
def test_gpu_aggregate_temporal_on_a_dask_cube():
    import pandas as pd

    from oracle import cubes as OC

    x = _cube(seed=3, shape=(60, 2, 6, 8))
    t = pd.date_range("2021-01-01T12:00:00", periods=60, freq="D").values
    iv = [["2021-01-01T00:00:00", "2021-02-01T00:00:00"], ["2021-02-01T00:00:00", "2021-03-01T00:00:00"]]
    cube = P.DataCube(_lazy(x, (15, 2, 3, 8)), ("t", "bands", "y", "x"), {"t": t})
    ref = P.ParameterReference(from_parameter="data")
    out = P.aggregate_temporal(data=cube, intervals=iv, reducer=functools.partial(P.process(P.mean), data=ref))
    want, _, _ = OC.aggregate_temporal(x, cube.dims, {"t": t}, iv, R.mean)
    np.testing.assert_array_equal(np.asarray(out.data), want)


@pytest.mark.gpu
def test_gpu_sort_of_a_dask_array_computes_like_the_reference():
    x = _cube(seed=4, shape=(30, 2, 5, 6))
    got = P.sort(_lazy(np.moveaxis(x, 0, -1).copy(), (2, 5, 6, 10)), nodata=True, axis=3)
    np.testing.assert_array_equal(np.asarray(got), np.sort(np.moveaxis(x, 0, -1), axis=3))
