"""Parity of the CUDA path against the oracle (numpy restatement of the reference), on a B200.

Everything goes through the reference-facing process functions, i.e. through the C ABI of
libeoc_sm100.so: numpy in -> ``eoc_*_host`` (strip pipeline), cuda tensor in -> ``eoc_*`` device
entry points.  Bars: bit-exact for min, max, count, median and NaN masks; for the (t,b,y,x)-style
reductions over a non-contiguous axis also bit-exact for sum/mean/sd/variance/quantiles (the kernels
reproduce numpy's operation order, see tests/test_hostmodel.py); rtol 1e-6 where numpy itself
switches to pairwise summation (reduced axis contiguous) -- the tolerance of BASELINE.json.
"""

import functools
import os
import warnings

import numpy as np
import pandas as pd
import pytest

from oracle import cubes as OC
from oracle import reducers as R
from oracle import synth

pytestmark = pytest.mark.gpu
warnings.simplefilter("ignore", RuntimeWarning)

RTOL = 1e-6  # BASELINE.json north_star: "within 1e-6 relative (fp32) for mean, sd and interpolated quantiles"


@pytest.fixture(scope="module")
def P():
    import openeo_processes_dask_b200.process_implementations as pi

    return pi


@pytest.fixture(scope="module")
def torch():
    import torch as t

    return t


def _mk(rng, shape, dtype, nan_prob):
    x = rng.normal(3000, 1000, size=shape).astype(dtype)
    if np.dtype(dtype).kind == "f":
        x[rng.random(shape) < nan_prob] = np.nan
        flat = x.reshape(-1)
        flat[7 % flat.size] = np.inf
        flat[11 % flat.size] = -np.inf
    return x


STREAM = ["sum", "mean", "min", "max", "sd", "variance", "product"]


def _fn(P, name):
    return {"sum": P._sum, "min": P._min, "max": P._max}.get(name) or getattr(P, name)


def _same(got, want, exact):
    got = np.asarray(got.cpu()) if hasattr(got, "cpu") else np.asarray(got)
    assert got.shape == np.shape(want), (got.shape, np.shape(want))
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
    if exact:
        np.testing.assert_array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=RTOL, equal_nan=True)


@pytest.mark.parametrize("kind", ["numpy", "cuda"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape,axis", [
    ((32, 4, 37, 64), 0),     # t of (t,b,y,x), vector path
    ((20, 3, 5, 7), 0),       # ragged inner: scalar path
    ((4, 32, 16, 16), 1),     # t of (b,t,y,x)
    ((9, 4, 6, 8), 1),        # bands of (t,b,y,x)
    ((30, 30, 20, 4), 2),     # the reference fixture layout (x,y,t,bands), reduce t
    ((12, 10, 6, 4), 3),      # ... reduce bands (contiguous axis)
    ((1, 1, 1, 1), 0),
])
@pytest.mark.parametrize("ignore", [True, False])
def test_streaming_reducers(P, torch, kind, dtype, shape, axis, ignore):
    rng = np.random.default_rng(abs(hash((shape, axis))) % 2**32)
    x = _mk(rng, shape, dtype, 0.3 if ignore else 0.02)
    arg = torch.from_numpy(x).cuda() if kind == "cuda" else x
    contiguous_axis = axis == len(shape) - 1
    for name in STREAM:
        got = _fn(P, name)(arg, ignore_nodata=ignore, axis=axis)
        want = R.BY_NAME[name](x, ignore_nodata=ignore, axis=axis)
        assert (got.cpu().numpy() if kind == "cuda" else got).dtype == want.dtype
        exact = name in ("min", "max") or not contiguous_axis
        _same(got, want, exact)
    got = P.count(arg, axis=axis)
    _same(got, R.count(x, axis=axis), True)
    assert (got.cpu().numpy() if kind == "cuda" else got).dtype == np.int64
    _same(P.count(arg, condition=True, axis=axis), R.count(x, condition=True, axis=axis), True)


@pytest.mark.parametrize("kind", ["numpy", "cuda"])
@pytest.mark.parametrize("dtype,T", [(np.float32, 5), (np.float32, 16), (np.float32, 33), (np.float32, 64),
                                     (np.float32, 100), (np.float32, 128), (np.float32, 129), (np.float32, 200), (np.float32, 256),
                                     (np.float32, 257), (np.float32, 365), (np.float32, 512),
                                     (np.float64, 12), (np.float64, 64), (np.float64, 65), (np.float64, 100),
                                     (np.float64, 128), (np.float64, 200), (np.float64, 256), (np.float64, 365),
                                     (np.float64, 512)])
@pytest.mark.parametrize("ignore", [True, False])
def test_median_and_quantiles(P, torch, kind, dtype, T, ignore):
    rng = np.random.default_rng(T)
    x = _mk(rng, (T, 2, 9, 33), dtype, 0.3 if ignore else 0.01)
    if ignore:
        x[:, 0, 0, 5] = np.nan       # all-NaN column
        x[1:, 0, 0, 1] = np.nan      # one valid sample
    x[:, 1, :, :8] = np.where(np.isnan(x[:, 1, :, :8]), np.nan, np.round(x[:, 1, :, :8] / 400))  # ties
    x[:, 0, 0, 0] = 1.0  # first slice valid: keeps numpy's quantile result buffer float64 (oracle quirk 2)
    arg = torch.from_numpy(x).cuda() if kind == "cuda" else x
    _same(P.median(arg, ignore_nodata=ignore, axis=0), R.median(x, ignore_nodata=ignore, axis=0), True)
    probs = [0.1, 0.5, 0.9]
    got = P.quantiles(arg, probabilities=probs, ignore_nodata=ignore, axis=0)
    want = R.quantiles(x.copy(), probabilities=probs, ignore_nodata=ignore, axis=0)
    _same(got, want, True)
    assert (got.cpu().numpy() if kind == "cuda" else got).dtype == np.float64


def test_quantiles_axis_position_and_probability_forms(P):
    rng = np.random.default_rng(3)
    x = _mk(rng, (6, 5, 20, 4), np.float32, 0.2)
    x[0, 0, :, 0] = 1.0
    for axis in (0, 2, 3):
        got = P.quantiles(x, probabilities=4, axis=axis)
        want = R.quantiles(x, probabilities=4, axis=axis)
        assert got.shape == want.shape and got.shape[axis] == 3
        np.testing.assert_allclose(got, want, rtol=RTOL, equal_nan=True)
    assert P.quantiles(x, probabilities=3, axis=2).shape[2] == 3          # arange quirk: p entries for p=3
    np.testing.assert_allclose(P.quantiles(x, q=4, axis=2), R.quantiles(x, q=4, axis=2), rtol=RTOL, equal_nan=True)
    v = np.array([2, 4, 4, 4, 5, 5, 7, 9])
    got = P.quantiles(v, probabilities=[0.005, 0.01, 0.02, 0.05, 0.1, 0.5])
    assert [round(float(q), 2) for q in got] == [2.07, 2.14, 2.28, 2.7, 3.4, 4.5]    # reference tests/test_math.py:9-14
    assert [round(float(q), 2) for q in P.quantiles(v, probabilities=4)] == [4, 4.5, 5.5]
    assert [round(float(q), 2) for q in P.quantiles(np.array([-1, -0.5, np.nan, 1]), probabilities=[2])] == [-0.5]
    q4 = P.quantiles(np.array([-1, -0.5, np.nan, 1]), q=4, ignore_nodata=False)
    assert len(q4) == 3 and np.all(np.isnan(q4))
    q5 = P.quantiles(np.array([]), probabilities=[0.1, 0.5])
    assert len(q5) == 2 and np.all(np.isnan(q5))
    with pytest.raises(P.exceptions.QuantilesParameterConflict):
        P.quantiles(v, probabilities=[0.5], q=2)
    with pytest.raises(P.exceptions.QuantilesParameterMissing):
        P.quantiles(v)


def test_spec_examples_through_the_gpu(P, spec_examples):
    """Every `examples` entry of the openEO specs the reference ships, through the CUDA path."""
    for name in ("mean", "median", "sd", "variance", "sum", "min", "max", "product"):
        for ex in spec_examples[name]:
            data = np.array([np.nan if v is None else v for v in ex["arguments"]["data"]], dtype=float)
            if len(data) == 0:
                continue  # empty input: host-side rule (nan / ValueError), covered in test_empty_inputs
            got = _fn(P, name)(data, ignore_nodata=ex["arguments"].get("ignore_nodata", True))
            want = ex["returns"]
            if want is None:
                if name in ("sum", "product") and ex["arguments"].get("ignore_nodata", True):
                    continue
                assert np.isnan(got), (name, ex)
            else:
                np.testing.assert_allclose(got, want, rtol=1e-12)


def test_empty_inputs(P):
    assert np.isnan(P._sum(np.array([]))) and np.isnan(P.product(np.array([])))
    with pytest.raises(ValueError):
        P._min(np.zeros((0, 4), np.float32), axis=0)
    assert np.isnan(P.mean(np.zeros((0, 4), np.float32), axis=0)).all()
    assert P.count(np.array([])) == 0


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint16, np.int32, np.int64])
def test_integer_inputs(P, dtype):
    rng = np.random.default_rng(5)
    x = rng.integers(0, 200, size=(20, 3, 6, 8)).astype(dtype)
    for name in ("sum", "mean", "min", "max", "sd", "variance", "median"):
        got = _fn(P, name)(x, axis=0)
        want = R.BY_NAME[name](x, axis=0)
        assert got.dtype == want.dtype, (name, got.dtype, want.dtype)
        np.testing.assert_allclose(got, want, rtol=1e-12)
    np.testing.assert_array_equal(P.count(x, axis=0), R.count(x, axis=0))


def test_golden_reference_outputs(P, golden):
    """The committed outputs of the reference's own functions (tests/golden/make_golden.py)."""
    for cname in ("int_valued_f32", "nan30_f32", "nan40_f64", "even_odd_f32"):
        x = golden[f"in/{cname}"]
        for axis in range(x.ndim):
            exact_sum = axis != x.ndim - 1
            for ign in (1, 0):
                for name in ("sum", "mean", "min", "max", "sd", "variance", "product", "median"):
                    want = golden[f"out/{cname}/{name}/axis{axis}/ign{ign}"]
                    got = _fn(P, name)(x, ignore_nodata=bool(ign), axis=axis)
                    _same(got, want, name in ("min", "max", "median") or exact_sum)
                want = golden[f"out/{cname}/quantiles3/axis{axis}/ign{ign}"]
                got = P.quantiles(x, probabilities=[0.1, 0.5, 0.9], ignore_nodata=bool(ign), axis=axis)
                np.testing.assert_allclose(got, want, rtol=RTOL, equal_nan=True)
            np.testing.assert_array_equal(P.count(x, axis=axis), golden[f"out/{cname}/count/axis{axis}/ign1"])
        nd = P.normalized_difference(np.ascontiguousarray(np.take(x, 0, axis=1)),
                                     np.ascontiguousarray(np.take(x, 1, axis=1)))
        _same(nd, golden[f"out/{cname}/nd01"], True)


def test_select_kernels_agree(P, torch, monkeypatch):
    """Register-network kernel vs the shared-memory sorter (two independent implementations)."""
    rng = np.random.default_rng(9)
    x = torch.from_numpy(_mk(rng, (100, 2, 16, 40), np.float32, 0.3)).cuda()
    a = P.median(x, axis=0)
    qa = P.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0)
    monkeypatch.setenv("EOC_SELECT_FORCE_SMEM", "1")
    b = P.median(x, axis=0)
    qb = P.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0)
    assert torch.equal(torch.nan_to_num(a, nan=-1.0), torch.nan_to_num(b, nan=-1.0))
    assert torch.equal(torch.nan_to_num(qa, nan=-1.0), torch.nan_to_num(qb, nan=-1.0))


def test_reduce_spatial_rows_kernel(P, torch):
    rng = np.random.default_rng(11)
    x = _mk(rng, (6, 3, 130, 257), np.float32, 0.3)
    d = torch.from_numpy(x).cuda()
    for name in ("sum", "mean", "min", "max", "sd", "variance"):
        for arg in (d, x):
            got = _fn(P, name)(arg, axis=(2, 3))
            want = R.BY_NAME[name](x.astype(np.float64), axis=(2, 3))
            _same(got, want.astype(np.float32), name in ("min", "max"))
    _same(P.count(d, axis=(2, 3)), R.count(x, axis=(2, 3)), True)
    xi = rng.integers(-100, 100, size=(30, 30, 20, 4)).astype(np.float32)   # reference tests/test_reduce.py:83-111
    np.testing.assert_array_equal(P._sum(xi, axis=(0, 1)), xi.sum(axis=(0, 1)))


def test_cube_processes_reference_tests(P, torch):
    """The reference's cube-level tests restated: tests/test_reduce.py:52-111, tests/test_arrays.py:582-657,
    tests/test_indices.py:17-83, tests/test_aggregate.py:22-96."""
    from openeo_processes_dask_b200.device_array import DeviceArray

    ref = P.ParameterReference(from_parameter="data")
    rng = np.random.default_rng(42)
    data = rng.integers(-100, 100, size=(30, 30, 20, 4)).astype(np.float32)
    t = pd.date_range("2018-05-01", "2018-06-01", periods=20).values
    for wrap in (lambda a: a, lambda a: DeviceArray.from_numpy(a)):
        cube = P.DataCube(wrap(data), ("x", "y", "t", "bands"),
                          {"x": np.arange(30.0), "y": np.arange(30.0), "t": t, "bands": ["B02", "B03", "B04", "B08"]},
                          attrs={"crs": "EPSG:4326"})
        out = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P.mean), ignore_nodata=True, data=ref),
                                 dimension="t")
        assert type(out.data) is type(cube.data) and out.dims == ("x", "y", "bands")
        np.testing.assert_array_equal(np.asarray(out.data), data.mean(axis=2))     # xr.testing.assert_equal there
        out = P.reduce_spatial(data=cube, reducer=functools.partial(P.process(P._sum), ignore_nodata=True, data=ref))
        np.testing.assert_array_equal(np.asarray(out.data), data.sum(axis=(0, 1)))
        cnt = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P.count), data=ref), dimension="bands")
        assert cnt.dims == ("x", "y", "t") and (np.asarray(cnt.data) == 4).all()
        cnt = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P.count), data=ref, condition=True),
                                 dimension="bands")
        assert (np.asarray(cnt.data) == 4).all()
        cnt = P.reduce_dimension(data=cube, dimension="bands", context={"y": -101},   # data >= -100
                                 reducer=functools.partial(P.process(P.count), data=ref, condition=P.process(P.gt)))
        assert (np.asarray(cnt.data) == 4).all()
        cnt = P.reduce_dimension(data=cube, dimension="bands",
                                 reducer=functools.partial(P.process(P.count), data=ref, condition=P.process(P.is_infinite)))
        assert (np.asarray(cnt.data) == 0).all()
        mx = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P._max), data=ref), dimension="t")
        np.testing.assert_array_equal(np.asarray(mx.data), data.max(axis=2))


def test_ndvi_and_fused_reduce(P, torch):
    from openeo_processes_dask_b200.device_array import DeviceArray, LazyNormalizedDifference

    rng = np.random.default_rng(42)
    data = rng.integers(-100, 100, size=(20, 20, 10, 2)).astype(np.float32)
    cube = P.DataCube(data, ("x", "y", "t", "b"), {"b": ["red", "nir"]}, attrs={"a": 1}, name="s2")
    cube = cube.assign_coords(common_name=("b", ["red", "nir"]))
    want = R.normalized_difference(data[..., 1], data[..., 0])
    out = P.ndvi(cube)
    assert "b" not in out.dims and out.attrs == cube.attrs
    np.testing.assert_allclose(out.data, want, rtol=RTOL, equal_nan=True)
    np.testing.assert_array_equal(np.isnan(out.data), np.isnan(want))
    renamed = cube.assign_coords(b=["blue", "yellow"])
    np.testing.assert_allclose(P.ndvi(renamed).data, want, rtol=RTOL, equal_nan=True)   # via common_name
    with pytest.raises(P.exceptions.DimensionAmbiguous):
        P.ndvi(out)
    with pytest.raises(P.exceptions.NirBandAmbiguous):
        P.ndvi(renamed.assign_coords(common_name=("b", ["blue", "red"])))
    with pytest.raises(P.exceptions.RedBandAmbiguous):
        P.ndvi(renamed.assign_coords(common_name=("b", ["nir", "yellow"])))
    with pytest.raises(KeyError):
        P.ndvi(renamed.drop_vars("common_name"))
    extra = P.ndvi(cube, target_band="yay")
    assert len(extra.dims) == len(out.dims) + 1 and len(extra.coords["b"]) == 3
    with pytest.raises(P.exceptions.BandExists):
        P.ndvi(cube, target_band="t" if "t" in cube.coords else "b")

    # fused ndvi -> max over t on a (t, bands, y, x) uint16 cube (config C4 in miniature)
    u = synth.synth_cube(45, (12, 4, 16, 32), np.uint16)
    dcube = P.DataCube(DeviceArray.from_numpy(u), ("t", "bands", "y", "x"), {"bands": ["b1", "red", "b3", "nir"]})
    lazy = P.ndvi(dcube.astype("float32"))   # the workflow's uint16 -> float32 cast, applied in the load stage
    assert isinstance(lazy.data, LazyNormalizedDifference)
    ref = P.ParameterReference(from_parameter="data")
    f = u.astype(np.float32)
    nd = R.normalized_difference(f[:, 3], f[:, 1])
    for name in ("max", "mean", "min"):
        red = P.reduce_dimension(data=lazy, reducer=functools.partial(P.process(_fn(P, name)), data=ref), dimension="t")
        _same(red.data.tensor, R.BY_NAME[name](nd, axis=0), True)
    _same(np.asarray(lazy.data), nd, True)


def test_aggregate_temporal_segmented(P, torch):
    """Config C5 in miniature: daily cube, monthly bins, mean + count, 40 % NaN; plus the on-edge quirk."""
    ref = P.ParameterReference(from_parameter="data")
    rng = np.random.default_rng(46)
    x = _mk(rng, (365, 2, 6, 16), np.float32, 0.4)
    months = pd.date_range("2021-01-01", "2022-01-01", freq="MS").strftime("%Y-%m-%dT%H:%M:%S")
    intervals = [[a, b] for a, b in zip(months[:-1], months[1:])]
    for hour in ("12:00:00", "00:00:00"):   # noon: no sample on an edge; midnight: Jan-01 falls out (right-closed)
        t = pd.date_range(f"2021-01-01T{hour}", periods=365, freq="D").values
        for kind in ("numpy", "cuda"):
            arr = torch.from_numpy(x).cuda() if kind == "cuda" else x
            cube = P.DataCube(arr, ("t", "bands", "y", "x"), {"t": t, "bands": ["a", "b"]}, attrs={"k": "v"})
            for name, oracle_fn in (("mean", R.mean), ("count", R.count), ("max", R._max), ("sd", R.sd)):
                red = functools.partial(P.process(_fn(P, name)), data=ref)
                out = P.aggregate_temporal(data=cube, intervals=intervals, reducer=red)
                want, _, wc = OC.aggregate_temporal(x, cube.dims, {"t": t}, intervals, oracle_fn)
                assert out.shape == (12, 2, 6, 16) and out.attrs == cube.attrs
                assert list(out["t"].values) == list(wc["t"]) and isinstance(out["t"].values[0], str)
                _same(out.data, want, True)
    # a gap between intervals and an empty bin -> NaN (count becomes float NaN like xarray's reindex)
    t = pd.date_range("2018-01-01", "2019-01-01", periods=100).values
    y = _mk(rng, (6, 5, 100, 4), np.float64, 0.1)
    cube = P.DataCube(y, ("x", "y", "t", "bands"), {"t": t})
    iv = [["2018-01-01T12:00:00", "2018-06-01T12:00:00"], ["2018-07-01T12:00:00", "2018-12-01T12:00:00"],
          ["2030-01-01T00:00:00", "2030-02-01T00:00:00"]]
    out = P.aggregate_temporal(data=cube, intervals=iv, reducer=functools.partial(P.process(P.mean), data=ref),
                               labels=["half-1", "half-2", "void"])
    want, _, _ = OC.aggregate_temporal(y, cube.dims, {"t": t}, iv, R.mean, ["half-1", "half-2", "void"])
    _same(out.data, want, True)
    assert np.isnan(out.data[:, :, 2]).all()
    med = P.aggregate_temporal(data=cube, intervals=iv[:2], reducer=functools.partial(P.process(P.median), data=ref))
    want, _, _ = OC.aggregate_temporal(y, cube.dims, {"t": t}, iv[:2], R.median)
    _same(med.data, want, True)
    per = P.aggregate_temporal_period(data=cube, reducer=functools.partial(P.process(P.mean), data=ref), period="month")
    assert per.shape[2] == 13


def test_device_array_numpy_protocol(P, torch):
    """Second seam: unmodified np.nan* calls on a DeviceArray run on the GPU."""
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(8)
    x = _mk(rng, (24, 3, 8, 16), np.float32, 0.3)
    d = DeviceArray.from_numpy(x)
    _same(np.nanmean(d, axis=0).tensor, np.nanmean(x, axis=0), True)
    _same(np.nanstd(d, axis=0, ddof=1).tensor, np.nanstd(x, axis=0, ddof=1), True)
    _same(np.nanmedian(d, axis=0).tensor, np.nanmedian(x, axis=0), True)
    _same(np.nanmax(d, axis=1).tensor, np.nanmax(x, axis=1), True)
    q = np.nanquantile(d, [0.25, 0.75], axis=0, method="linear")
    np.testing.assert_allclose(np.asarray(q), np.nanquantile(x, [0.25, 0.75], axis=0), rtol=RTOL, equal_nan=True)
    with pytest.raises(TypeError):
        np.cumsum(d)


def test_synth_fill_matches_numpy_twin(torch):
    from openeo_processes_dask_b200 import engine

    T, B, Y, X = 5, 3, 40, 33
    y0, y1 = 7, 29
    for dt, tdt, p in ((np.float32, torch.float32, 0.3), (np.uint16, torch.uint16, 0.0), (np.float64, torch.float64, 0.4)):
        out = torch.empty((T * B, (y1 - y0) * X), dtype=tdt, device="cuda")
        engine.synth_fill(out, T * B, Y * X, y0 * X, (y1 - y0) * X, seed=44, nan_prob=p)
        want = synth.synth_window(44, (T, B, Y, X), (slice(0, T), slice(0, B), slice(y0, y1), slice(0, X)), dt, p)
        got = out.cpu().view(torch.int16).numpy().view(np.uint16) if dt == np.uint16 else out.cpu().numpy()
        np.testing.assert_array_equal(got.reshape(T, B, y1 - y0, X), want)


def test_full_size_properties(P, torch):
    """BASELINE-sized columns (T=100, 30 % NaN) at a strip large enough to fill the GPU: properties
    that need no CPU oracle of the whole strip, plus oracle parity on random windows of it."""
    from openeo_processes_dask_b200 import engine

    T, B, Y, X = 100, 4, 10980, 10980
    rows = 24
    strip = torch.empty((T, B, rows, X), dtype=torch.float32, device="cuda")
    engine.synth_fill(strip.view(T * B, rows * X), T * B, Y * X, 5000 * X, rows * X, seed=43, nan_prob=0.30)
    med = P.median(strip, axis=0)
    mn, mx, mean = P._min(strip, axis=0), P._max(strip, axis=0), P.mean(strip, axis=0)
    q = P.quantiles(strip, probabilities=[0.1, 0.5, 0.9], axis=0)
    ok = ~torch.isnan(med)
    assert torch.equal(torch.isnan(med), torch.isnan(mean))
    assert bool((mn[ok] <= med[ok]).all()) and bool((med[ok] <= mx[ok]).all())
    assert bool((q[0][ok] <= q[1][ok]).all()) and bool((q[1][ok] <= q[2][ok]).all())
    assert torch.equal(q[1][ok].float(), med[ok])                         # the 0.5 quantile is the median
    # permutation invariance along t (order statistics) and linearity of the mean under negation
    perm = torch.randperm(T, device="cuda")
    assert torch.equal(torch.nan_to_num(P.median(strip[perm].contiguous(), axis=0), nan=-1),
                       torch.nan_to_num(med, nan=-1))
    assert torch.equal(torch.nan_to_num(P.mean(-strip, axis=0), nan=7.0), torch.nan_to_num(-mean, nan=7.0))
    cnt = P.count(strip, axis=0)
    assert abs(float(cnt.double().mean()) / T - 0.70) < 0.005             # 30 % NaN by construction
    # oracle parity on random 128-pixel windows regenerated from the hash on the CPU
    rng = np.random.default_rng(0)
    for _ in range(8):
        r, c = int(rng.integers(0, rows)), int(rng.integers(0, X - 128))
        win = synth.synth_window(43, (T, B, Y, X), (slice(0, T), slice(0, B), slice(5000 + r, 5001 + r), slice(c, c + 128)),
                                 np.float32, 0.30)
        np.testing.assert_array_equal(med[:, r, c:c + 128].cpu().numpy(), np.nanmedian(win, axis=0)[:, 0])
        np.testing.assert_array_equal(mean[:, r, c:c + 128].cpu().numpy(), np.nanmean(win, axis=0)[:, 0])


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.uint16, np.int32])
@pytest.mark.parametrize("shape,axis", [
    ((100, 2, 9, 128), 0),     # t of (t,b,y,x): one box per tile
    ((3, 100, 7, 52), 1),      # t of (b,t,y,x): outer > 1, ragged last tile (inner = 364)
    ((365, 1, 3, 200), 0),     # R > 256: two stacked TMA boxes, narrow tile
    ((2, 5, 40), 0),           # R = 2
    ((33, 1000), 0),
])
@pytest.mark.parametrize("ignore", [True, False])
def test_tma_tile_kernel(P, torch, monkeypatch, dtype, shape, axis, ignore):
    """variance/sd through the TMA tile kernel (forced on), against the oracle and against the
    two-pass column kernel (tile kernel off): bit-exact, every co-selected reducer included."""
    from openeo_processes_dask_b200 import _lib as L
    from openeo_processes_dask_b200 import engine

    rng = np.random.default_rng(abs(hash((shape, axis, str(dtype)))) % 2**32)
    x = _mk(rng, shape, dtype, 0.3 if ignore else 0.01)
    d = torch.from_numpy(x.view(np.int16) if dtype == np.uint16 else x).cuda()
    if dtype == np.uint16:
        d = d.view(torch.uint16)
    ops = [L.OP_SD, L.OP_VARIANCE, L.OP_MEAN, L.OP_SUM, L.OP_MIN, L.OP_MAX, L.OP_COUNT, L.OP_PRODUCT]
    monkeypatch.setenv("EOC_TILE", "1")
    engine.launch_count(reset=True)
    tile = engine.reduce(d, ops, axis, ignore_nodata=ignore)
    monkeypatch.setenv("EOC_TILE", "0")
    cols = engine.reduce(d, ops, axis, ignore_nodata=ignore)
    for op in ops:
        a, b = tile[op].cpu().numpy(), cols[op].cpu().numpy()
        np.testing.assert_array_equal(a, b, err_msg=f"op {op}")
    xf = x if np.dtype(dtype).kind == "f" else x.astype(np.float64)
    _same(tile[L.OP_SD], R.sd(xf, ignore_nodata=ignore, axis=axis), np.dtype(dtype).kind == "f")
    _same(tile[L.OP_VARIANCE], R.variance(xf, ignore_nodata=ignore, axis=axis), np.dtype(dtype).kind == "f")
    _same(tile[L.OP_MEAN], R.mean(xf, ignore_nodata=ignore, axis=axis), True)


def test_tma_tile_kernel_segmented(P, torch, monkeypatch):
    """aggregate_temporal-style segments (contiguous runs and gathered members) out of the resident tile."""
    from openeo_processes_dask_b200 import _lib as L
    from openeo_processes_dask_b200 import engine

    rng = np.random.default_rng(12)
    x = _mk(rng, (90, 3, 5, 64), np.float32, 0.35)
    d = torch.from_numpy(x).cuda()
    groups = [list(range(0, 31)), [], list(range(31, 59)), [80, 3, 64, 65, 7], list(range(59, 90))]
    ops = [L.OP_SD, L.OP_MEAN, L.OP_COUNT, L.OP_MAX]
    monkeypatch.setenv("EOC_TILE", "1")
    tile = engine.segmented_reduce(d, 0, groups, ops)
    monkeypatch.setenv("EOC_TILE", "0")
    cols = engine.segmented_reduce(d, 0, groups, ops)
    for op in ops:
        np.testing.assert_array_equal(tile[op].cpu().numpy(), cols[op].cpu().numpy())
    for gi, g in enumerate(groups):
        if g:
            _same(tile[L.OP_SD][gi], R.sd(x[g], axis=0), True)
        else:
            assert bool(torch.isnan(tile[L.OP_SD][gi]).all())


def test_tma_tile_full_size_sd(P, torch, monkeypatch):
    """C2/C3-shaped strip: sd from the single-pass TMA tile kernel == sd from the two-pass column kernel,
    and both equal the oracle on random windows."""
    from openeo_processes_dask_b200 import engine

    T, B, Y, X = 100, 4, 10980, 10980
    rows = 16
    strip = torch.empty((T, B, rows, X), dtype=torch.float32, device="cuda")
    engine.synth_fill(strip.view(T * B, rows * X), T * B, Y * X, 777 * X, rows * X, seed=43, nan_prob=0.30)
    monkeypatch.delenv("EOC_TILE", raising=False)
    engine.launch_count(reset=True)
    sd_tile = P.sd(strip, axis=0)
    monkeypatch.setenv("EOC_TILE", "0")
    sd_cols = P.sd(strip, axis=0)
    assert torch.equal(torch.nan_to_num(sd_tile, nan=-1.0), torch.nan_to_num(sd_cols, nan=-1.0))
    rng = np.random.default_rng(1)
    for _ in range(6):
        r, c = int(rng.integers(0, rows)), int(rng.integers(0, X - 128))
        win = synth.synth_window(43, (T, B, Y, X), (slice(0, T), slice(0, B), slice(777 + r, 778 + r), slice(c, c + 128)),
                                 np.float32, 0.30)
        np.testing.assert_array_equal(sd_tile[:, r, c:c + 128].cpu().numpy(), np.nanstd(win, axis=0, ddof=1)[:, 0])


@pytest.mark.parametrize("kind", ["numpy", "cuda"])
@pytest.mark.parametrize("dtype,shape,axis", [
    (np.float32, (100, 2, 5, 96), 0),    # N = 100 register kernel, fused
    (np.float32, (37, 3, 4, 33), 0),     # N = 64 with padding slots, ragged inner
    (np.float32, (3, 16, 130), 1),       # outer > 1
    (np.float32, (200, 2, 40), 0),       # R > 128: selection in shared memory + separate streaming launch
    (np.float64, (24, 2, 50), 0),        # float64: two launches
    (np.uint16, (12, 300), 0),
])
@pytest.mark.parametrize("ignore", [True, False])
def test_fused_select_and_streaming(P, torch, kind, dtype, shape, axis, ignore):
    """eoc_select_fused: quantiles + median + streaming reducers of the same column in one call equal the
    single processes (which equal the oracle), bit for bit."""
    from openeo_processes_dask_b200.process_implementations.math import reduce_many

    rng = np.random.default_rng(abs(hash((shape, axis, str(dtype), ignore))) % 2**32)
    x = _mk(rng, shape, dtype, 0.3 if ignore else 0.01)
    arg = torch.from_numpy(x).cuda() if kind == "cuda" else x
    probs = [0.1, 0.5, 0.9]
    specs = ["sd", ("quantiles", {"probabilities": probs}), "median", "mean", "max", "min", "sum", "variance"]
    got = reduce_many(arg, specs, ignore_nodata=ignore, axis=axis)
    xf = x if np.dtype(dtype).kind == "f" else x.astype(np.float64)
    exact = np.dtype(dtype).kind == "f"
    want = [R.sd(xf, ignore_nodata=ignore, axis=axis), R.quantiles(x, probabilities=probs, ignore_nodata=ignore, axis=axis),
            R.median(x, ignore_nodata=ignore, axis=axis), R.mean(xf, ignore_nodata=ignore, axis=axis),
            R._max(x, ignore_nodata=ignore, axis=axis), R._min(x, ignore_nodata=ignore, axis=axis),
            R._sum(x, ignore_nodata=ignore, axis=axis), R.variance(xf, ignore_nodata=ignore, axis=axis)]
    for name, g, w in zip(["sd", "quantiles", "median", "mean", "max", "min", "sum", "variance"], got, want):
        if name == "quantiles":
            g = np.asarray(g.cpu()) if hasattr(g, "cpu") else np.asarray(g)
            np.testing.assert_allclose(g, w, rtol=RTOL, equal_nan=True, err_msg=name)
        else:
            _same(g, w, exact or name in ("max", "min", "median", "sum"))
    # and the single processes agree exactly with the fused call
    single = [P.sd(arg, ignore_nodata=ignore, axis=axis), P.quantiles(arg, probabilities=probs, ignore_nodata=ignore, axis=axis),
              P.median(arg, ignore_nodata=ignore, axis=axis)]
    for g, sng in zip(got[:3], single):
        a = np.asarray(g.cpu()) if hasattr(g, "cpu") else np.asarray(g)
        b = np.asarray(sng.cpu()) if hasattr(sng, "cpu") else np.asarray(sng)
        np.testing.assert_array_equal(a, b)


def test_reduce_dimension_many(P, torch):
    from openeo_processes_dask_b200.process_implementations.cubes.reduce import reduce_dimension_many

    rng = np.random.default_rng(77)
    x = _mk(rng, (100, 3, 6, 64), np.float32, 0.3)
    cube = P.DataCube(torch.from_numpy(x).cuda(), ("t", "bands", "y", "x"), {"bands": ["a", "b", "c"]}, attrs={"k": 1})
    res = reduce_dimension_many(cube, {"q": ("quantiles", {"probabilities": [0.1, 0.5, 0.9]}), "s": "sd"}, "t")
    assert res["q"].dims == ("t", "bands", "y", "x") and res["q"].shape == (3, 3, 6, 64)
    assert res["s"].dims == ("bands", "y", "x") and res["s"].attrs["k"] == 1
    np.testing.assert_allclose(res["q"].values, R.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0), rtol=RTOL, equal_nan=True)
    _same(res["s"].data, R.sd(x, axis=0), True)


def test_sharded_spatial_reduce_single_rank(P, torch):
    """sharding.spatial_reduce with the kernels' partials (world size 1: no collective issued); the N = 2
    combine step itself is covered on the CPU over gloo (tests/test_sharding_gloo.py) and on two GPUs over
    NCCL by tests/multi_gpu_check.py."""
    from openeo_processes_dask_b200 import sharding

    rng = np.random.default_rng(21)
    x = _mk(rng, (5, 3, 70, 90), np.float32, 0.3)
    d = torch.from_numpy(x).cuda()
    flat = x.reshape(5, 3, -1).astype(np.float64)
    want = {"sum": np.nansum(flat, axis=2), "mean": np.nanmean(flat, axis=2), "min": np.nanmin(flat, axis=2),
            "max": np.nanmax(flat, axis=2), "sd": np.nanstd(flat, axis=2, ddof=1),
            "variance": np.nanvar(flat, axis=2, ddof=1), "count": np.sum(np.isfinite(flat), axis=2)}
    for op, w in want.items():
        got = sharding.spatial_reduce(op, d, (2, 3)).cpu().numpy()
        np.testing.assert_allclose(got, w, rtol=RTOL, equal_nan=True, err_msg=op)
    med = sharding.spatial_reduce("median", d[:, :, :10, :40].contiguous(), (2, 3))
    _same(med, np.nanmedian(x[:, :, :10, :40].reshape(5, 3, -1), axis=2), True)


@pytest.mark.parametrize("shape,axis", [
    ((100, 2, 9, 128), 0),     # N = 100, FULL
    ((3, 64, 7, 52), 1),       # outer > 1, N = 64 FULL, ragged last tile
    ((77, 1, 5, 200), 0),      # N = 100 with padding slots
    ((120, 260), 0),           # N = 128
    ((40, 3, 1000), 0),        # N = 64 with padding slots
])
@pytest.mark.parametrize("ignore", [True, False])
def test_tma_staged_selection_kernel(P, torch, monkeypatch, shape, axis, ignore):
    """k_select_tile (TMA-staged, forced on) == k_select_regs (forced off) == oracle, for median, quantiles
    and the fused streaming reducers."""
    from openeo_processes_dask_b200 import _lib as L
    from openeo_processes_dask_b200 import engine

    rng = np.random.default_rng(abs(hash((shape, axis, ignore))) % 2**32)
    x = _mk(rng, shape, np.float32, 0.3 if ignore else 0.01)
    d = torch.from_numpy(x).cuda()
    probs = [0.1, 0.5, 0.9, 1.0]
    ops = [L.OP_SD, L.OP_MEAN, L.OP_MAX, L.OP_COUNT]
    res = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("EOC_SELECT_TMA", flag)
        med, _ = engine.select(d, axis, ignore, True, None)
        _, qs = engine.select(d, axis, ignore, False, probs)
        fmed, fqs, fouts = engine.select(d, axis, ignore, True, probs, ops=ops)
        res[flag] = [med, qs, fmed, fqs] + [fouts[o] for o in ops]
    for a, b in zip(res["1"], res["0"]):
        np.testing.assert_array_equal(a.cpu().numpy(), b.cpu().numpy())
    med, qs, fmed, fqs, sd, mean, mx, cnt = res["1"]
    _same(med, R.median(x, ignore_nodata=ignore, axis=axis), True)
    _same(fmed, R.median(x, ignore_nodata=ignore, axis=axis), True)
    np.testing.assert_allclose(qs.cpu().numpy(), R.quantiles(x, probabilities=probs, ignore_nodata=ignore, axis=axis),
                               rtol=RTOL, equal_nan=True)
    _same(sd, R.sd(x, ignore_nodata=ignore, axis=axis), True)
    _same(mean, R.mean(x, ignore_nodata=ignore, axis=axis), True)
    _same(mx, R._max(x, ignore_nodata=ignore, axis=axis), True)
    _same(cnt, R.count(x, axis=axis), True)


def test_unsigned_normalized_difference_is_ieee_exhaustive(torch):
    """The uint8/uint16 fast path of eoc_nd_reduce (2^23 conversion trick + unguarded Newton division) is
    correctly rounded for EVERY pair of uint16 samples: compared on the device with __fdiv_rn, 2^32 pairs."""
    import ctypes as C

    from openeo_processes_dask_b200 import _lib as L

    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    L.check(L.lib().eoc_selftest_nd_u16(C.c_void_p(bad.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert int(bad.item()) == 0


# ---- first / last / _any / _all / array_element -------------------------------------------
def test_first_last_any_all_golden(P, torch, golden):
    """The committed outputs of the reference's own arrays.first / last, logic._any / _all (incl. the
    nan_to_num clamp quirk of first / last) on host and device inputs, every axis, both nodata modes."""
    fns = {"first": P.first, "last": P.last, "any": P._any, "all": P._all}
    for cname in ("truth_f32", "nan30_f32", "nan40_f64", "int_valued_f32"):
        x = golden[f"in/{cname}"]
        d = torch.from_numpy(x).cuda()
        for axis in range(x.ndim):
            for ign in (1, 0):
                for name, fn in fns.items():
                    want = golden[f"out/{cname}/{name}/axis{axis}/ign{ign}"]
                    for arg in (x, d):
                        got = fn(arg, ignore_nodata=bool(ign), axis=axis)
                        got = got.cpu().numpy() if hasattr(got, "cpu") else np.asarray(got)
                        assert got.dtype == want.dtype, (cname, name, axis, ign, got.dtype, want.dtype)
                        np.testing.assert_array_equal(got, want, err_msg=f"{cname}/{name}/axis{axis}/ign{ign}")
            want = golden[f"out/{cname}/element1/axis{axis}/ign1"]
            np.testing.assert_array_equal(P.array_element(x, index=1, axis=axis), want)
            np.testing.assert_array_equal(P.array_element(d, index=1, axis=axis).cpu().numpy(), want)


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.uint8, np.int16, np.uint16, np.int32])
def test_first_last_any_all_layouts(P, torch, dtype):
    rng = np.random.default_rng(21)
    if np.dtype(dtype).kind == "f":
        x = rng.choice(np.array([0.0, 1.0, -3.5, np.nan], dtype=dtype), size=(23, 3, 9, 40), p=[0.4, 0.3, 0.1, 0.2])
    else:
        x = rng.integers(0, 3, size=(23, 3, 9, 40)).astype(dtype)
    d = torch.from_numpy(x).cuda()
    for axis in (0, 1, 3):
        for ign in (True, False):
            for name, fn, ofn in (("first", P.first, R.first), ("last", P.last, R.last), ("any", P._any, R._any),
                                  ("all", P._all, R._all)):
                want = np.asarray(ofn(x, ignore_nodata=ign, axis=axis))
                got = fn(d, ignore_nodata=ign, axis=axis).cpu().numpy()
                assert got.dtype == want.dtype, (name, axis, ign, got.dtype, want.dtype)
                np.testing.assert_array_equal(got, want, err_msg=f"{name} axis{axis} ign{ign}")
    # first + last + any + all + mean + max of the same cube in one launch (engine level, C ABI eoc_reduce)
    from openeo_processes_dask_b200 import _lib as L, engine

    if np.dtype(dtype).kind == "f":
        ops = [L.OP_FIRST, L.OP_LAST, L.OP_ANY, L.OP_ALL, L.OP_MEAN, L.OP_MAX, L.OP_SD]
        out = engine.reduce(d, ops, 0)
        np.testing.assert_array_equal(out[L.OP_FIRST].cpu().numpy(), R.first(x, axis=0))
        np.testing.assert_array_equal(out[L.OP_LAST].cpu().numpy(), R.last(x, axis=0))
        np.testing.assert_array_equal(out[L.OP_ANY].cpu().numpy().astype(bool), R._any(x, axis=0))
        np.testing.assert_array_equal(out[L.OP_MAX].cpu().numpy(), R._max(x, axis=0))
        _same(out[L.OP_MEAN], R.mean(x, axis=0), True)
        _same(out[L.OP_SD], R.sd(x, axis=0), True)


def test_array_element_errors_and_labels(P, torch):
    x = np.arange(24, dtype=np.float32).reshape(4, 3, 2)
    d = torch.from_numpy(x).cuda()
    labels = np.array(["B02", "B03", "B04"])
    np.testing.assert_array_equal(P.array_element(d, label="B03", axis=1, dim_labels=labels).cpu().numpy(), x[:, 1])
    with pytest.raises(P.exceptions.ArrayElementParameterMissing):
        P.array_element(x)
    with pytest.raises(P.exceptions.ArrayElementParameterConflict):
        P.array_element(x, index=0, label="B02")
    with pytest.raises(P.exceptions.ArrayNotLabeled):
        P.array_element(x, label="B02", axis=1)
    with pytest.raises(P.exceptions.ArrayElementNotAvailable):
        P.array_element(d, index=7, axis=1)
    got = P.array_element(d, index=7, axis=1, return_nodata=True)
    assert tuple(got.shape) == (4, 2) and bool(got.isnan().all())
    got = P.array_element(x, label="B99", axis=1, dim_labels=labels, return_nodata=True)
    assert got.shape == (4, 2) and np.isnan(got).all()


def test_reduce_dimension_with_order_reducers(P, torch):
    """Process-graph style: @process-wrapped reducers with a ParameterReference, as pg-parser would call them."""
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(8)
    x = _mk(rng, (12, 4, 6, 50), np.float32, 0.4)
    ref = P.ParameterReference(from_parameter="data")
    for wrap in (lambda a: a, lambda a: DeviceArray.from_numpy(a)):
        cube = P.DataCube(wrap(x), ("t", "bands", "y", "x"),
                          {"t": np.arange(12), "bands": np.array(["B02", "B03", "B04", "B08"])})
        out = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P.last), data=ref), dimension="t")
        np.testing.assert_array_equal(np.asarray(out.data), R.last(x, axis=0))
        out = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P.first), data=ref, ignore_nodata=False),
                                 dimension="t")
        np.testing.assert_array_equal(np.asarray(out.data), x[0])
        out = P.reduce_dimension(data=cube, dimension="bands",
                                 reducer=functools.partial(P.process(P.array_element), data=ref, label="B08"))
        np.testing.assert_array_equal(np.asarray(out.data), x[:, 3])
        out = P.reduce_dimension(data=cube, reducer=functools.partial(P.process(P._any), data=ref), dimension="bands")
        np.testing.assert_array_equal(np.asarray(out.data), R._any(x, axis=1))
        out = P.reduce_dimension(data=cube, dimension="bands",
                                 reducer=functools.partial(P.process(P._all), data=ref, ignore_nodata=False))
        np.testing.assert_array_equal(np.asarray(out.data), R._all(x, ignore_nodata=False, axis=1))


# ---- mask fused into the reduce -------------------------------------------------------------
def _mask_case(rng, shape=(40, 3, 12, 64), nan_prob=0.2, cloud=0.35):
    x = _mk(rng, shape, np.float32, nan_prob)
    T_, B_, Y_, X_ = shape
    m_tyx = (rng.random((T_, Y_, X_)) < cloud)
    return x, m_tyx


@pytest.mark.parametrize("mask_dims", [("t", "y", "x"), ("t", "bands", "y", "x"), ("y", "x"), ("bands", "y", "x")])
@pytest.mark.parametrize("mask_dtype", [np.bool_, np.float32, np.uint8])
def test_mask_process_materialised(P, torch, mask_dims, mask_dtype):
    """mask() against the oracle's restatement of data.where(_not(mask), replacement): device, DeviceArray
    and numpy inputs, masks lacking bands / t, NaN inside a float mask (kept), integer data."""
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(31)
    dims = ("t", "bands", "y", "x")
    x = _mk(rng, (9, 3, 6, 20), np.float32, 0.2)
    mshape = tuple(x.shape[dims.index(d)] for d in mask_dims)
    m = (rng.random(mshape) < 0.4).astype(mask_dtype)
    if np.dtype(mask_dtype).kind == "f":
        m[rng.random(mshape) < 0.1] = np.nan          # NaN in the mask: keep the sample
        m[rng.random(mshape) < 0.1] = 2.5             # any non-zero value masks
    coords = {"t": np.arange(9), "bands": np.array(["B02", "B03", "B04"])}
    for wrap in (lambda a: a, lambda a: torch.from_numpy(a).cuda(), lambda a: DeviceArray.from_numpy(a)):
        for repl in (None, -1.0):
            cube = P.DataCube(wrap(x), dims, coords)
            mcube = P.DataCube(wrap(m), mask_dims, {k: v for k, v in coords.items() if k in mask_dims})
            out = P.mask(data=cube, mask=mcube, replacement=repl)
            want, want_dims = OC.mask(x, dims, m, mask_dims, repl)
            # same array kind out as in (reference tests/general_checks.py:19); a bare CUDA tensor -- not a kind the
            # reference knows -- comes back as a DeviceArray-typed lazy result
            assert out.dims == want_dims
            assert isinstance(out.data, DeviceArray if torch.is_tensor(cube.data) else type(cube.data))
            got = out.values
            assert got.dtype == want.dtype
            np.testing.assert_array_equal(got, want)
    xi = rng.integers(0, 10000, size=(9, 3, 6, 20)).astype(np.uint16)
    for repl in (None, 0):
        out = P.mask(data=P.DataCube(torch.from_numpy(xi).cuda(), dims, coords),
                     mask=P.DataCube(torch.from_numpy(m).cuda(), mask_dims), replacement=repl)
        want, _ = OC.mask(xi, dims, m, mask_dims, repl)
        assert out.values.dtype == want.dtype
        np.testing.assert_array_equal(out.values, want)


def test_mask_process_errors(P, torch):
    x = np.zeros((4, 2, 3, 5), np.float32)
    dims = ("t", "bands", "y", "x")
    cube = P.DataCube(x, dims, {"t": np.arange(4)})
    with pytest.raises(P.exceptions.DimensionMismatch):       # spatial dims differ (tests/test_mask.py:66-68)
        P.mask(data=cube, mask=P.DataCube(np.zeros((4, 2, 3), bool), ("t", "bands", "y")))
    with pytest.raises(P.exceptions.DimensionLabelCountMismatch):
        P.mask(data=cube, mask=P.DataCube(np.zeros((3, 3, 5), bool), ("t", "y", "x"), {"t": np.arange(3)}))
    with pytest.raises(P.exceptions.LabelMismatch):
        P.mask(data=cube, mask=P.DataCube(np.zeros((4, 3, 5), bool), ("t", "y", "x"), {"t": np.arange(4) + 1}))


@pytest.mark.parametrize("mask_dims", [("t", "y", "x"), ("t", "bands", "y", "x"), ("y", "x")])
def test_mask_fused_into_reducers(P, torch, mask_dims):
    """mask(..., fuse=True) -> reduce_dimension: streaming reducers (eoc_masked_reduce), median / quantiles
    (eoc_masked_select, TMA tile + cp.async mask tile) and the materialise-first fallback all equal the
    oracle reducing the oracle's masked cube; bit-exact wherever the unfused path is."""
    rng = np.random.default_rng(41)
    dims = ("t", "bands", "y", "x")
    x = _mk(rng, (100, 3, 10, 96), np.float32, 0.2)
    mshape = tuple(x.shape[dims.index(d)] for d in mask_dims)
    m = rng.random(mshape) < 0.35
    masked, _ = OC.mask(x, dims, m, mask_dims, None)
    cube = P.DataCube(torch.from_numpy(x).cuda(), dims, {"t": np.arange(100)})
    mcube = P.DataCube(torch.from_numpy(m).cuda(), mask_dims, {"t": np.arange(100)} if "t" in mask_dims else {})
    lazy = P.mask(data=cube, mask=mcube, fuse=True)
    from openeo_processes_dask_b200.device_array import LazyMasked

    if "t" in mask_dims:   # (a mask without t transposes the cube; dims stay (t,bands,y,x) here anyway)
        assert isinstance(lazy.data, LazyMasked)
    ref = P.ParameterReference(from_parameter="data")
    for dim, axis in (("t", 0), ("bands", 1)):
        for name in ("mean", "sum", "min", "max", "sd", "variance", "median"):
            for ign in (True, False):
                red = P.reduce_dimension(data=lazy, dimension=dim,
                                         reducer=functools.partial(P.process(_fn(P, name)), data=ref, ignore_nodata=ign))
                want = R.BY_NAME[name](masked, ignore_nodata=ign, axis=axis)
                _same(red.data, want, True)
        red = P.reduce_dimension(data=lazy, dimension=dim, reducer=functools.partial(P.process(P.count), data=ref))
        np.testing.assert_array_equal(red.data.cpu().numpy(), R.count(masked, axis=axis))
    q = P.quantiles(lazy.data, probabilities=[0.1, 0.5, 0.9], axis=0)
    want = R.quantiles(masked, probabilities=[0.1, 0.5, 0.9], axis=0)
    assert tuple(q.shape) == want.shape
    np.testing.assert_allclose(q.cpu().numpy(), want, rtol=RTOL, equal_nan=True)
    np.testing.assert_array_equal(np.asarray(lazy.values), masked)


def test_masked_median_full_width_strip(P, torch):
    """The composite workflow at BASELINE width: (t=100, bands=4, rows, x=10980) strip, (t, y, x) cloud mask;
    fused one-pass median == median of the materialised masked cube, bit for bit, and == oracle on windows."""
    from openeo_processes_dask_b200 import engine

    T_, B_, Y_, X_ = 100, 4, 8, 10980
    x = torch.empty((T_, B_, Y_, X_), dtype=torch.float32, device="cuda")
    engine.synth_fill(x.view(T_ * B_, Y_ * X_), T_ * B_, 10980 * X_, 0, Y_ * X_, 43, 0.10)
    g = torch.Generator(device="cuda").manual_seed(5)
    m = torch.rand((T_, Y_, X_), device="cuda", generator=g) < 0.25
    cube = P.DataCube(x, ("t", "bands", "y", "x"), {"t": np.arange(T_)})
    mcube = P.DataCube(m, ("t", "y", "x"), {"t": np.arange(T_)})
    fused = P.median(P.mask(data=cube, mask=mcube, fuse=True).data, axis=0)
    plain = P.median(P.mask(data=cube, mask=mcube).data, axis=0)
    assert torch.equal(torch.nan_to_num(fused, nan=-7.0), torch.nan_to_num(plain, nan=-7.0))
    xs = x[:, :, :2, 5000:5200].cpu().numpy()
    ms = m[:, :2, 5000:5200].cpu().numpy()
    want = np.nanmedian(np.where(ms[:, None], np.nan, xs), axis=0)
    np.testing.assert_array_equal(fused[:, :2, 5000:5200].cpu().numpy(), want)


# ---- apply_dimension with the cumulative array processes -----------------------------------
def test_cumulative_processes_golden(P, torch, golden):
    """cumsum / cumproduct / cummin / cummax against the reference's own outputs: bit-exact (the scan kernel
    keeps numpy's sequential order), host and device inputs, every axis, both nodata modes."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2] in ("cumsum", "cumproduct", "cummin", "cummax"):
            x = golden["in/" + parts[1]]
            kw = dict(ignore_nodata=bool(int(parts[4][3:])), axis=int(parts[3][4:]))
            want = golden[key]
            got = getattr(P, parts[2])(torch.from_numpy(x).cuda(), **kw).cpu().numpy()
            assert got.dtype == want.dtype
            np.testing.assert_array_equal(got, want, err_msg=key)
            if n % 8 == 0:
                np.testing.assert_array_equal(getattr(P, parts[2])(x, **kw), want, err_msg=key + " (numpy in)")
            n += 1
    assert n == 96


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.int16, np.uint16, np.int32])
def test_cumulative_processes_layouts(P, torch, dtype):
    rng = np.random.default_rng(17)
    x = _mk(rng, (33, 3, 7, 52), dtype, 0.3) if np.dtype(dtype).kind == "f" else \
        rng.integers(0, 50, size=(33, 3, 7, 52)).astype(dtype)
    d = torch.from_numpy(x).cuda()
    for axis in (0, 1, 3):
        for name in ("cumsum", "cummin", "cummax") + (("cumproduct",) if np.dtype(dtype).kind == "f" else ()):
            for ign in (True, False):
                want = getattr(R, name)(x, ignore_nodata=ign, axis=axis)
                got = getattr(P, name)(d, ignore_nodata=ign, axis=axis).cpu().numpy()
                assert got.dtype == want.dtype, (name, got.dtype, want.dtype)
                np.testing.assert_array_equal(got, want, err_msg=f"{name} axis{axis} ign{ign}")
    sl = d[:, 1]                                   # a band selection: strided view, no copy
    np.testing.assert_array_equal(P.cumsum(sl, axis=0).cpu().numpy(), R.cumsum(x[:, 1], axis=0))
    np.testing.assert_array_equal(P.cumsum(d, axis=None).cpu().numpy(), R.cumsum(x, axis=None))  # numpy flattens


def test_apply_dimension(P, torch):
    """reference tests/test_apply.py:54-181 in miniature: axis kept (cumsum), axis resized (quantiles with
    several probabilities -> labels 0..n-1), target_dimension renaming, reducers through apply_dimension."""
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(23)
    x = _mk(rng, (6, 5, 30, 4), np.float32, 0.2)          # (x, y, t, bands) like the reference fixture
    ref = P.ParameterReference(from_parameter="data")
    t = np.arange("2018-05-01", "2018-05-31", dtype="datetime64[D]")
    for wrap in (lambda a: a, lambda a: DeviceArray.from_numpy(a)):
        cube = P.DataCube(wrap(x), ("x", "y", "t", "bands"), {"t": t, "bands": ["B02", "B03", "B04", "B08"]})
        out = P.apply_dimension(data=cube, process=functools.partial(P.process(P.cumsum), data=ref), dimension="t")
        assert out.dims == cube.dims and type(out.data) is type(cube.data)
        np.testing.assert_array_equal(out["t"].values, t)
        np.testing.assert_array_equal(out.values, R.cumsum(x, axis=2))
        # quantiles p=4 -> 3 values along t, relabelled 0..2 (tests/test_apply.py:158-181: shape (6,5,3,4))
        out = P.apply_dimension(data=cube, dimension="t",
                                process=functools.partial(P.process(P.quantiles), data=ref, probabilities=4))
        assert out.shape == (6, 5, 3, 4) and out.dims == cube.dims
        np.testing.assert_array_equal(out["t"].values, np.arange(3))
        np.testing.assert_allclose(out.values, R.quantiles(x, probabilities=4, axis=2), rtol=RTOL, equal_nan=True)
        # a reducer through apply_dimension with a target dimension: one label, renamed
        out = P.apply_dimension(data=cube, dimension="t", target_dimension="stat",
                                process=functools.partial(P.process(P.mean), data=ref))
        assert out.dims == ("x", "y", "stat", "bands") and out.shape == (6, 5, 1, 4)
        _same(out.values[:, :, 0], R.mean(x, axis=2), False)
        out = P.apply_dimension(data=cube, dimension="bands",
                                process=functools.partial(P.process(P.cummax), data=ref, ignore_nodata=False))
        np.testing.assert_array_equal(out.values, R.cummax(x, ignore_nodata=False, axis=3))
    with pytest.raises(P.exceptions.DimensionNotAvailable):
        P.apply_dimension(data=cube, process=P.process(P.cumsum), dimension="nope")


def test_sort_golden_and_layouts(P, torch, golden):
    """eoc_sort against the reference's outputs (apply_dimension convention) and the oracle: TMA tile kernel,
    register kernel, shared-memory sorter (float64 R > 64, integers), asc / desc, NaNs last / first / dropped."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2].startswith("sort_"):
            x = golden["in/" + parts[1]]
            kw = dict(asc=parts[2][8] == "1", nodata=parts[2][-1] == "1", axis=int(parts[3][4:]))
            got = P.sort(torch.from_numpy(x).cuda(), **kw).cpu().numpy()
            assert got.dtype == golden[key].dtype
            np.testing.assert_array_equal(got, golden[key], err_msg=key)
            n += 1
    assert n == 32
    rng = np.random.default_rng(29)
    for shape, dtype, axis in (((100, 2, 8, 300), np.float32, 0), ((70, 2, 40), np.float64, 0), ((5, 40, 30), np.float32, 1),
                               ((12, 6, 50), np.int16, 0), ((200, 3, 33), np.float32, 0)):
        x = _mk(rng, shape, dtype, 0.3) if np.dtype(dtype).kind == "f" else rng.integers(-50, 50, size=shape).astype(dtype)
        d = torch.from_numpy(x).cuda()
        for asc in (True, False):
            for nd in (True, False):
                got = P.sort(d, asc=asc, nodata=nd, axis=axis).cpu().numpy()
                want = R.sort(x, asc=asc, nodata=nd, axis=axis)
                assert got.dtype == want.dtype
                np.testing.assert_array_equal(got, want, err_msg=f"{shape} {dtype} asc={asc} nodata={nd}")
    v = np.array([6, -1, 2, np.nan, 7, 4, np.nan, 8, 3, 9, 9])
    np.testing.assert_array_equal(P.sort(v), [-1, 2, 3, 4, 6, 7, 8, 9, 9])                      # nodata=None drops
    np.testing.assert_array_equal(P.sort(torch.from_numpy(v).cuda(), asc=False).cpu().numpy(), [9, 9, 8, 7, 6, 4, 3, 2, -1])
    big = torch.from_numpy(_mk(rng, (100, 4, 40, 512), np.float32, 0.3)).cuda()   # large enough for the TMA tile kernel
    got = P.sort(big, nodata=True, axis=0)
    want = torch.sort(torch.where(big.isnan(), torch.full_like(big, float("inf")), big), dim=0).values
    nan_count = big.isnan().sum(dim=0)
    idx = torch.arange(100, device="cuda").view(100, 1, 1, 1)
    assert torch.equal(torch.where(idx < 100 - nan_count, got, torch.zeros_like(got)),
                       torch.where(idx < 100 - nan_count, want, torch.zeros_like(want)))
    assert bool(torch.where(idx >= 100 - nan_count, got.isnan(), torch.ones_like(got, dtype=torch.bool)).all())
    ref = P.ParameterReference(from_parameter="data")
    cube = P.DataCube(d, ("t", "y", "x"), {"t": np.arange(200)})
    out = P.apply_dimension(data=cube, dimension="t",
                            process=functools.partial(P.process(P.sort), data=ref, asc=False, nodata=True))
    np.testing.assert_array_equal(out.values, R.sort(x, asc=False, nodata=True, axis=0))


def test_array_interpolate_linear(P, torch, golden):
    """eoc_interpolate_linear: bit-exact vs the reference's outputs and vs the oracle on other layouts (np.interp
    in float64, cast back), unequal label spacing, host and device inputs, apply_dimension over t."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2] == "interp":
            x = golden["in/" + parts[1]]
            axis = int(parts[3][4:])
            got = P.array_interpolate_linear(torch.from_numpy(x).cuda(), axis=axis).cpu().numpy()
            assert got.dtype == golden[key].dtype
            np.testing.assert_array_equal(got, golden[key], err_msg=key)
            n += 1
    assert n == 9
    rng = np.random.default_rng(37)
    for shape, dtype, axis, p in (((100, 3, 9, 140), np.float32, 0, 0.3), ((365, 2, 70), np.float32, 0, 0.6),
                                  ((6, 50, 33), np.float64, 1, 0.4), ((4, 5, 130), np.float32, 2, 0.9)):
        x = _mk(rng, shape, dtype, p)
        want = R.array_interpolate_linear(x, axis=axis)
        got = P.array_interpolate_linear(torch.from_numpy(x).cuda(), axis=axis).cpu().numpy()
        np.testing.assert_array_equal(got, want, err_msg=str(shape))
        np.testing.assert_array_equal(P.array_interpolate_linear(x, axis=axis), want)
    x = _mk(rng, (40, 6, 64), np.float32, 0.4)
    days = np.sort(rng.choice(np.arange(400), size=40, replace=False)).astype(float)   # irregular acquisition dates
    want = R.array_interpolate_linear(x, axis=0, positions=days)
    got = P.array_interpolate_linear(torch.from_numpy(x).cuda(), axis=0, dim_labels=list(days)).cpu().numpy()
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(P.array_interpolate_linear([np.nan, 1, np.nan, 6, np.nan, -8]), [np.nan, 1, 3.5, 6, -1, -8])
    xi = rng.integers(0, 9, size=(5, 7)).astype(np.int16)
    assert P.array_interpolate_linear(xi, axis=1) is xi
    ref = P.ParameterReference(from_parameter="data")
    x = _mk(rng, (30, 4, 8, 96), np.float32, 0.35)
    cube = P.DataCube(torch.from_numpy(x).cuda(), ("t", "bands", "y", "x"), {"t": np.arange(30) * 5})
    out = P.apply_dimension(data=cube, dimension="t", process=functools.partial(P.process(P.array_interpolate_linear), data=ref))
    np.testing.assert_array_equal(out.values, R.array_interpolate_linear(x, axis=0))   # positions 0..n-1, like the reference
    np.testing.assert_array_equal(out["t"].values, np.arange(30) * 5)


def test_reduce_spatial_sd_with_nodata_corner(P, torch):
    """the long-row variance shifts its sums by the first FINITE sample of the row; with a NaN corner the
    round-1 kernel shifted by 0 and lost the variance of data with |mean| / sd ~ 1e6 to cancellation."""
    rng = np.random.default_rng(77)
    x = (1.0e6 + rng.normal(size=(3, 2, 300, 400))).astype(np.float64)
    x[:, :, :40, :50] = np.nan               # nodata corner: the first sample of every plane is NaN
    x[1, 0, 0, :] = np.inf                   # and a first row of infinities
    cube = P.DataCube(torch.from_numpy(x).cuda(), ("t", "bands", "y", "x"), {})
    ref = P.ParameterReference(from_parameter="data")
    for name in ("sd", "variance", "mean"):
        out = P.reduce_spatial(data=cube, reducer=functools.partial(P.process(_fn(P, name)), data=ref))
        flat = x.reshape(3, 2, -1)
        with np.errstate(invalid="ignore"):
            want = R.BY_NAME[name](flat, axis=2)
        np.testing.assert_allclose(np.asarray(out.data.cpu()), want, rtol=1e-6, equal_nan=True)


def test_reduce_row_slice_without_copy(P, torch):
    """A row slice cube[:, :, y0:y1] of a (t,bands,y,x) cube reduced over t is ONE strided view once the bands
    join `outer` (views.plan_view_pair): no contiguous copy, results identical."""
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.views import plan_view

    rng = np.random.default_rng(51)
    x = _mk(rng, (60, 3, 20, 128), np.float32, 0.3)
    d = torch.from_numpy(x).cuda()
    sl = d[:, :, 4:12, :]
    assert plan_view(tuple(sl.shape), tuple(sl.stride()), 0) is None
    _, v = engine.prepare(sl, 0, allow_tail=True)
    assert (v.outer, v.R, v.inner, v.stride_outer, v.stride_R) == (3, 60, 8 * 128, 20 * 128, 3 * 20 * 128)
    for name in ("mean", "sd", "max", "sum"):
        _same(_fn(P, name)(sl, axis=0), R.BY_NAME[name](x[:, :, 4:12, :], axis=0), True)
    np.testing.assert_array_equal(P.count(sl, axis=0).cpu().numpy(), R.count(x[:, :, 4:12, :], axis=0))
    np.testing.assert_array_equal(P.median(sl, axis=0).cpu().numpy(), R.median(x[:, :, 4:12, :], axis=0))


def test_elementwise_apply(P, torch, golden):
    """clip / linear_scale_range through eoc_elementwise: bit-exact vs the reference's outputs (numpy's float32
    operation order with weak python scalars), reference-test vectors, and `apply` over a cube."""
    cases = (("lsr_0_5000", lambda x: P.linear_scale_range(x, 0, 5000)),
             ("lsr_swapped_255", lambda x: P.linear_scale_range(x, 5000, 0, 0, 255)),
             ("lsr_float", lambda x: P.linear_scale_range(x, -0.3, 2500.7, 0.1, 0.9)),
             ("clip_1000_4000", lambda x: P.clip(x, 1000, 4000)), ("clip_float", lambda x: P.clip(x, -0.25, 0.75)))
    for cname in ("nan30_f32", "nan40_f64", "int_valued_f32"):
        x = golden["in/" + cname]
        for key, fn in cases:
            want = golden[f"out/{cname}/{key}"]
            for arg in (torch.from_numpy(x).cuda(), x):
                got = fn(arg)
                got = got.cpu().numpy() if hasattr(got, "cpu") else got
                assert got.dtype == want.dtype, (cname, key)
                np.testing.assert_array_equal(got, want, err_msg=f"{cname}/{key}")
    np.testing.assert_array_equal(P.linear_scale_range(np.array([5, 0]), inputMin=0, inputMax=5), [1, 0])
    np.testing.assert_array_equal(P.linear_scale_range(np.array([5, 0]), inputMin=5, inputMax=0), [0, 1])
    got = P.clip(np.array([5, 0]), min=0, max=2)
    assert got.dtype == np.array([5, 0]).dtype
    np.testing.assert_array_equal(got, [2, 0])
    with pytest.raises(P.exceptions.MinMaxSwapped):
        P.clip(np.array([5, 0]), 10, 0)
    rng = np.random.default_rng(61)
    x = _mk(rng, (7, 3, 11, 33), np.float32, 0.2)       # odd length: vector body + scalar tail
    cube = P.DataCube(torch.from_numpy(x).cuda(), ("t", "bands", "y", "x"), {"t": np.arange(7)})
    ref = P.ParameterReference(from_parameter="x")
    out = P.apply(data=cube, process=functools.partial(P.process(P.linear_scale_range), x=ref, inputMin=0, inputMax=8000,
                                                       outputMax=255))
    assert out.dims == cube.dims
    np.testing.assert_array_equal(out.values, R.linear_scale_range(x, 0, 8000, 0.0, 255))


def _scalar_or_array(a):
    return a.item() if a.ndim == 0 else a


BINARY_PROCESSES = {"add": "add", "subtract": "subtract", "multiply": "multiply", "divide": "divide", "eq": "eq",
                    "neq": "neq", "gt": "gt", "gte": "gte", "lt": "lt", "lte": "lte", "and": "_and", "or": "_or",
                    "xor": "xor"}


def test_binary_processes_equal_reference_outputs(P, torch, golden_binary):
    """eoc_binary through the process functions: bit-exact (values, NaN masks, dtypes) vs the outputs of the
    reference's own add .. xor on equal shapes, broadcast planes / slabs, python scalars, wrapping integers and
    strided views; numpy in -> numpy out and CUDA in -> CUDA out."""
    g = golden_binary
    cases = sorted({k.split("/")[1] for k in g.files if k.startswith("in/")})
    for c in cases:
        x, y = _scalar_or_array(g[f"in/{c}/x"]), _scalar_or_array(g[f"in/{c}/y"])
        for name, pname in BINARY_PROCESSES.items():
            want = g[f"out/{c}/{name}"]
            fn = getattr(P, pname)
            dev = lambda v: torch.from_numpy(np.ascontiguousarray(v)).cuda() if isinstance(v, np.ndarray) else v
            for a, b in ((x, y), (dev(x), dev(y))):
                got = fn(a, b)
                if hasattr(got, "cpu"):
                    got = got.cpu().numpy()
                assert got.dtype == want.dtype, (c, name, got.dtype, want.dtype)
                np.testing.assert_array_equal(got, want, err_msg=f"{c}/{name}")


def test_binary_views_and_oracle(P, torch):
    """Non-contiguous device views (no copy), odd lengths (vector body + tail), 5-D broadcasting vs the oracle."""
    rng = np.random.default_rng(77)
    x = _mk(rng, (6, 3, 5, 37), np.float32, 0.2)
    y = _mk(rng, (6, 3, 5, 37), np.float32, 0.2)
    tx, ty = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for name, pname in BINARY_PROCESSES.items():
        fn, ofn = getattr(P, pname), R.BINARY[name]
        for (a, b), (oa, ob) in (((tx[:, 1], ty[:, 2]), (x[:, 1], y[:, 2])),
                                 ((tx[1:, :, :, 3:], ty[1:, :1, :, 3:]), (x[1:, :, :, 3:], y[1:, :1, :, 3:])),
                                 ((tx.permute(1, 0, 2, 3), ty[0, :, None, :, :]), (x.transpose(1, 0, 2, 3), y[0, :, None, :, :])),
                                 ((tx[None], ty[2, 1, 3]), (x[None], y[2, 1, 3]))):
            got = fn(a, b).cpu().numpy()
            want = np.asarray(ofn(oa, ob))
            assert got.dtype == want.dtype and got.shape == want.shape, (name, got.shape, want.shape)
            np.testing.assert_array_equal(got, want, err_msg=name)
    # DeviceArray in -> DeviceArray out; scalar op scalar stays on the device path
    from openeo_processes_dask_b200.device_array import DeviceArray

    out = P.add(DeviceArray(tx), 1.5)
    assert isinstance(out, DeviceArray)
    np.testing.assert_array_equal(np.asarray(out), x + np.float32(1.5))
    assert P.add(2, 3) == 5 and P.divide(1, 4) == 0.25


def _cube_of(P, torch, arr, dims, coords, device=True):
    data = torch.from_numpy(np.ascontiguousarray(arr)).cuda() if device else arr
    return P.DataCube(data, dims, {k: np.asarray(v) for k, v in coords.items()}, attrs={"crs": "EPSG:32633"})


def test_merge_cubes(P, torch):
    """merge_cubes (cubes/merge.py:54-295): the four spec examples the reference tests (tests/test_merge.py:22-150,
    test_logic.py:172-222, test_comparison.py:377-433) vs the oracle restatement, device and numpy cubes."""
    rng = np.random.default_rng(31)
    dims = ("x", "y", "t", "bands")                    # the reference fixture's layout
    full = _mk(rng, (6, 5, 4, 4), np.float64, 0.1)
    coords = {"x": np.arange(6) * 10.0, "y": np.arange(5) * 10.0, "t": np.arange("2021-01-01", "2021-01-05", dtype="datetime64[D]"),
              "bands": np.array(["B02", "B03", "B04", "--324"])}
    ref_x, ref_y = P.ParameterReference(from_parameter="x"), P.ParameterReference(from_parameter="y")
    add = functools.partial(P.process(P.add), x=ref_x, y=ref_y)

    def sub(b_idx):
        c = dict(coords)
        c["bands"] = coords["bands"][b_idx]
        return full[..., b_idx], c

    for device in (True, False):
        mk = lambda arr, d, c: _cube_of(P, torch, arr, d, c, device)
        # example 1: disjoint bands -> the original cube back (test_merge_cubes_type_1)
        (a1, c1), (a2, c2) = sub([0, 1]), sub([2, 3])
        out = P.merge_cubes(mk(a1, dims, c1), mk(a2, dims, c2))
        want, wdims, wcoords = OC.merge_cubes((a1, dims, c1), (a2, dims, c2))
        assert out.dims == wdims == dims
        np.testing.assert_array_equal(out["bands"].values, coords["bands"])
        np.testing.assert_array_equal(out.values, full)
        np.testing.assert_array_equal(out.values, want)
        assert out.attrs == {"crs": "EPSG:32633"}
        # disjoint time steps, given in reverse order: joined ascending by label
        ta, tb = full[:, :, 2:], full[:, :, :2]
        ca, cb = dict(coords, t=coords["t"][2:]), dict(coords, t=coords["t"][:2])
        out = P.merge_cubes(mk(ta, dims, ca), mk(tb, dims, cb))
        np.testing.assert_array_equal(out.values, full)
        np.testing.assert_array_equal(out["t"].values, coords["t"])
        # example 2: bands overlap partially -> resolver on the common band (test_merge_cubes_type_2)
        (a1, c1), (a2, c2) = sub([0, 1, 2]), sub([1, 2, 3])
        with pytest.raises(P.exceptions.OverlapResolverMissing):
            P.merge_cubes(mk(a1, dims, c1), mk(a2, dims, c2))
        out = P.merge_cubes(mk(a1, dims, c1), mk(a2, dims, c2), overlap_resolver=add)
        want, wdims, wcoords = OC.merge_cubes((a1, dims, c1), (a2, dims, c2), R.add)
        np.testing.assert_array_equal(out["bands"].values, wcoords["bands"])
        np.testing.assert_array_equal(out.values, want)
        b03 = list(out["bands"].values).index("B03")
        np.testing.assert_array_equal(out.values[..., b03], full[..., 1] * 2)
        # example 3: identical cubes (test_merge_cubes_type_3, test_merge_cubes_eq, test_logic.test_merge_cubes)
        out = P.merge_cubes(mk(full, dims, coords), mk(full, dims, coords))
        assert out.dims == ("__cubes__",) + dims and list(out["__cubes__"].values) == ["cube1", "cube2"]
        np.testing.assert_array_equal(out.values, np.stack([full, full]))
        out = P.merge_cubes(mk(full, dims, coords), mk(full, dims, coords), add)
        np.testing.assert_array_equal(out.values, full * 2)
        eq = P.merge_cubes(mk(full, dims, coords), mk(full, dims, coords), functools.partial(P.process(P.eq), x=ref_x, y=ref_y))
        neq = P.merge_cubes(mk(full, dims, coords), mk(full, dims, coords), functools.partial(P.process(P.neq), x=ref_x, y=ref_y))
        lt = P.merge_cubes(mk(full, dims, coords), mk(full, dims, coords), functools.partial(P.process(P.lt), x=ref_x, y=ref_y))
        np.testing.assert_array_equal(eq.values, neq.values + 1)
        np.testing.assert_array_equal(lt.values, neq.values)
        np.testing.assert_array_equal(eq.values, R.eq(full, full))
        ones, zeros = np.ones((6, 5, 4)), np.zeros((6, 5, 4))
        c3 = {k: v for k, v in coords.items() if k != "bands"}
        for pname, expect in (("_and", zeros), ("_or", ones), ("xor", ones)):
            res = functools.partial(P.process(getattr(P, pname)), x=ref_x, y=ref_y)
            out = P.merge_cubes(mk(ones, dims[:3], c3), mk(zeros, dims[:3], c3), overlap_resolver=res)
            assert out.dims == ("x", "y", "t")
            np.testing.assert_array_equal(out.values, expect)
        # same labels in another dimension / label order: aligned to cube1 like xr.concat does
        perm = [2, 0, 3, 1]
        shuffled = mk(np.ascontiguousarray(full[..., perm].transpose(3, 0, 1, 2)), ("bands", "x", "y", "t"),
                      dict(coords, bands=coords["bands"][perm]))
        out = P.merge_cubes(mk(full, dims, coords), shuffled, functools.partial(P.process(P.subtract), x=ref_x, y=ref_y))
        np.testing.assert_array_equal(out.values, R.subtract(full, full))
        # example 4: (x, y) plane broadcast over the cube, both argument orders (test_merge_cubes_type_4)
        plane = _mk(rng, (6, 5), np.float64, 0.0)
        cp = {"x": coords["x"], "y": coords["y"]}
        with pytest.raises(P.exceptions.OverlapResolverMissing):
            P.merge_cubes(mk(full, dims, coords), mk(plane, ("x", "y"), cp))
        sub_r = functools.partial(P.process(P.subtract), x=ref_x, y=ref_y)
        m1 = P.merge_cubes(mk(full, dims, coords), mk(plane, ("x", "y"), cp), overlap_resolver=sub_r)
        m2 = P.merge_cubes(mk(plane, ("x", "y"), cp), mk(full, dims, coords), overlap_resolver=sub_r)
        assert m1.dims == m2.dims == dims
        np.testing.assert_array_equal(m1.values, full - plane[:, :, None, None])
        np.testing.assert_array_equal(m2.values, plane[:, :, None, None] - full)
        want, _, _ = OC.merge_cubes((plane, ("x", "y"), cp), (full, dims, coords), R.subtract)
        np.testing.assert_array_equal(m2.values, want)
        # a (y, x)-ordered plane: transposed view, still no copy of the cube
        m3 = P.merge_cubes(mk(full, dims, coords), mk(np.ascontiguousarray(plane.T), ("y", "x"), cp), overlap_resolver=sub_r)
        np.testing.assert_array_equal(m3.values, m1.values)
        # float coordinates that differ by < 1e-6 are aligned to cube1's (test_merge_float_coord_alignment)
        near = dict(coords, x=coords["x"] + 1e-9)
        out = P.merge_cubes(mk(full, dims, coords), mk(full, dims, near), add)
        np.testing.assert_array_equal(out.values, full * 2)
        np.testing.assert_array_equal(out["x"].values, coords["x"])
        with pytest.raises(ValueError):
            P.merge_cubes(mk(full, dims, coords), mk(plane[0], ("x",), {"x": coords["x"]}), add)


def test_unary_processes_equal_reference_outputs(P, torch, golden_binary):
    """absolute / sgn / sqrt / floor / ceil / int through eoc_elementwise: values, NaN masks, dtypes and the sign of
    zeros equal the outputs of the reference's own functions (float32 / float64 cubes, special values incl.
    subnormals, 16- and 32-bit integers)."""
    g = golden_binary
    names = {"absolute": "absolute", "sgn": "sgn", "sqrt": "sqrt", "floor": "floor", "ceil": "ceil", "int": "_int"}
    for c in sorted(k.split("/")[1] for k in g.files if k.startswith("uin/")):
        x = g[f"uin/{c}"]
        for name, pname in names.items():
            want = g[f"uout/{c}/{name}"]
            for arg in (x.copy(), torch.from_numpy(x.copy()).cuda()):
                got = getattr(P, pname)(arg)
                got = got.cpu().numpy() if hasattr(got, "cpu") else np.asarray(got)
                assert got.dtype == want.dtype and got.shape == want.shape, (c, name, got.dtype, want.dtype)
                np.testing.assert_array_equal(got, want, err_msg=f"{c}/{name}")
                if got.dtype.kind == "f":
                    np.testing.assert_array_equal(np.signbit(got) & ~np.isnan(got), np.signbit(want) & ~np.isnan(want),
                                                  err_msg=f"{c}/{name} (sign of zero)")


@pytest.mark.parametrize("T", [130, 255, 300, 511])
def test_wide_selection_kernel_layouts_and_sort(P, torch, T):
    """k_select_wide (128 < R <= 512, K threads per column): (bands,t,y,x) layout (outer > 1, strided column),
    reduction over the middle axis, odd widths, sort along t in both orders, real infinities next to the NaN pads,
    and agreement with the shared-memory sorter it replaces (EOC_SELECT_FORCE_SMEM is read per call)."""
    rng = np.random.default_rng(900 + T)
    x = _mk(rng, (3, T, 5, 41), np.float32, 0.25)          # (bands, t, y, x): the load_stac layout
    x[:, :, 0, 0] = np.nan                                  # all-NaN columns
    x[0, : T // 2, 1, 1] = np.inf
    x[0, T // 2:, 1, 1] = -np.inf
    x[1, :, 2, 2] = 7.0                                     # all ties
    tx = torch.from_numpy(x).cuda()
    for ign in (True, False):
        _same(P.median(tx, ignore_nodata=ign, axis=1), R.median(x, ignore_nodata=ign, axis=1), True)
    got = P.quantiles(tx, probabilities=[0.0, 0.25, 0.5, 1.0], axis=1).cpu().numpy()
    want = R.quantiles(x.copy(), probabilities=[0.0, 0.25, 0.5, 1.0], axis=1)
    np.testing.assert_allclose(got, want, rtol=RTOL, equal_nan=True)
    view = tx[:, :, 1:4, 3:36]                               # a window of the cube: strides of the parent
    _same(P.median(view, axis=1), R.median(x[:, :, 1:4, 3:36], axis=1), True)
    moved = np.ascontiguousarray(np.moveaxis(x, 1, -1))
    for asc in (True, False):
        for nd in (True, False):
            want = np.moveaxis(np.asarray(R.sort(moved.copy(), asc=asc, nodata=nd, axis=moved.ndim - 1)), -1, 1)
            got = P.sort(tx, asc=asc, nodata=nd, axis=1).cpu().numpy()
            np.testing.assert_array_equal(got, want, err_msg=f"sort asc={asc} nodata={nd}")
    os.environ["EOC_SELECT_FORCE_SMEM"] = "1"
    try:
        old = P.median(tx, axis=1).cpu().numpy()
    finally:
        del os.environ["EOC_SELECT_FORCE_SMEM"]
    np.testing.assert_array_equal(P.median(tx, axis=1).cpu().numpy(), old)


@pytest.mark.parametrize("kind", ["numpy", "cuda"])
@pytest.mark.parametrize("dtype,hi", [(np.uint16, 65536), (np.int16, 32768), (np.uint8, 256)])
@pytest.mark.parametrize("T", [5, 24, 64, 100, 128, 200, 365])
def test_integer_cubes_on_the_float32_networks(P, torch, kind, dtype, hi, T):
    """Raw digital numbers (uint16 / int16 / uint8): median and quantiles run the float32 register networks with a
    conversion on load and return float64 like np.nanmedian / np.nanquantile of an integer array -- bit for bit
    (every sample, the mean of the two middle ones and the float64 interpolation are exact)."""
    rng = np.random.default_rng(T * 7 + hi)
    lo = -hi if np.dtype(dtype).kind == "i" else 0
    x = rng.integers(lo, hi, size=(T, 2, 7, 45)).astype(dtype)
    x[:, 0, 0, :8] = rng.integers(0, 3, size=(T, 8)).astype(dtype)       # heavy ties
    x[:, 1, 1, 0] = np.iinfo(dtype).max
    x[:, 1, 1, 1] = np.iinfo(dtype).min
    arg = torch.from_numpy(x).cuda() if kind == "cuda" else x
    for axis in ((0, 1) if T <= 24 else (0,)):
        got = P.median(arg, axis=axis)
        got = got.cpu().numpy() if hasattr(got, "cpu") else got
        want = R.median(x, axis=axis)
        assert got.dtype == want.dtype == np.float64
        np.testing.assert_array_equal(got, want)
        probs = [0.0, 0.1, 0.5, 0.77, 1.0]
        gq = P.quantiles(arg, probabilities=probs, axis=axis)
        gq = gq.cpu().numpy() if hasattr(gq, "cpu") else gq
        wq = R.quantiles(x, probabilities=probs, axis=axis)
        assert gq.dtype == np.float64 and gq.shape == wq.shape
        np.testing.assert_array_equal(gq, wq)
    # a band-sliced view (strided columns) and ignore_nodata = False (integers hold no NaN: same result)
    if kind == "cuda":
        np.testing.assert_array_equal(P.median(arg[:, 1], ignore_nodata=False, axis=0).cpu().numpy(), R.median(x[:, 1], axis=0))


@pytest.mark.parametrize("pinned", [False, True])
def test_host_pipeline_many_strips(P, torch, pinned, monkeypatch):
    from openeo_processes_dask_b200 import _lib as L_

    """The host strip pipeline with many small strips (EOC_STRIP_BYTES): pageable input goes through the pinned
    staging strips gathered by host threads, pinned input through cudaMemcpy2DAsync; strips along inner, chunks of
    outer slices, deferred result copies -- all bit-exact vs the oracle."""
    monkeypatch.setenv("EOC_STRIP_BYTES", str(1 << 20))
    rng = np.random.default_rng(404)
    for shape, axis in (((40, 3, 50, 300), 0), ((6, 40, 20, 1500), 1), ((3, 17, 9000), 1)):
        x = _mk(rng, shape, np.float32, 0.25)
        if pinned:
            t = torch.empty(shape, dtype=torch.float32, pin_memory=True)
            t.numpy()[...] = x
            arg = t.numpy()
        else:
            arg = x
        for name in ("mean", "sd", "max"):
            _same(_fn(P, name)(arg, axis=axis), R.BY_NAME[name](x, axis=axis), True)
        _same(P.median(arg, axis=axis), R.median(x, axis=axis), True)
        np.testing.assert_array_equal(P.count(arg, axis=axis), R.count(x, axis=axis))
        got = P.quantiles(arg, probabilities=[0.25, 0.5], axis=axis)
        np.testing.assert_allclose(got, R.quantiles(x.copy(), probabilities=[0.25, 0.5], axis=axis), rtol=RTOL, equal_nan=True)
    u = rng.integers(0, 10000, size=(30, 2, 40, 700)).astype(np.uint16)     # integer cube, pageable
    np.testing.assert_array_equal(P.median(u, axis=0), R.median(u, axis=0))
    if pinned:   # a source pitch beyond cudaMemcpy2D's limit (forced): the strips are copied row by row
        monkeypatch.setenv("EOC_MAX_PITCH", "1024")
        x = _mk(rng, (40, 3, 50, 300), np.float32, 0.25)
        t = torch.empty(x.shape, dtype=torch.float32, pin_memory=True)
        t.numpy()[...] = x
        _same(P.median(t.numpy(), axis=0), R.median(x, axis=0), True)
        _same(P.mean(t.numpy(), axis=1), R.mean(x, axis=1), True)
        monkeypatch.delenv("EOC_MAX_PITCH")
    # round 2: the other array processes on the same two-lane pipeline (eoc_segmented_reduce_host, eoc_nd_reduce_host,
    # eoc_scan_host, eoc_sort_host): numpy in -> numpy out, many strips, pinned and pageable
    from openeo_processes_dask_b200 import engine

    def host(a):
        if not pinned:
            return a
        t = torch.empty(a.shape, dtype=getattr(torch, str(a.dtype)), pin_memory=True)
        t.numpy()[...] = a
        return t.numpy()

    x = _mk(rng, (60, 3, 40, 700), np.float32, 0.3)
    arg = host(x)
    np.testing.assert_array_equal(P.cumsum(arg, axis=0), R.cumsum(x, axis=0))
    np.testing.assert_array_equal(P.cummax(arg, ignore_nodata=False, axis=0), R.cummax(x, ignore_nodata=False, axis=0))
    y = np.ascontiguousarray(np.moveaxis(x, 0, -1))
    np.testing.assert_array_equal(P.sort(host(y), nodata=True, axis=3), R.sort(y, nodata=True, axis=3))
    # round 2, late: gap filling, elementwise processes and recorded chains of numpy chunks on the same pipeline
    np.testing.assert_array_equal(P.array_interpolate_linear(host(x.copy()), axis=0), R.array_interpolate_linear(x, axis=0))
    days = np.sort(rng.choice(np.arange(400), size=60, replace=False)).astype(float)
    np.testing.assert_array_equal(P.array_interpolate_linear(host(x.copy()), axis=0, dim_labels=list(days)),
                                  R.array_interpolate_linear(x, axis=0, positions=days))
    np.testing.assert_array_equal(P.clip(host(x), min=-500, max=900), np.clip(x, -500, 900))
    np.testing.assert_array_equal(P.linear_scale_range(host(x), inputMin=-1000, inputMax=4000, outputMin=0, outputMax=255),
                                  R.linear_scale_range(x, -1000, 4000, 0, 255))
    with np.errstate(all="ignore"):
        np.testing.assert_allclose(np.asarray(P.exp(host(x / 1000))), np.exp(x / 1000), rtol=1e-6, equal_nan=True)
        np.testing.assert_array_equal(np.asarray(P.absolute(host(x))), np.abs(x))
    np.testing.assert_array_equal(P.sort(arg, asc=False, nodata=True, axis=0), -np.sort(-x, axis=0))
    groups = [list(range(0, 20)), [], list(range(25, 60, 2))]
    res = engine.segmented_reduce(arg, 0, groups, [L_.OP_MEAN, L_.OP_COUNT, L_.OP_SD])
    for k, g in enumerate(groups):
        if g:
            np.testing.assert_array_equal(res[L_.OP_MEAN][k], R.mean(x[g], axis=0))
            np.testing.assert_array_equal(res[L_.OP_COUNT][k], R.count(x[g], axis=0))
            np.testing.assert_array_equal(res[L_.OP_SD][k], R.sd(x[g], axis=0))
        else:
            assert np.isnan(res[L_.OP_MEAN][k]).all() and (res[L_.OP_COUNT][k] == 0).all()
    a16, b16 = u[:, 0], u[:, 1]
    with np.errstate(all="ignore"):
        np.testing.assert_array_equal(P.normalized_difference(host(a16), host(b16)), R.normalized_difference(a16, b16))
        nd = R.normalized_difference(x[:, 0], x[:, 1])
        np.testing.assert_array_equal(P.normalized_difference(host(x[:, 0].copy()), host(x[:, 1].copy())), nd)
        outs, _ = engine.nd_reduce(host(x[:, 0].copy()), host(x[:, 1].copy()), 0, [L_.OP_MAX, L_.OP_MEAN])
        np.testing.assert_array_equal(outs[L_.OP_MAX], R._max(nd, axis=0))
        np.testing.assert_array_equal(outs[L_.OP_MEAN], R.mean(nd, axis=0))
    assert L_.lib().eoc_host_release() == 0     # frees this thread's staging; the next call re-creates it
    np.testing.assert_array_equal(P.cumsum(arg, axis=0), R.cumsum(x, axis=0))


# ---- round 2: no host arithmetic in normalized_difference; fused paths are the default ------------------------
def test_normalized_difference_every_operand_kind_runs_on_the_device(P, torch, golden):
    """math.py:412-414 for integer / broadcast / scalar / mixed operands: eoc_nd_reduce(integer_wrap) or three
    eoc_binary launches -- never numpy.  Golden nd_uint16 (uint16 1000, 2000 -> 64536 / 3000 = 21.512) is an output
    of the reference's own function."""
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.device_array import DeviceArray

    got = P.normalized_difference(np.array([1000], dtype=np.uint16), np.array([2000], dtype=np.uint16))
    np.testing.assert_array_equal(got, golden["out/test_math/nd_uint16"])
    assert got.dtype == np.float64 and abs(float(got[0]) - 21.512) < 1e-12
    rng = np.random.default_rng(77)
    cases = {
        "u16": (rng.integers(0, 65536, (5, 6, 33)).astype(np.uint16), rng.integers(0, 65536, (5, 6, 33)).astype(np.uint16)),
        "i16": (rng.integers(-32768, 32768, (5, 6, 33)).astype(np.int16), rng.integers(-32768, 32768, (5, 6, 33)).astype(np.int16)),
        "u8": (rng.integers(0, 256, (7, 40)).astype(np.uint8), rng.integers(0, 256, (7, 40)).astype(np.uint8)),
        "i32": (rng.integers(-2**31, 2**31 - 1, (4, 33)).astype(np.int32), rng.integers(-2**31, 2**31 - 1, (4, 33)).astype(np.int32)),
        "f32_plane": (_mk(rng, (6, 3, 5, 9), np.float32, 0.2), _mk(rng, (5, 9), np.float32, 0.2)),
        "f32_scalar": (_mk(rng, (6, 3, 5, 9), np.float32, 0.2), 1000),
        "scalar_f64": (2.5, _mk(rng, (4, 7), np.float64, 0.2)),
        "mixed": (_mk(rng, (4, 7), np.float64, 0.2), _mk(rng, (4, 7), np.float32, 0.2)),
        "u16_scalar": (rng.integers(0, 65536, (5, 33)).astype(np.uint16), 3),
        "zeros": (np.zeros((3, 8), np.uint16), np.zeros((3, 8), np.uint16)),
    }
    for name, (x, y) in cases.items():
        with np.errstate(all="ignore"):
            want = np.asarray(R.normalized_difference(x, y))
        for kind in ("numpy", "device"):
            wrap = (lambda a: a) if kind == "numpy" else (lambda a: DeviceArray.from_numpy(a) if isinstance(a, np.ndarray) else a)
            engine.launch_count(reset=True)
            out = P.normalized_difference(wrap(x), wrap(y))
            assert engine.launch_count() >= 1, f"{name}/{kind}: no kernel ran"
            res = np.asarray(out)
            assert res.dtype == want.dtype and res.shape == want.shape, (name, kind, res.dtype, want.dtype)
            np.testing.assert_array_equal(res, want, err_msg=f"{name}/{kind}")
    assert float(np.asarray(P.normalized_difference(5, 3))) == 0.25     # tests/test_math.py:78-87
    assert float(np.asarray(P.normalized_difference(1, 1))) == 0.0


@pytest.mark.parametrize("dtype", [np.float32, np.uint16, np.int16, np.uint8, np.float64])
def test_ndvi_fused_is_the_default_and_equals_unfused(P, torch, dtype):
    """ndvi -> reduce_dimension as an unmodified graph runs it: the lazy result feeds eoc_nd_reduce.  Fused ==
    materialise-then-reduce bit for bit, and both equal the reference's numpy arithmetic: integer cubes WRAP
    (float64 result); the same cube cast with astype("float32") is converted exactly in the load stage."""
    from openeo_processes_dask_b200.device_array import DeviceArray, LazyCast, LazyNormalizedDifference

    rng = np.random.default_rng(12)
    if np.dtype(dtype).kind == "f":
        u = _mk(rng, (24, 4, 6, 40), dtype, 0.2)
    else:
        info = np.iinfo(dtype)
        u = rng.integers(info.min, int(info.max) + 1, (24, 4, 6, 40)).astype(dtype)
    bands = ["b1", "red", "b3", "nir"]
    ref = P.ParameterReference(from_parameter="data")
    variants = [("raw", lambda c: c, u)]
    if np.dtype(dtype) in LazyCast.CASTABLE:
        variants.append(("cast", lambda c: c.astype("float32"), u.astype(np.float32)))
    for vname, prep, host in variants:
        cube = prep(P.DataCube(DeviceArray.from_numpy(u), ("t", "bands", "y", "x"), {"bands": bands}))
        if vname == "cast":
            assert isinstance(cube.data, LazyCast)
        with np.errstate(all="ignore"):
            nd = np.asarray(R.normalized_difference(host[:, 3], host[:, 1]))
        lazy = P.ndvi(cube)                     # default: fused
        eager = P.ndvi(cube, fuse=False)
        assert isinstance(lazy.data, LazyNormalizedDifference) and isinstance(lazy.data, DeviceArray)
        assert not isinstance(eager.data, LazyNormalizedDifference)
        assert lazy.data.dtype == nd.dtype == eager.data.dtype
        np.testing.assert_array_equal(np.asarray(eager.data), nd)
        for name in ("max", "min", "mean", "sum", "sd", "median"):
            for ign in (True, False):
                red = functools.partial(P.process(_fn(P, name)), data=ref, ignore_nodata=ign)
                a = P.reduce_dimension(data=lazy, reducer=red, dimension="t")
                b = P.reduce_dimension(data=eager, reducer=red, dimension="t")
                np.testing.assert_array_equal(np.asarray(a.data), np.asarray(b.data), err_msg=f"{vname}/{name}/{ign}")
                with np.errstate(all="ignore"):
                    want = R.BY_NAME[name](nd, ignore_nodata=ign, axis=0)
                _same(a.data.tensor, want, True)
        np.testing.assert_array_equal(np.asarray(lazy.data), nd)   # touching it materialises with the same kernel


@pytest.mark.parametrize("replacement", [0, -1.5, None])
def test_mask_default_is_fused_and_honours_the_replacement(P, torch, replacement):
    """mask(replacement=0) -> median / quantiles must treat replaced samples as 0, not NaN.  The
    fused selection kernel only covers replacement = NaN; anything else materialises first -- same answer as the
    oracle's data.where(_not(mask), replacement) followed by the reducer, for ignore_nodata True and False."""
    from openeo_processes_dask_b200.device_array import DeviceArray, LazyMasked

    rng = np.random.default_rng(19)
    dims = ("t", "bands", "y", "x")
    x = _mk(rng, (100, 2, 6, 64), np.float32, 0.15)
    m = rng.random((100, 6, 64)) < 0.3
    masked, _ = OC.mask(x, dims, m, ("t", "y", "x"), replacement)
    cube = P.DataCube(DeviceArray.from_numpy(x), dims, {"t": np.arange(100)})
    mcube = P.DataCube(DeviceArray.from_numpy(m), ("t", "y", "x"), {"t": np.arange(100)})
    lazy = P.mask(data=cube, mask=mcube, replacement=replacement)
    assert isinstance(lazy.data, LazyMasked) and isinstance(lazy.data, DeviceArray)
    ref = P.ParameterReference(from_parameter="data")
    for ign in (True, False):
        for name in ("median", "mean", "max", "sd"):
            red = functools.partial(P.process(_fn(P, name)), data=ref, ignore_nodata=ign)
            out = P.reduce_dimension(data=lazy, reducer=red, dimension="t")
            _same(out.data.tensor, R.BY_NAME[name](masked, ignore_nodata=ign, axis=0), True)
        q = P.quantiles(lazy.data, probabilities=[0.1, 0.5, 0.9], ignore_nodata=ign, axis=0)
        want = R.quantiles(masked, probabilities=[0.1, 0.5, 0.9], ignore_nodata=ign, axis=0)
        np.testing.assert_allclose(np.asarray(q), want, rtol=RTOL, equal_nan=True)
    # a consumer that is not a reducer sees the materialised cube
    nd = P.ndvi(P.DataCube(lazy.data, dims, {"bands": ["red", "nir"], "t": np.arange(100)}))
    with np.errstate(all="ignore"):
        np.testing.assert_array_equal(np.asarray(nd.data), R.normalized_difference(masked[:, 1], masked[:, 0]))


# ---- round 2: the bench configurations at their own geometry ("What's weak" 1) ----------------------
def _windows(rng, rows, x_len, n, width=128):
    for _ in range(n):
        yield int(rng.integers(0, rows)), int(rng.integers(0, x_len - width))


def test_bench_geometry_quantiles_and_fused_sd_at_full_width(P, torch):
    """C3 geometry: (100, 12 bands, rows, 10980): x = 10980 leaves a ragged last tile (10980 mod 128 = 100, mod 256 =
    228), band planes 12 apart.  Quantile VALUES (not only their ordering) and the fused quantiles + sd pass through
    the DEFAULT kernel choice against the oracle on windows regenerated from the hash, including the last columns."""
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.process_implementations.math import reduce_many

    T_, B_, Y_, X_ = 100, 12, 6, 10980
    x = torch.empty((T_, B_, Y_, X_), dtype=torch.float32, device="cuda")
    engine.synth_fill(x.view(T_ * B_, Y_ * X_), T_ * B_, 10980 * X_, 0, Y_ * X_, 44, 0.30)
    probs = [0.1, 0.5, 0.9]
    q = P.quantiles(x, probabilities=probs, axis=0)
    fq, fsd = reduce_many(x, [("quantiles", {"probabilities": probs}), "sd"], axis=0)
    assert torch.equal(torch.nan_to_num(q, nan=-7.0), torch.nan_to_num(fq, nan=-7.0))
    sd = P.sd(x, axis=0)
    assert torch.equal(torch.nan_to_num(sd, nan=-7.0), torch.nan_to_num(fsd, nan=-7.0))
    rng = np.random.default_rng(3)
    wins = list(_windows(rng, Y_, X_, 20)) + [(Y_ - 1, X_ - 128), (0, 0)]
    for y, x0 in wins:
        for b in (0, 7, 11):
            w = synth.synth_window(44, (T_, B_, 10980, X_), (slice(0, T_), slice(b, b + 1), slice(y, y + 1),
                                                              slice(x0, x0 + 128)), np.float32, 0.30)
            want = R.quantiles(w, probabilities=probs, axis=0)[:, 0, 0]
            np.testing.assert_allclose(q[:, b, y, x0:x0 + 128].cpu().numpy(), want, rtol=RTOL, equal_nan=True)
            np.testing.assert_array_equal(fsd[b, y, x0:x0 + 128].cpu().numpy(), R.sd(w, axis=0)[0, 0])


def test_bench_geometry_ndvi_max_uint16_12_bands(P, torch):
    """C4 geometry: (200, 12, rows, 10980) uint16, nir = band 8, red = band 4 read with a 12-plane stride; the cube
    is cast with astype(float32) (LazyCast) and ndvi -> reduce_dimension(max) runs fused by default."""
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.device_array import DeviceArray

    T_, B_, Y_, X_ = 200, 12, 4, 10980
    u = torch.empty((T_, B_, Y_, X_), dtype=torch.uint16, device="cuda")
    engine.synth_fill(u.view(T_ * B_, Y_ * X_), T_ * B_, 10980 * X_, 0, Y_ * X_, 45, 0.0)
    names = [f"B{k:02d}" for k in range(1, 13)]
    cube = P.DataCube(DeviceArray(u), ("t", "bands", "y", "x"), {"bands": names}).astype("float32")
    ref = P.ParameterReference(from_parameter="data")
    nd = P.ndvi(cube, nir="B08", red="B04")
    for name in ("max", "mean"):
        out = P.reduce_dimension(data=nd, reducer=functools.partial(P.process(_fn(P, name)), data=ref), dimension="t")
        got = np.asarray(out.data)
        rng = np.random.default_rng(4)
        for y, x0 in list(_windows(rng, Y_, X_, 12)) + [(Y_ - 1, X_ - 128)]:
            w = synth.synth_window(45, (T_, B_, 10980, X_), (slice(0, T_), slice(0, B_), slice(y, y + 1),
                                                              slice(x0, x0 + 128)), np.uint16, 0.0).astype(np.float32)
            with np.errstate(all="ignore"):
                want = R.BY_NAME[name](R.normalized_difference(w[:, 7], w[:, 3]), axis=0)[0]
            np.testing.assert_array_equal(got[y, x0:x0 + 128], want, err_msg=name)


def test_bench_geometry_aggregate_temporal_x5490(P, torch):
    """C5 geometry: (365, 4, rows, 5490) float32, 40 % NaN, 12 monthly bins.  A row is 21960 bytes -- not a
    multiple of 16 -- so whole-row vector loads are only possible on every other row (SURVEY 7 names the hazard)."""
    from openeo_processes_dask_b200 import engine

    T_, B_, Y_, X_ = 365, 4, 5, 5490
    c = torch.empty((T_, B_, Y_, X_), dtype=torch.float32, device="cuda")
    engine.synth_fill(c.view(T_ * B_, Y_ * X_), T_ * B_, 5490 * X_, 0, Y_ * X_, 46, 0.40)
    days = pd.date_range("2021-01-01T12:00:00", periods=T_, freq="D").values
    months = pd.date_range("2021-01-01", "2022-01-01", freq="MS").strftime("%Y-%m-%dT%H:%M:%S")
    intervals = [[a, b] for a, b in zip(months[:-1], months[1:])]
    ref = P.ParameterReference(from_parameter="data")
    # the second cube is one band x one row: its t stride is 21960 bytes, so NO row after the first is 16-byte aligned
    for sub in (c, c[:, :1, :1, :].contiguous()):
        cube = P.DataCube(sub, ("t", "bands", "y", "x"), {"t": days})
        host = sub.cpu().numpy()
        for name, ofn in (("mean", R.mean), ("count", R.count), ("sd", R.sd), ("median", R.median)):
            out = P.aggregate_temporal(data=cube, intervals=intervals,
                                       reducer=functools.partial(P.process(_fn(P, name)), data=ref))
            want, _, _ = OC.aggregate_temporal(host[:, :, :, -700:], cube.dims, {"t": days}, intervals, ofn)
            _same(out.data[:, :, :, -700:], want, True)


@pytest.mark.parametrize("R_", [129, 160, 161, 177, 200, 201, 256, 257, 300, 301, 320, 321, 365, 384, 385, 400, 401, 512])
@pytest.mark.parametrize("width", ["64", "128", "multi"])
def test_split_selection_long_columns(P, torch, monkeypatch, R_, width):
    """128 < R <= 512 (forced on for a small cube): k_select_multi (eight sentinel runs, several threads per column:
    the default) and k_select_split with K = 2 .. 4 runs (both tile widths): median, quantiles and
    ignore_nodata = False against the oracle, bit for bit; ragged last tile; every run length."""
    from openeo_processes_dask_b200 import engine

    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    monkeypatch.setenv("EOC_SELECT_MULTI", "1" if width == "multi" else "0")
    if width != "multi":
        monkeypatch.setenv("EOC_SPLIT_WIDE_W", width)
    rng = np.random.default_rng(R_)
    x = _mk(rng, (R_, 2, 3, 100), np.float32, 0.3)
    x[:, 0, 0, 1] = np.nan
    x[5:, 0, 0, 2] = np.nan
    x[: R_ - 3, 0, 0, 3] = np.nan
    x[:, 1, 2, 96:] = np.round(x[:, 1, 2, 96:] / 700)
    d = torch.from_numpy(x).cuda()
    engine.launch_count(reset=True)
    for ign in (True, False):
        _same(P.median(d, ignore_nodata=ign, axis=0), R.median(x, ignore_nodata=ign, axis=0), True)
    assert engine.launch_count() == 2
    probs = [0.1, 0.5, 0.9, 1.0]
    x[:, 0, 0, 0] = 1.0
    d = torch.from_numpy(x).cuda()
    q = P.quantiles(d, probabilities=probs, axis=0)
    np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(x, probabilities=probs, axis=0))
    y = np.ascontiguousarray(np.moveaxis(x, 0, 1))          # (b, t, y, x): outer > 1
    _same(P.median(torch.from_numpy(y).cuda(), axis=1), R.median(y, axis=1), True)


@pytest.mark.parametrize("R_", [17, 24, 25, 32, 33, 40, 48, 49, 52, 53, 56, 57, 64, 65, 72, 80, 81, 96, 99, 100, 101, 112, 127, 128])
@pytest.mark.parametrize("width", ["64", "128"])
def test_quad_selection_kernel(P, torch, monkeypatch, R_, width):
    """k_select_quad (the plain median of 64 < R <= 128 samples as four raw-sorted runs with NaN sentinels and the
    PTX rank elimination), forced on for a small cube: == oracle bit for bit, == the round-1 median cone, both tile
    widths, ragged last tile, outer > 1, NaN payloads incl. signalling and negative NaNs, infinities, ties,
    all-NaN / one-sample columns, both nodata modes."""
    from openeo_processes_dask_b200 import engine

    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    monkeypatch.setenv("EOC_QUAD_W", width)
    monkeypatch.setenv("EOC_QUAD_SHORT", "1")   # 17 .. 32 samples: off by default (the register kernel is ahead)
    rng = np.random.default_rng(R_ * 7 + int(width))
    x = _mk(rng, (R_, 2, 3, 300), np.float32, 0.3)
    x[:, 0, 0, 1] = np.nan
    x[1:, 0, 0, 2] = np.nan
    x[: R_ - 3, 0, 0, 3] = np.nan
    x[: R_ // 2, 0, 0, 4] = np.nan
    x[::3, 0, 1, :8] = np.inf
    x[1::3, 0, 1, :8] = -np.inf
    x[:, 1, 2, 200:] = np.where(np.isnan(x[:, 1, 2, 200:]), np.nan, np.round(x[:, 1, 2, 200:] / 700))
    bits = x.view(np.uint32)
    nanmask = np.isnan(x)
    bits[nanmask] = rng.choice(np.array([0x7FC00000, 0xFFC00000, 0x7FC12345, 0xFFFFFFFF, 0x7F800001], dtype=np.uint32),
                               size=int(nanmask.sum()))
    d = torch.from_numpy(x).cuda()
    y = np.ascontiguousarray(np.moveaxis(x, 0, 1))          # (b, t, y, x): outer > 1
    dy = torch.from_numpy(y).cuda()
    engine.launch_count(reset=True)
    for ign in (True, False):
        monkeypatch.setenv("EOC_SELECT_QUAD", "1")
        got = P.median(d, ignore_nodata=ign, axis=0)
        _same(got, R.median(x, ignore_nodata=ign, axis=0), True)
        _same(P.median(dy, ignore_nodata=ign, axis=1), R.median(y, ignore_nodata=ign, axis=1), True)
        monkeypatch.setenv("EOC_SELECT_QUAD", "0")
        cone = P.median(d, ignore_nodata=ign, axis=0)
        assert bool(((got.view(torch.int32) == cone.view(torch.int32)) | (torch.isnan(got) & torch.isnan(cone))).all())
    assert engine.launch_count() == 6


@pytest.mark.parametrize("R_", [33, 48, 52, 56, 64, 65, 80, 81, 99, 100, 101, 127, 128])
def test_quad_selection_quantiles_fused_masked(P, torch, monkeypatch, R_):
    """The other uses of k_select_quad, forced on for a small cube, against the oracle: quantiles (ascending and
    unordered probabilities: the rank cursor moves on or starts over), quantiles + sd / median + mean + max + count out
    of one read (eoc_select_fused), mask -> median / quantiles (eoc_masked_select)."""
    from openeo_processes_dask_b200 import _lib as L
    from openeo_processes_dask_b200 import engine

    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    rng = np.random.default_rng(R_ * 13)
    x = _mk(rng, (R_, 2, 3, 300), np.float32, 0.3)
    x[:, 0, 0, 1] = np.nan
    x[1:, 0, 0, 2] = np.nan
    x[: R_ - 3, 0, 0, 3] = np.nan
    x[:, 0, 0, 0] = 1.0     # (numpy sizes the quantile buffer's dtype from the first slices)
    x[::3, 0, 1, :8] = np.inf
    x[:, 1, 2, 200:] = np.where(np.isnan(x[:, 1, 2, 200:]), np.nan, np.round(x[:, 1, 2, 200:] / 700))
    d = torch.from_numpy(x).cuda()
    for probs in ([0.1, 0.5, 0.9], [0.9, 0.25, 0.5, 1.0], [0.0, 1 / 3, 0.999]):
        for ign in (True, False):
            q = P.quantiles(d, probabilities=probs, ignore_nodata=ign, axis=0)
            np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(x, probabilities=probs, ignore_nodata=ign, axis=0))
    for ign in (True, False):
        ops = [L.OP_SD, L.OP_MEAN, L.OP_MAX, L.OP_COUNT]
        med, qs, outs = engine.select(d, 0, ign, True, [0.1, 0.5, 0.9], ops=ops)
        _same(med, R.median(x, ignore_nodata=ign, axis=0), True)
        np.testing.assert_array_equal(qs.cpu().numpy(), R.quantiles(x, probabilities=[0.1, 0.5, 0.9], ignore_nodata=ign, axis=0))
        _same(outs[L.OP_SD], R.sd(x, ignore_nodata=ign, axis=0), True)
        _same(outs[L.OP_MEAN], R.mean(x, ignore_nodata=ign, axis=0), True)
        _same(outs[L.OP_MAX], R._max(x, ignore_nodata=ign, axis=0), True)
        _same(outs[L.OP_COUNT], R.count(x, axis=0), True)
        _, qs1, outs1 = engine.select(d, 0, ign, False, [0.5], ops=[L.OP_SD])   # fused level 1
        _same(outs1[L.OP_SD], R.sd(x, ignore_nodata=ign, axis=0), True)
    dims = ("t", "bands", "y", "x")
    m = rng.random((R_, 3, 300)) < 0.35
    masked, _ = OC.mask(x, dims, m, ("t", "y", "x"), None)
    cube = P.DataCube(d, dims, {"t": np.arange(R_)})
    mcube = P.DataCube(torch.from_numpy(m).cuda(), ("t", "y", "x"), {"t": np.arange(R_)})
    lazy = P.mask(data=cube, mask=mcube)
    for ign in (True, False):
        _same(P.median(lazy.data, ignore_nodata=ign, axis=0), R.median(masked, ignore_nodata=ign, axis=0), True)
    q = P.quantiles(lazy.data, probabilities=[0.1, 0.5, 0.9], axis=0)
    np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(masked, probabilities=[0.1, 0.5, 0.9], axis=0))


@pytest.mark.parametrize("R_", [33, 48, 52, 56, 64, 65, 80, 81, 100, 101, 128, 129, 160, 200, 256, 300, 365, 400, 512])
@pytest.mark.parametrize("dtype", [np.uint16, np.int16, np.uint8])
def test_quad_selection_integer_cubes(P, torch, monkeypatch, R_, dtype):
    """k_select_quad (R <= 128) / k_select_multi (R <= 512) on raw digital numbers (TMA box of uint16 / int16 / uint8 samples landing inside the run's float
    rows, exact conversion in the register fill, float64 results), forced on for a small cube: median and quantiles ==
    numpy bit for bit, == the register kernels; ragged last tile; outer > 1; int16 quantile wrap-around."""
    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    rng = np.random.default_rng(R_ * 3 + np.dtype(dtype).itemsize)
    info = np.iinfo(dtype)
    x = rng.integers(info.min, int(info.max) + 1, size=(R_, 2, 3, 336), dtype=np.int64).astype(dtype)
    x[:, 0, 0, :16] = info.max
    x[: R_ // 2, 0, 1, :16] = info.min      # int16: neighbours 65535 apart around the middle
    x[R_ // 2:, 0, 1, :16] = info.max
    d = torch.from_numpy(x).cuda()
    want_m = R.median(x, axis=0)
    want_q = R.quantiles(x, probabilities=[0.1, 0.5, 0.9, 1.0], axis=0)
    for flag in ("1", "0"):
        monkeypatch.setenv("EOC_QUAD_INT", flag)
        got = P.median(d, axis=0)
        assert got.dtype == torch.float64
        _same(got, want_m, True)
        q = P.quantiles(d, probabilities=[0.1, 0.5, 0.9, 1.0], axis=0)
        np.testing.assert_array_equal(q.cpu().numpy(), want_q)
    monkeypatch.setenv("EOC_QUAD_INT", "1")
    y = np.ascontiguousarray(np.moveaxis(x, 0, 1))          # (b, t, y, x): outer > 1
    _same(P.median(torch.from_numpy(y).cuda(), axis=1), R.median(y, axis=1), True)


@pytest.mark.parametrize("R_", [33, 52, 64, 80, 100, 128])
@pytest.mark.parametrize("dtype", [np.uint16, np.int16, np.uint8])
def test_quad_selection_masked_integer_cubes(P, torch, monkeypatch, R_, dtype):
    """mask -> median / quantiles of RAW digital numbers in one pass (a cloud mask over a uint16 Sentinel-2 cube):
    k_select_quad<..., MASKED, In> == the oracle reducing the oracle's masked cube (float64, like numpy's where()),
    == the materialise-first fallback; bit for bit."""
    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    rng = np.random.default_rng(R_ * 5 + np.dtype(dtype).itemsize)
    info = np.iinfo(dtype)
    x = rng.integers(info.min, int(info.max) + 1, size=(R_, 2, 3, 336), dtype=np.int64).astype(dtype)
    dims = ("t", "bands", "y", "x")
    m = rng.random((R_, 3, 336)) < 0.35
    m[:, 0, 0] = True            # everything masked: NaN
    m[1:, 0, 1] = True           # one sample left
    masked, _ = OC.mask(x, dims, m, ("t", "y", "x"), None)
    cube = P.DataCube(torch.from_numpy(x).cuda(), dims, {"t": np.arange(R_)})
    mcube = P.DataCube(torch.from_numpy(m).cuda(), ("t", "y", "x"), {"t": np.arange(R_)})
    want_q = R.quantiles(masked, probabilities=[0.1, 0.5, 0.9], axis=0)
    for flag in ("1", "0"):
        monkeypatch.setenv("EOC_QUAD_INT", flag)
        lazy = P.mask(data=cube, mask=mcube)
        for ign in (True, False):
            got = P.median(lazy.data, ignore_nodata=ign, axis=0)
            _same(got, R.median(masked, ignore_nodata=ign, axis=0), True)
            assert np.asarray(got.cpu() if hasattr(got, "cpu") else got).dtype == np.float64
        q = P.quantiles(lazy.data, probabilities=[0.1, 0.5, 0.9], axis=0)
        np.testing.assert_array_equal(np.asarray(q.cpu() if hasattr(q, "cpu") else q), want_q)


@pytest.mark.parametrize("R_", [33, 48, 49, 64, 65, 80, 99, 100, 101, 128, 129, 160, 200, 256, 300, 365, 400, 512])
@pytest.mark.parametrize("g", ["1", "2", "4"])
def test_quad64_selection_kernel(P, torch, monkeypatch, R_, g):
    """k_select_quad64 (float64 columns: NaN counted and padded with +inf in the register fill, a one-DSETP comparator,
    four runs with sentinels, 64-bit rank elimination; 1 / 2 / 4 threads per column), forced on for a small cube: median
    and quantiles == numpy bit for bit in both nodata modes; real infinities next to the pads; ties; all-NaN and
    one-sample columns; ragged last tile; outer > 1."""
    monkeypatch.setenv("EOC_SELECT_TMA", "1")
    monkeypatch.setenv("EOC_QUAD64_G", g)
    monkeypatch.setenv("EOC_MULTI64_G", g)      # 129 .. 512 samples: k_select_multi64 (eight runs)
    rng = np.random.default_rng(R_ * 11 + int(g))
    x = _mk(rng, (R_, 2, 3, 150), np.float64, 0.3)
    x[:, 0, 0, 1] = np.nan
    x[1:, 0, 0, 2] = np.nan
    x[: R_ - 3, 0, 0, 3] = np.nan
    x[:, 0, 0, 0] = 1.0
    x[::3, 0, 1, :8] = np.inf
    x[1::3, 0, 1, :8] = -np.inf
    x[:, 1, 2, 100:] = np.where(np.isnan(x[:, 1, 2, 100:]), np.nan, np.round(x[:, 1, 2, 100:] / 700))
    d = torch.from_numpy(x).cuda()
    probs = [0.1, 0.5, 0.9, 1.0]
    for ign in (True, False):
        _same(P.median(d, ignore_nodata=ign, axis=0), R.median(x, ignore_nodata=ign, axis=0), True)
        q = P.quantiles(d, probabilities=probs, ignore_nodata=ign, axis=0)
        np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(x, probabilities=probs, ignore_nodata=ign, axis=0))
    y = np.ascontiguousarray(np.moveaxis(x, 0, 1))          # (b, t, y, x): outer > 1
    _same(P.median(torch.from_numpy(y).cuda(), axis=1), R.median(y, axis=1), True)
    monkeypatch.setenv("EOC_QUAD64", "0")                     # the earlier float64 kernels give the same
    _same(P.median(d, axis=0), R.median(x, axis=0), True)


def test_fixture_layout_with_short_inner_extent(P, torch):
    """The reference's test-fixture layout (x, y, t, bands): reducing t leaves inner = bands = 4.  Enough columns for
    the TMA-staged kernels' default gate, but their 128-column tiles would be 3 % full -- the dispatch must leave such
    views to the per-thread-load kernels (tma_tiles_fill); results against the oracle either way."""
    from openeo_processes_dask_b200 import engine

    rng = np.random.default_rng(2024)
    x = _mk(rng, (200, 200, 100, 4), np.float32, 0.3)
    d = torch.from_numpy(x).cuda()
    _same(P.median(d, axis=2), R.median(x, axis=2), True)
    q = P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=2)
    np.testing.assert_allclose(q.cpu().numpy(), R.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=2), rtol=RTOL, equal_nan=True)
    _same(P.sd(d, axis=2), R.sd(x, axis=2), True)
    _same(P.mean(d, axis=2), R.mean(x, axis=2), True)
    _same(P.array_interpolate_linear(d[:60, :60].contiguous(), axis=2), R.array_interpolate_linear(x[:60, :60], axis=2), True)
    # and it is not slow: the same cube through the forced TMA path takes several times longer
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    P.median(d, axis=2)
    ev[1].record()
    torch.cuda.synchronize()
    assert ev[0].elapsed_time(ev[1]) < 2.0, "median of a 64 MB cube with inner = 4 took more than 2 ms"


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.uint16])
def test_warp_per_row_kernel_for_many_short_contiguous_rows(P, torch, monkeypatch, dtype):
    """Reducing the LAST axis of a cube with many rows (x of (t, bands, y, x)): k_rows_warp (a warp per row, float64
    partials) == oracle within the contiguous-axis tolerance, == the column kernel for the exact reducers; rows that are
    all NaN / hold one sample; unaligned rows (odd row length); both nodata modes."""
    rng = np.random.default_rng(31)
    for X in (1000, 1023, 256):
        if np.dtype(dtype).kind == "f":
            x = _mk(rng, (7, 3, 40, X), dtype, 0.3)
            x[0, 0, 0, :] = np.nan
            x[0, 0, 1, 1:] = np.nan
        else:
            x = rng.integers(0, 10000, size=(7, 3, 40, X)).astype(dtype)
        d = torch.from_numpy(x).cuda()
        for ign in (True, False):
            for name in ("mean", "sum", "sd", "variance"):
                with np.errstate(all="ignore"):
                    want = R.BY_NAME[name](x, ignore_nodata=ign, axis=3)
                got = _fn(P, name)(d, ignore_nodata=ign, axis=3)
                np.testing.assert_allclose(np.asarray(got.cpu()), want, rtol=2e-6, equal_nan=True, err_msg=f"{name} X={X}")
            for name in ("min", "max"):
                _same(_fn(P, name)(d, ignore_nodata=ign, axis=3), R.BY_NAME[name](x, ignore_nodata=ign, axis=3), True)
        np.testing.assert_array_equal(np.asarray(P.count(d, axis=3).cpu()), R.count(x, axis=3))
        monkeypatch.setenv("EOC_ROWS_WARP", "0")
        _same(P._max(d, axis=3), R._max(x, axis=3), True)
        monkeypatch.delenv("EOC_ROWS_WARP")


def test_selection_over_a_contiguous_axis_of_up_to_1024_samples(P, torch):
    """median / quantiles over the LAST axis (x of a (t, bands, y, x) cube: inner = 1) with 513 .. 1024 samples: TMA
    cannot address the view, so the host layer reduces a transposed copy with the eight-run kernel instead of falling
    to the any-length general path; == oracle, quantile axis in place."""
    rng = np.random.default_rng(99)
    x = _mk(rng, (6, 2, 64, 1000), np.float32, 0.3)       # 768 columns ... too few: forced below
    big = np.tile(x, (16, 8, 1, 1))                          # (96, 16, 64, 1000): 98304 columns of 1000 samples
    d = torch.from_numpy(big).cuda()
    got = P.median(d, axis=3)
    _same(got[:6, :2], R.median(x, axis=3), True)
    q = P.quantiles(d, probabilities=[0.25, 0.75], axis=3)
    assert tuple(q.shape) == (96, 16, 64, 2)
    np.testing.assert_array_equal(q[:6, :2].cpu().numpy(), R.quantiles(x, probabilities=[0.25, 0.75], axis=3))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    P.median(d, axis=3)
    ev[1].record()
    torch.cuda.synchronize()
    assert ev[0].elapsed_time(ev[1]) < 10.0, "median over a contiguous 1000-sample axis took the general path"


@pytest.mark.parametrize("dtype", [np.uint16, np.int16, np.uint8])
def test_multi_selection_up_to_1024_integer_samples(P, torch, monkeypatch, dtype):
    """513 .. 1024 raw integer samples on k_select_multi<128, ..., In>: median / quantiles == numpy (float64), forced on
    for a small cube, and the general path when the library declines."""
    info = np.iinfo(dtype)
    rng = np.random.default_rng(7 + np.dtype(dtype).itemsize)
    for R_ in (513, 1024):
        x = rng.integers(info.min, int(info.max) + 1, size=(R_, 2, 160), dtype=np.int64).astype(dtype)
        d = torch.from_numpy(x).cuda()
        for flag in ("1", "0"):
            monkeypatch.setenv("EOC_SELECT_TMA", flag)
            got = P.median(d, axis=0)
            assert got.dtype == torch.float64
            _same(got, R.median(x, axis=0), True)
            q = P.quantiles(d, probabilities=[0.1, 0.9], axis=0)
            np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(x, probabilities=[0.1, 0.9], axis=0))


@pytest.mark.parametrize("R_", [513, 730, 1000, 1024])
def test_multi_selection_up_to_1024_samples(P, torch, monkeypatch, R_):
    """513 .. 1024 float32 samples (two to three years of daily scenes) on k_select_multi (eight runs of 128), forced on
    for a small cube: median, quantiles, both nodata modes == oracle == the any-length general path, bit for bit."""
    from openeo_processes_dask_b200 import engine

    rng = np.random.default_rng(R_)
    x = _mk(rng, (R_, 2, 160), np.float32, 0.3)
    x[:, 0, 1] = np.nan
    x[3:, 0, 2] = np.nan
    x[:, 0, 0] = 1.0
    d = torch.from_numpy(x).cuda()
    res = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("EOC_SELECT_TMA", flag)      # "0": the library declines, the host layer takes the general path
        outs = []
        for ign in (True, False):
            outs.append(P.median(d, ignore_nodata=ign, axis=0))
            _same(outs[-1], R.median(x, ignore_nodata=ign, axis=0), True)
        q = P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=0)
        np.testing.assert_array_equal(q.cpu().numpy(), R.quantiles(x, probabilities=[0.1, 0.5, 0.9], axis=0))
        res[flag] = outs
    for a, b in zip(res["1"], res["0"]):
        _same(a, b.cpu().numpy(), True)


@pytest.mark.parametrize("R_,dtype", [(513, np.float32), (730, np.float32), (1500, np.float64), (600, np.uint16),
                                      (2049, np.float32)])
def test_any_length_selection_general_path(P, torch, R_, dtype):
    """median / quantiles / sort along an axis longer than 512 (two years of daily scenes, a
    spatial dimension) go through eoc_select_general instead of raising; numpy and device input; bit-exact."""
    rng = np.random.default_rng(R_)
    if np.dtype(dtype).kind == "f":
        x = _mk(rng, (R_, 2, 37), dtype, 0.3)
        x[:, 0, 1] = np.nan
        x[3:, 0, 2] = np.nan
    else:
        x = rng.integers(0, 10000, size=(R_, 2, 37)).astype(dtype)
    for data in (x, torch.from_numpy(x).cuda()):
        for ign in (True, False):
            _same(P.median(data, ignore_nodata=ign, axis=0), R.median(x, ignore_nodata=ign, axis=0), True)
        q = P.quantiles(data, probabilities=[0.1, 0.5, 0.9], axis=0)
        want = R.quantiles(x.astype(np.float64) if np.dtype(dtype).kind != "f" else x, probabilities=[0.1, 0.5, 0.9], axis=0)
        np.testing.assert_allclose(np.asarray(q.cpu() if hasattr(q, "cpu") else q)[:, :, 3:], want[:, :, 3:], rtol=RTOL,
                                   equal_nan=True)
    y = np.ascontiguousarray(np.moveaxis(x, 0, -1))           # sort the way apply_dimension calls it: last axis
    for asc in (True, False):
        for nd in (True, False):
            got = P.sort(torch.from_numpy(y).cuda(), asc=asc, nodata=nd, axis=2)
            _same(got, R.sort(y, asc=asc, nodata=nd, axis=2), True)
    z = np.ascontiguousarray(np.moveaxis(x, 0, 1))            # (2, R, 37): outer > 1
    _same(P.median(torch.from_numpy(z).cuda(), axis=1), R.median(z, axis=1), True)


@pytest.mark.parametrize("shape,dtype,axis", [((100, 3, 9, 140), np.float32, 0), ((3, 100, 1000), np.float32, 1),
                                              ((301, 2, 260), np.float32, 0), ((64, 5, 132), np.float64, 0),
                                              ((365, 1, 128), np.float32, 0), ((50, 7, 129), np.float32, 0)])
@pytest.mark.parametrize("labels", [False, True])
def test_interpolate_linear_tma_kernel(P, torch, monkeypatch, shape, dtype, axis, labels):
    """k_interp_tma (TMA in, bulk tensor store out, Markstein slope when the positions are indices), forced on for
    small cubes, == oracle == the per-thread-load kernel, bit for bit: ragged last tile (clipped store), outer > 1,
    two row boxes (R = 301), float64, label positions (exact division), columns that stay as they are (all finite,
    fewer than two finite, infinities), an inner extent TMA cannot address (falls back)."""
    rng = np.random.default_rng(sum(shape) + int(labels))
    x = _mk(rng, shape, dtype, 0.35)
    idx = [slice(None)] * len(shape)
    idx[axis] = slice(None)
    moved = np.moveaxis(x, axis, 0)
    moved[:, ..., 0] = 1.0                 # nothing missing
    moved[1:, ..., 1] = np.nan             # one finite sample: stays as it is
    moved[:, ..., 2] = np.nan              # none
    moved[3, ..., 3] = np.inf              # an infinity is a gap slot
    moved[0, ..., 4] = -np.inf
    n = shape[axis]
    pos = np.sort(rng.choice(np.arange(4 * n), size=n, replace=False)).astype(float) if labels else None
    want = R.array_interpolate_linear(x, axis=axis, positions=pos)
    d = torch.from_numpy(x).cuda()
    kw = {"dim_labels": list(pos)} if labels else {}
    monkeypatch.setenv("EOC_INTERP_TMA", "1")
    got = P.array_interpolate_linear(d, axis=axis, **kw)
    _same(got, want, True)
    monkeypatch.setenv("EOC_INTERP_TMA", "0")
    old = P.array_interpolate_linear(d, axis=axis, **kw)
    _same(old, want, True)


@pytest.mark.parametrize("R_,dtype", [(500, np.float32), (300, np.float64), (1200, np.float32)])
def test_interpolate_linear_long_axis(P, torch, R_, dtype):
    """array_interpolate_linear along an axis too long for the shared-memory tile: the global-memory variant."""
    rng = np.random.default_rng(R_)
    x = rng.normal(size=(3, 41, R_)).astype(dtype)
    x[rng.random(x.shape) < 0.4] = np.nan
    x[0, 0, :] = np.nan
    x[0, 1, 5:] = np.nan
    x[0, 2, 7] = np.inf
    got = P.array_interpolate_linear(torch.from_numpy(x).cuda(), axis=2)
    _same(got, R.array_interpolate_linear(x, axis=2), True)
    y = np.ascontiguousarray(np.moveaxis(x, 2, 1))            # axis 1: inner = 41
    _same(P.array_interpolate_linear(torch.from_numpy(y).cuda(), axis=1), R.array_interpolate_linear(y, axis=1), True)
