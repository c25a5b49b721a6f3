"""Host-side logic of the drop-in layer (CPU only): view planning, y-strips, temporal bins, the
@process calling convention, the cube stand-in, and that the C-ABI library loads and exports every
symbol include/eoc.h declares.  No kernel is launched here."""

import ctypes
import functools
import os
import re

import numpy as np
import pandas as pd
import pytest

from oracle import cubes as OC
from oracle import synth
from openeo_processes_dask_b200 import _lib as L
from openeo_processes_dask_b200.process_implementations import core
from openeo_processes_dask_b200.process_implementations.core import ParameterReference, process
from openeo_processes_dask_b200.process_implementations.cubes import DataCube
from openeo_processes_dask_b200.process_implementations.cubes import _periods
from openeo_processes_dask_b200.process_implementations.cubes import aggregate as agg
from openeo_processes_dask_b200.process_implementations.cubes._xr_interop import OpenEODims
from openeo_processes_dask_b200.process_implementations.exceptions import (
    DimensionNotAvailable, OpenEOException, ProcessParameterMissing, TooManyDimensions)
from openeo_processes_dask_b200.views import plan_view, strip_bounds, contiguous_strides

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- C ABI ------------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "eoc.h")).read()
    declared = set(re.findall(r"^\s*(?:int|size_t|int64_t|const char\*)\s+(eoc_\w+)\s*\(", header, re.M))
    assert declared == set(L.EXPORTS), declared ^ set(L.EXPORTS)
    lib = L.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.eoc_abi_version() == 3
    assert isinstance(lib.eoc_last_error(), bytes)


def test_invalid_arguments_are_reported_not_crashed():
    lib = L.lib()
    v = L.make_view(0, 99, 1, 4, 4, 16, 4)
    rc = lib.eoc_reduce(ctypes.byref(v), 1, 1, 1, L.outs_array({}), None, 0, None)
    assert rc == -1 and b"dtype" in lib.eoc_last_error()
    v = L.make_view(0, L.F32, 1, 4, 4, 16, 4)
    rc = lib.eoc_select(ctypes.byref(v), 1, None, 0, None, None, L.F64, None)
    assert rc == -1


# ---- views ------------------------------------------------------------------------------------------
def _check_view(shape, axis, transpose=None, index=None):
    a = np.arange(int(np.prod(shape)), dtype=np.int64).reshape(shape)
    if transpose:
        a = a.transpose(transpose)
    if index is not None:
        a = a[index]
    strides = tuple(s // a.itemsize for s in a.strides)
    v = plan_view(a.shape, strides, axis)
    if v is None:
        return None
    base = a.ravel()[0] if a.size else 0
    flat = np.lib.stride_tricks.as_strided(a, shape=(v.outer, v.R, v.inner),
                                           strides=(v.stride_outer * 8, v.stride_R * 8, 8)) if a.size else None
    want = a.sum(axis=axis)
    assert v.out_shape == want.shape
    if a.size:
        np.testing.assert_array_equal(flat.sum(axis=1).reshape(v.out_shape), want)
    return v


def test_plan_view_layouts():
    v = _check_view((5, 3, 4, 6), 0)           # t of (t,b,y,x)
    assert (v.outer, v.R, v.inner, v.stride_R) == (1, 5, 72, 72)
    v = _check_view((3, 5, 4, 6), 1)           # t of (b,t,y,x)
    assert (v.outer, v.R, v.inner, v.stride_outer, v.stride_R) == (3, 5, 24, 120, 24)
    v = _check_view((4, 6, 5, 3), 2)           # t of (x,y,t,b): the reference fixture layout
    assert (v.outer, v.R, v.inner) == (24, 5, 3)
    v = _check_view((4, 6, 5, 3), 3)           # bands, contiguous
    assert (v.outer, v.R, v.inner, v.stride_R) == (120, 3, 1, 1)
    v = _check_view((5, 3, 4, 6), (2, 3))      # reduce_spatial on (t,b,y,x)
    assert (v.outer, v.R, v.inner) == (15, 24, 1)
    v = _check_view((4, 6, 5, 3), (0, 1))      # reduce_spatial on the fixture layout
    assert (v.outer, v.R, v.inner) == (1, 24, 15)
    v = _check_view((5, 3, 4, 6), None)
    assert (v.outer, v.R, v.inner) == (1, 360, 1)
    # a band selection cube[:, 1] is reducible over t without a copy
    v = _check_view((5, 3, 4, 6), 0, index=(slice(None), 1))
    assert (v.outer, v.R, v.inner, v.stride_R) == (1, 5, 24, 72)
    # layouts that need a copy
    assert _check_view((5, 3, 4, 6), 0, transpose=(0, 1, 3, 2)) is None
    assert _check_view((5, 3, 4, 6), (0, 2)) is None
    with pytest.raises(ValueError):
        plan_view((2, 3), (3, 1), 2)
    assert contiguous_strides((5, 3, 4, 6)) == (72, 24, 6, 1)


def test_strip_bounds():
    # 10980 rows over 8 GPUs -> 4 x 1373 + 4 x 1372
    sizes = [b - a for a, b in (strip_bounds(10980, 8, r) for r in range(8))]
    assert sizes == [1373] * 4 + [1372] * 4
    edges = [strip_bounds(5490, 8, r) for r in range(8)]
    assert edges[0][0] == 0 and edges[-1][1] == 5490
    assert all(edges[i][1] == edges[i + 1][0] for i in range(7))
    assert strip_bounds(3, 8, 7) == (3, 3)
    with pytest.raises(ValueError):
        strip_bounds(10, 2, 2)


# ---- temporal bins ------------------------------------------------------------------------------------
def test_plan_bins_matches_pandas_cut_right_closed():
    t = pd.date_range("2021-01-01", "2021-12-31", freq="D").values  # midnight: ON the edges (quirk 8)
    months = pd.date_range("2021-01-01", "2022-01-01", freq="MS").strftime("%Y-%m-%dT%H:%M:%S")
    intervals = [[a, b] for a, b in zip(months[:-1], months[1:])]
    groups, labels = agg.plan_bins(t, intervals)
    assert [len(g) for g in groups] == [31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 30]
    assert labels[0] == "2021-01-01T00:00:00" and isinstance(labels[0], str)
    edges, labels_all, labels_sel = OC.temporal_bins(intervals)
    codes = OC.bin_codes(t, edges)
    for k, g in enumerate(groups):
        assert g == np.nonzero(codes == k)[0].tolist()
    assert 0 not in groups[0]  # Jan-01T00:00 equals the first interval's START: excluded by (start, end]


def test_plan_bins_gap_and_labels():
    # the reference test's intervals (tests/test_aggregate.py:28-34): a gap between the two halves
    t = pd.date_range("2018-01-01", "2019-01-01", periods=100).values
    intervals = [["2018-01-01T12:00:00", "2018-06-01T12:00:00"], ["2018-07-01T12:00:00", "2018-12-01T12:00:00"]]
    groups, labels = agg.plan_bins(t, intervals, ["half-1", "half-2"])
    assert labels == ["half-1", "half-2"] and len(groups) == 2
    edges, labels_all, _ = OC.temporal_bins(intervals, ["half-1", "half-2"])
    assert list(labels_all) == ["half-1", "1", "half-2"]
    codes = OC.bin_codes(t, edges)
    assert groups[0] == np.nonzero(codes == 0)[0].tolist()
    assert groups[1] == np.nonzero(codes == 2)[0].tolist()
    with pytest.raises(NotImplementedError):
        agg.plan_bins(t, [["2018-01-01", "2018-06-01"], ["2018-05-01", "2018-12-01"]])


@pytest.mark.parametrize("extent,period,expected", [
    (["2018-05-01", "2018-05-02"], "hour", 25), (["2018-05-01", "2018-06-01"], "day", 32),
    (["2018-05-01", "2018-06-01"], "week", 5), (["2018-05-01", "2018-08-31"], "dekad", 12),
    (["2018-05-15", "2018-08-14"], "dekad", 10), (["2018-05-01", "2018-06-01"], "month", 2),
    (["2018-01-01", "2018-12-31"], "season", 5), (["2018-05-01", "2019-04-30"], "tropical-season", 2),
    (["2018-01-01", "2018-12-31"], "year", 1), (["2019-01-01", "2022-12-31"], "decade", 2),
    (["2019-01-01", "2022-12-31"], "decade-ad", 2),
])
def test_period_label_counts(extent, period, expected):
    """The label counts the reference asserts (tests/test_aggregate.py:99-150), 4 samples per cube."""
    t = pd.date_range(np.datetime64(extent[0]), np.datetime64(extent[1]), periods=4).values
    if period in agg._PANDAS_FREQ:
        groups, labels = _periods.calendar_groups(t, agg._PANDAS_FREQ[period])
    else:
        intervals, labels = _periods.period_intervals(t, period)
        groups, labels = agg.plan_bins(t, intervals, labels)
    assert len(labels) == expected == len(groups)
    assert sum(len(g) for g in groups) <= 4


def test_period_labels():
    t = pd.date_range("2018-05-15", "2018-08-14", periods=4).values
    _, labels = _periods.period_intervals(t, "dekad")
    assert labels[0] == "2018-13" and labels[-1] == "2018-22"
    _, labels = _periods.period_intervals(pd.date_range("2018-05-01", "2019-04-30", periods=4).values,
                                          "tropical-season")
    assert labels == ["2018-mjjaso", "2018-ndjfma"]
    _, labels = _periods.period_intervals(pd.date_range("2019-01-01", "2022-12-31", periods=4).values, "decade-ad")
    assert labels == ["2011", "2021"]
    with pytest.raises(NotImplementedError):
        _periods.period_intervals(t, "fortnight")


# ---- @process (mirrors the reference's tests/test_core.py) ---------------------------------------------
def test_process_decorator_resolves_references():
    def f(param3, param4, param5):
        return param3, param4, param5

    result = process(f)(1, 5, param3=3, param4=ParameterReference(from_parameter="test_param_ref_4"),
                        param5=ParameterReference(from_parameter="test_param_ref_5"),
                        named_parameters={"test_param_ref_2": 2, "test_param_ref_4": 4},
                        positional_parameters={"test_param_ref_5": 1})
    assert result == (3, 4, 5)


def test_process_decorator_missing_parameter():
    def f(param1, param2=6):
        return param1 * param2

    with pytest.raises(ProcessParameterMissing):
        process(f)(param1=ParameterReference(from_parameter="test_param_ref"), parameters={"wrong_param": 2})
    with pytest.raises(ProcessParameterMissing):
        process(f)(ParameterReference(from_parameter="test_param_ref"), parameters={"wrong_param": 2})
    assert issubclass(ProcessParameterMissing, OpenEOException)


def test_process_decorator_axis():
    def with_axis(param1, param2=6, axis=None):
        return param1, param2, axis

    assert process(with_axis)(param1=1, param2=2) == (1, 2, None)

    def without_axis(param1, param2=6):
        return param1, param2

    assert process(without_axis)(param1=1, param2=2, axis=None, context={}, dim_labels=[1]) == (1, 2)


# ---- cube stand-in + cube-level processes with a numpy reducer -------------------------------------------
def _cube():
    data = np.arange(2 * 3 * 4 * 2, dtype=np.float32).reshape(2, 3, 4, 2)
    return DataCube(data, ("x", "y", "t", "bands"),
                    {"x": [10.0, 11.0], "y": [1.0, 2.0, 3.0],
                     "t": pd.date_range("2018-05-01", "2018-06-01", periods=4).values, "bands": ["B04", "B08"]},
                    attrs={"crs": "EPSG:4326"})


def test_openeo_dims():
    d = OpenEODims(("bands", "Time", "lat", "lon", "foo"))
    assert d.temporal_dims == ("Time",) and d.band_dims == ("bands",)
    assert d.spatial_dims == ("lon", "lat") and d.x_dim == "lon" and d.y_dim == "lat"
    assert d.other_dims == ("foo",)


def test_reduce_dimension_glue_with_plain_reducer():
    from openeo_processes_dask_b200.process_implementations.cubes import reduce_dimension, reduce_spatial

    cube = _cube()
    reducer = functools.partial(process(lambda data, axis=None: np.mean(data, axis=axis)),
                                data=ParameterReference(from_parameter="data"))
    out = reduce_dimension(data=cube, reducer=reducer, dimension="t")
    assert out.dims == ("x", "y", "bands") and out.attrs["crs"] == "EPSG:4326"
    np.testing.assert_array_equal(out.data, cube.data.mean(axis=2))
    assert out.attrs["reduced_dimensions_min_values"]["t"] == np.datetime64("2018-05-01")
    with pytest.raises(DimensionNotAvailable):
        reduce_dimension(data=cube, reducer=reducer, dimension="time")
    sp = reduce_spatial(data=cube, reducer=reducer)
    assert sp.dims == ("t", "bands")
    np.testing.assert_array_equal(sp.data, cube.data.mean(axis=(0, 1)))
    want, dims, _ = OC.reduce_dimension(cube.data, cube.dims, {"t": cube["t"].values},
                                        lambda a, axis: np.mean(a, axis=axis), "t")
    np.testing.assert_array_equal(out.data, want)


def test_aggregate_temporal_glue_with_plain_reducer():
    """Per-group path (reducer not fusable) against the oracle's restatement of groupby_bins."""
    from openeo_processes_dask_b200.process_implementations.cubes import aggregate_temporal

    rng = np.random.default_rng(1)
    t = pd.date_range("2018-01-01", "2019-01-01", periods=100).values
    cube = DataCube(rng.normal(size=(3, 2, 100, 2)), ("x", "y", "t", "bands"),
                    {"t": t, "bands": ["a", "b"]}, attrs={"k": 1})
    intervals = [["2018-01-01T12:00:00", "2018-06-01T12:00:00"], ["2018-07-01T12:00:00", "2018-12-01T12:00:00"]]
    reducer = functools.partial(process(lambda data, axis=None: np.nanmean(data, axis=axis)),
                                data=ParameterReference(from_parameter="data"))
    out = aggregate_temporal(data=cube, intervals=intervals, reducer=reducer, labels=["half-1", "half-2"])
    assert out.dims == cube.dims and out.shape == (3, 2, 2, 2) and out.attrs == cube.attrs
    assert list(out["t"].values) == ["half-1", "half-2"] and isinstance(out["t"].values[0], str)
    want, _, _ = OC.aggregate_temporal(cube.data, cube.dims, {"t": t}, intervals,
                                       lambda a, axis: np.nanmean(a, axis=axis), ["half-1", "half-2"])
    np.testing.assert_array_equal(out.data, want)
    with pytest.raises(DimensionNotAvailable):
        aggregate_temporal(data=cube, intervals=intervals, reducer=reducer, dimension="nope")
    two = DataCube(np.zeros((2, 2)), ("t", "time"), {})
    with pytest.raises(TooManyDimensions):
        aggregate_temporal(data=two, intervals=intervals, reducer=reducer)


def test_fusable_detection():
    from openeo_processes_dask_b200.process_implementations import count, mean, median

    ref = ParameterReference(from_parameter="data")
    assert agg._fusable(functools.partial(process(mean), data=ref, ignore_nodata=False)) == (L.OP_MEAN, False)
    assert agg._fusable(functools.partial(process(mean), data=ref)) == (L.OP_MEAN, True)
    assert agg._fusable(functools.partial(process(count), data=ref)) == (L.OP_COUNT, True)
    assert agg._fusable(functools.partial(process(count), data=ref, condition=True)) is None
    assert agg._fusable(functools.partial(process(median), data=ref)) is None
    assert agg._fusable(lambda data, axis=None: 0) is None


# ---- synthetic cube twin ----------------------------------------------------------------------------------
def test_synth_strip_addressable():
    full = synth.synth_cube(43, (3, 2, 8, 5), np.float32, 0.3)
    win = synth.synth_window(43, (3, 2, 8, 5), (slice(0, 3), slice(0, 2), slice(2, 6), slice(0, 5)), np.float32, 0.3)
    np.testing.assert_array_equal(win, full[:, :, 2:6, :])
    frac = np.isnan(synth.synth_cube(1, (200, 500), np.float32, 0.3)).mean()
    assert 0.29 < frac < 0.31
    u = synth.synth_cube(45, (4, 100), np.uint16)
    assert u.dtype == np.uint16 and u.max() < 10000


# ---- paired views (mask fused into a reduce) and the oracle's mask ------------------------------------------
def test_plan_view_pair_broadcast_mask():
    from openeo_processes_dask_b200.views import contiguous_strides, plan_view_pair

    T, B, Y, X = 100, 4, 30, 40
    shape = (T, B, Y, X)
    sa = contiguous_strides(shape)
    # (t,y,x) mask over a (t,bands,y,x) cube reduced over t: bands become `outer` with mask stride 0
    va, vm, tail = plan_view_pair(shape, sa, (Y * X, 0, X, 1), 0)
    assert (va.outer, va.R, va.inner, va.stride_outer, va.stride_R) == (B, T, Y * X, Y * X, B * Y * X)
    assert (vm.outer, vm.R, vm.inner, vm.stride_outer, vm.stride_R) == (B, T, Y * X, 0, Y * X)
    assert tail == 1 and va.out_shape == (B, Y, X) and va.axis_pos == 0
    # same mask, reduced over bands: the mask's R stride is 0
    va, vm, tail = plan_view_pair(shape, sa, (Y * X, 0, X, 1), 1)
    assert (vm.outer, vm.R, vm.stride_outer, vm.stride_R, tail) == (T, B, Y * X, 0, 0)
    # identical layouts keep the largest inner
    va, vm, tail = plan_view_pair(shape, sa, sa, 0)
    assert (va.outer, va.inner, tail) == (1, B * Y * X, 0) and va == vm
    # (y,x) mask: broadcast over t and bands
    va, vm, tail = plan_view_pair(shape, sa, (0, 0, X, 1), 0)
    assert (vm.stride_outer, vm.stride_R, vm.inner) == (0, 0, Y * X)
    # a layout that cannot be paired (outer would need two strides)
    shape5 = (2, T, B, Y, X)
    assert plan_view_pair(shape5, contiguous_strides(shape5), (T * Y * X, Y * X, 0, X, 1), 1) is None
    # every element is addressed exactly once by both views
    shape = (5, 3, 2, 4)
    sa, sm = contiguous_strides(shape), (8, 0, 4, 1)
    va, vm, _ = plan_view_pair(shape, sa, sm, 0)
    data = np.arange(np.prod(shape)).reshape(shape)
    m = np.arange(5 * 2 * 4).reshape(5, 1, 2, 4)
    for o in range(va.outer):
        for r in range(va.R):
            for i in range(va.inner):
                assert data.reshape(-1)[o * va.stride_outer + r * va.stride_R + i] == data[r, o, i // 4, i % 4]
                assert m.reshape(-1)[o * vm.stride_outer + r * vm.stride_R + i] == m[r, 0, i // 4, i % 4]


def test_oracle_mask_semantics():
    from oracle import cubes as OC

    x = np.arange(24, dtype=np.float32).reshape(2, 3, 2, 2)
    dims = ("t", "bands", "y", "x")
    m = np.array([[1, 0], [np.nan, 2.0]], dtype=np.float32)   # NaN in the mask keeps the sample
    out, d = OC.mask(x, dims, m, ("y", "x"))
    assert d == dims and out.dtype == np.float32
    assert np.isnan(out[:, :, 0, 0]).all() and np.isnan(out[:, :, 1, 1]).all()
    np.testing.assert_array_equal(out[:, :, 0, 1], x[:, :, 0, 1])
    np.testing.assert_array_equal(out[:, :, 1, 0], x[:, :, 1, 0])
    out, _ = OC.mask(x.astype(np.uint16), dims, m, ("y", "x"), replacement=0)
    assert out.dtype == np.uint16 and out[0, 0, 0, 0] == 0
    out, _ = OC.mask(x.astype(np.uint16), dims, m, ("y", "x"))
    assert out.dtype == np.float64
    # a mask without t on a (bands,t,y,x) cube leaves the result in (t,bands,y,x) order (reference quirk)
    out, d = OC.mask(np.transpose(x, (1, 0, 2, 3)), ("bands", "t", "y", "x"), np.zeros((3, 2, 2)), ("bands", "y", "x"))
    assert d == dims and out.shape == x.shape


def test_merge_cubes_glue_with_plain_resolver():
    """merge_cubes host logic (cubes/merge.py:54-295) on numpy cubes with a plain python overlap resolver -- no
    kernel involved: label bookkeeping, join order, the `__cubes__` stack, broadcasting views, errors."""
    from openeo_processes_dask_b200.process_implementations.cubes.merge import merge_cubes
    from openeo_processes_dask_b200.process_implementations.exceptions import OverlapResolverMissing

    rng = np.random.default_rng(5)
    dims = ("x", "y", "t", "bands")
    full = rng.normal(size=(4, 3, 5, 4))
    coords = {"x": np.arange(4.0), "y": np.arange(3.0), "t": np.arange("2021-01-01", "2021-01-06", dtype="datetime64[D]"),
              "bands": np.array(["B02", "B03", "B04", "B08"])}

    def cube(arr, d=dims, **over):
        c = {k: v for k, v in {**coords, **over}.items() if k in d}
        return DataCube(arr, d, c, attrs={"a": 1})

    def resolver(f):
        @process
        def fn(x, y):
            return f(x, y)
        return functools.partial(fn, x=ParameterReference(from_parameter="x"), y=ParameterReference(from_parameter="y"))

    # example 1: disjoint bands, cube2 first in label order -> cube1's bands, then cube2's
    out = merge_cubes(cube(full[..., 2:], bands=coords["bands"][2:]), cube(full[..., :2], bands=coords["bands"][:2]))
    assert list(out["bands"].values) == ["B04", "B08", "B02", "B03"] and out.dims == dims and out.attrs == {"a": 1}
    np.testing.assert_array_equal(out.values, full[..., [2, 3, 0, 1]])
    want, _, wc = OC.merge_cubes((full[..., 2:], dims, {**coords, "bands": coords["bands"][2:]}),
                                 (full[..., :2], dims, {**coords, "bands": coords["bands"][:2]}))
    np.testing.assert_array_equal(out.values, want)
    assert list(wc["bands"]) == list(out["bands"].values)
    # disjoint time steps: ascending labels whatever the argument order
    out = merge_cubes(cube(full[:, :, 3:], t=coords["t"][3:]), cube(full[:, :, :3], t=coords["t"][:3]))
    np.testing.assert_array_equal(out["t"].values, coords["t"])
    np.testing.assert_array_equal(out.values, full)
    # example 2: partial overlap of one dimension
    a, b = cube(full[:, :, :4], t=coords["t"][:4]), cube(full[:, :, 2:] * 10, t=coords["t"][2:])
    with pytest.raises(OverlapResolverMissing):
        merge_cubes(a, b)
    out = merge_cubes(a, b, resolver(lambda x, y: x + y))
    expect = np.concatenate([full[:, :, :2], full[:, :, 2:4] * 11, full[:, :, 4:] * 10], axis=2)
    np.testing.assert_allclose(out.values, expect, rtol=1e-15)
    want, _, _ = OC.merge_cubes((full[:, :, :4], dims, {**coords, "t": coords["t"][:4]}),
                                (full[:, :, 2:] * 10, dims, {**coords, "t": coords["t"][2:]}), lambda x, y: x + y)
    np.testing.assert_array_equal(out.values, want)
    # example 3: equal cubes; no resolver stacks, cube2 aligned to cube1's dimension and label order
    perm = [3, 1, 0, 2]
    other = DataCube(np.ascontiguousarray(full[..., perm].transpose(3, 2, 0, 1)), ("bands", "t", "x", "y"),
                     {**coords, "bands": coords["bands"][perm]})
    out = merge_cubes(cube(full), other)
    assert out.dims == ("__cubes__",) + dims and list(out["__cubes__"].values) == ["cube1", "cube2"]
    np.testing.assert_array_equal(out.values, np.stack([full, full]))
    out = merge_cubes(cube(full), other, resolver(lambda x, y: x - y))
    np.testing.assert_array_equal(out.values, np.zeros_like(full))
    # example 4: lower-dimensional cube broadcast as a VIEW (no expanded copy), x is always cube1
    plane = rng.normal(size=(3, 4))
    seen = {}

    def sub(x, y):
        seen["shapes"] = (x.shape, y.shape)
        return x - y

    out = merge_cubes(DataCube(plane, ("y", "x"), {"x": coords["x"], "y": coords["y"]}), cube(full), resolver(sub))
    assert seen["shapes"] == ((4, 3, 1, 1), (4, 3, 5, 4)) and out.dims == dims
    np.testing.assert_array_equal(out.values, plane.T[:, :, None, None] - full)
    with pytest.raises(OverlapResolverMissing):
        merge_cubes(cube(full), DataCube(plane, ("y", "x"), {"x": coords["x"], "y": coords["y"]}))
    with pytest.raises(ValueError):
        merge_cubes(cube(full), DataCube(plane[0], ("x",), {"x": coords["x"]}), resolver(sub))
    with pytest.raises(Exception, match="incompatible types"):
        merge_cubes(cube(full), full)
    # more than one partially overlapping dimension
    with pytest.raises(ValueError):
        merge_cubes(cube(full[:3, :, :4], x=coords["x"][:3], t=coords["t"][:4]),
                    cube(full[1:, :, 2:], x=coords["x"][1:], t=coords["t"][2:]), resolver(sub))
