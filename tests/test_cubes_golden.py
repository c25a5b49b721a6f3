"""Cube-level temporal aggregation against outputs of the REFERENCE's own code.

tests/golden/reference_cubes.{json,npz} were produced by tests/golden/make_golden_cubes.py, which loads the
reference's unmodified ``cubes/aggregate.py`` and CALLS ``get_intervals``, ``aggregate_temporal`` and
``aggregate_temporal_period`` (the last two on a duck cube whose ``groupby_bins`` is ``pandas.cut`` and whose
``resample`` is ``pandas.Series.resample``).  Here:

* not gpu: the oracle restatements (oracle/cubes.py) and the product's host logic (``_periods.period_intervals``,
  ``plan_bins``, ``calendar_groups``) reproduce those outputs case by case;
* gpu: the product's ``aggregate_temporal`` / ``aggregate_temporal_period`` (one segmented CUDA launch per call)
  return the reference's VALUES and labels -- bit-exact (the reduced axis is not contiguous, so the kernels follow
  numpy's operation order), for every reducer, numpy and device input.
"""

import functools
import json
import os
import warnings

import numpy as np
import pytest

from oracle import cubes as OC
from oracle import reducers as R

warnings.simplefilter("ignore", RuntimeWarning)
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(HERE, "golden", "reference_cubes.json")) as f:
        doc = json.load(f)
    return doc, np.load(os.path.join(HERE, "golden", "reference_cubes.npz"))


ORACLE_REDUCERS = {
    "mean": R.mean, "max": R._max, "median": R.median, "sd": R.sd, "count": R.count,
    "mean_keepnan": functools.partial(R.mean, ignore_nodata=False),
}
DIMS = ("t", "bands", "y", "x")


def _t(entry):
    return np.array(entry["t"], dtype="datetime64[ns]")


def _interval_result(fn, r):
    t = np.array([r["start"], r["end"]], dtype="datetime64[ns]")
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            iv, labels = fn(t, r["period"])
        return {"intervals": np.asarray(iv).tolist(), "labels": [str(x) for x in labels]}
    except ValueError as e:  # pandas' DateParseError is a ValueError
        return {"error": type(e).__name__}


def test_get_intervals_oracle_and_product_match_reference(gold):
    """569 extents x {decade, decade-ad, tropical-season, dekad}: bounds, labels and the raising cases."""
    from openeo_processes_dask_b200.process_implementations.cubes import _periods

    doc, _ = gold
    assert len(doc["get_intervals"]) > 500
    for r in doc["get_intervals"]:
        want = {k: r[k] for k in ("intervals", "labels", "error") if k in r}
        assert _interval_result(OC.get_intervals, r) == want, ("oracle", r["period"], r["start"], r["end"])
        assert _interval_result(_periods.period_intervals, r) == want, ("product", r["period"], r["start"], r["end"])


def test_oracle_aggregate_temporal_matches_reference(gold):
    doc, arrays = gold
    for name, e in doc["aggregate_temporal"].items():
        x = arrays[f"at/{name}/in"]
        for rname, fn in ORACLE_REDUCERS.items():
            out, dims, coords = OC.aggregate_temporal(x, DIMS, {"t": _t(e)}, e["intervals"], fn, e["labels_in"])
            np.testing.assert_array_equal(out, arrays[f"at/{name}/{rname}"], err_msg=f"{name}/{rname}")
            assert [str(v) for v in coords["t"]] == e["labels_out"]
    assert doc["overlap_error"] == "NotImplementedError"


def test_oracle_aggregate_temporal_period_matches_reference(gold):
    doc, arrays = gold
    for name, e in doc["aggregate_temporal_period"].items():
        key = name.replace("/", "__")
        assert "error" not in e
        for rname in ("mean", "max", "count", "median"):
            out, _, coords = OC.aggregate_temporal_period(arrays[f"atp/{key}/in"], DIMS, {"t": _t(e)},
                                                          ORACLE_REDUCERS[rname], e["period"])
            np.testing.assert_array_equal(out, arrays[f"atp/{key}/{rname}"], err_msg=f"{name}/{rname}")
            assert [str(v) for v in coords["t"]] == e["labels_out"], name


def test_product_bins_match_reference_membership(gold):
    """plan_bins (searchsorted, right-closed) and calendar_groups pick the members pandas.cut / resample pick."""
    from openeo_processes_dask_b200.process_implementations.cubes import _periods
    from openeo_processes_dask_b200.process_implementations.cubes import aggregate as agg

    doc, _ = gold
    for name, e in doc["aggregate_temporal"].items():
        groups, labels = agg.plan_bins(_t(e), e["intervals"], e["labels_in"])
        edges, labels_all, labels_sel = OC.temporal_bins(e["intervals"], e["labels_in"])
        codes = OC.bin_codes(_t(e), edges)
        where = {lab: i for i, lab in enumerate(labels_all)}
        assert labels == e["labels_out"] == [str(v) for v in labels_sel]
        for g, lab in zip(groups, labels):
            assert g == np.nonzero(codes == where[lab])[0].tolist(), (name, lab)
    for name, e in doc["aggregate_temporal_period"].items():
        if e["period"] in OC._PANDAS_FREQ:
            groups, labels = _periods.calendar_groups(_t(e), OC._PANDAS_FREQ[e["period"]])
            want_groups, want_labels = OC.resample_groups(_t(e), OC._PANDAS_FREQ[e["period"]])
            assert [list(g) for g in groups] == [g.tolist() for g in want_groups]
            assert [str(v) for v in labels] == e["labels_out"] == [str(v) for v in want_labels]


# ---- the CUDA path ---------------------------------------------------------------------------------------------
def _product_reducers(P):
    ref = P.ParameterReference(from_parameter="data")
    mk = lambda f, **kw: functools.partial(P.process(f), data=ref, **kw)  # noqa: E731
    return {"mean": mk(P.mean), "max": mk(P._max), "median": mk(P.median), "sd": mk(P.sd), "count": mk(P.count),
            "mean_keepnan": mk(P.mean, ignore_nodata=False)}


def _host(a):
    a = getattr(a, "tensor", a)
    return a.cpu().numpy() if hasattr(a, "cpu") else np.asarray(a)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["numpy", "cuda"])
def test_gpu_aggregate_temporal_values_match_reference(gold, kind):
    import torch

    import openeo_processes_dask_b200.process_implementations as P

    doc, arrays = gold
    reds = _product_reducers(P)
    for name, e in doc["aggregate_temporal"].items():
        x = arrays[f"at/{name}/in"]
        data = torch.from_numpy(x).cuda() if kind == "cuda" else x
        cube = P.DataCube(data, DIMS, {"t": _t(e), "bands": ["B04", "B08"]}, attrs={"crs": "EPSG:32633"})
        for rname, red in reds.items():
            out = P.aggregate_temporal(data=cube, intervals=e["intervals"], reducer=red, labels=e["labels_in"])
            want = arrays[f"at/{name}/{rname}"]
            got = _host(out.data)
            assert out.dims == DIMS and out.attrs == cube.attrs
            assert [str(v) for v in out["t"].values] == e["labels_out"]
            np.testing.assert_array_equal(got, want, err_msg=f"{name}/{rname}/{kind}")
            assert got.dtype == want.dtype, (name, rname, got.dtype, want.dtype)
    with pytest.raises(NotImplementedError):
        P.aggregate_temporal(data=cube, intervals=[["2021-01-01", "2021-02-01"], ["2021-01-15", "2021-03-01"]],
                             reducer=reds["mean"])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["numpy", "cuda"])
def test_gpu_aggregate_temporal_period_values_match_reference(gold, kind):
    """All ten periods (hour, day, week, dekad, month, season, tropical-season, year, decade, decade-ad)."""
    import torch

    import openeo_processes_dask_b200.process_implementations as P

    doc, arrays = gold
    reds = _product_reducers(P)
    seen = set()
    for name, e in doc["aggregate_temporal_period"].items():
        key = name.replace("/", "__")
        x = arrays[f"atp/{key}/in"]
        data = torch.from_numpy(x).cuda() if kind == "cuda" else x
        cube = P.DataCube(data, DIMS, {"t": _t(e), "bands": ["B04", "B08"]})
        for rname in ("mean", "max", "count", "median"):
            out = P.aggregate_temporal_period(data=cube, reducer=reds[rname], period=e["period"])
            want = arrays[f"atp/{key}/{rname}"]
            assert [str(v) for v in out["t"].values] == e["labels_out"], name
            np.testing.assert_array_equal(_host(out.data), want, err_msg=f"{name}/{rname}/{kind}")
        seen.add(e["period"])
    assert len(seen) == 10
