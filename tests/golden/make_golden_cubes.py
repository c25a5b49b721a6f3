"""Golden vectors for the CUBE-LEVEL temporal aggregation, produced by the REFERENCE's own code.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_cubes.py

``cubes/aggregate.py`` of the reference is loaded *unmodified* under its real module name with inert
stand-ins for the third-party modules that are absent here (geopandas, shapely, xvec, dask_geopandas,
xarray, openeo_pg_parser_networkx) -- the same loader as ``make_golden.py``.  Then

1. ``get_intervals`` (cubes/aggregate.py:104-203: decade / decade-ad / tropical-season / dekad interval and
   label synthesis -- plain numpy + pandas) is CALLED on a grid of temporal extents; its outputs (or the
   exception type it raises) are stored.  pandas here is 3.0.2; the function's output depends on pandas'
   ``date_range`` / ``period_range`` / ``strftime`` (e.g. ``period_range(freq="6M").strftime`` yields the period
   END dates, so the reference's tropical seasons start on Nov-30 / May-31).
2. ``aggregate_temporal`` (cubes/aggregate.py:29-101) is CALLED on a duck cube that implements the xarray
   surface the function touches: ``.openeo.temporal_dims``, ``cube[t].values``, ``cube[t] = ...``,
   ``copy.deepcopy``, ``groupby_bins(t, bins=...)`` -> ``.reduce(reducer, ...)``, ``cube[t + "_bins"] = labels``
   (length-checked like xarray), ``.sel`` and ``.rename``.  Bin membership inside the duck is ``pandas.cut`` --
   the very call xarray's ``groupby_bins`` makes -- so the reference's own edge / filler-label / label-selection
   arithmetic (lines 56-99) runs for real and decides which members land under which output label.  What the
   duck restates (marked [U] in oracle/cubes.py): groups reduced in index order, an empty bin re-indexed to NaN,
   the new axis at the position of ``t``.
3. ``aggregate_temporal_period`` (cubes/aggregate.py:206-253) is CALLED for all ten periods; the six
   pandas-frequency ones go through the duck's ``resample`` (= ``pandas.Series.resample`` with pandas' defaults per
   frequency, what xarray's resample builds its Grouper from [U]).

Reducers are the reference's own ``math.py`` / ``arrays.py`` functions.

Output (committed): tests/golden/reference_cubes.json, tests/golden/reference_cubes.npz
"""

import copy
import importlib.util
import json
import os
import sys
import warnings

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402


def load_reference_aggregate():
    math, arrays, logic = mg.load_reference_modules()

    class _Never:
        pass

    for name, attrs in (("geopandas", dict(GeoDataFrame=_Never)), ("dask_geopandas", dict(GeoDataFrame=_Never)),
                        ("shapely", {}), ("xvec", {})):
        mg._stub(name, **attrs)
    sys.modules["xarray"].Dataset = _Never
    ps = sys.modules["openeo_pg_parser_networkx.pg_schema"]
    ps.TemporalInterval = _Never
    ps.TemporalIntervals = _Never

    def load(rel, modname):
        spec = importlib.util.spec_from_file_location(modname, os.path.join(mg.REF, rel))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    pi = "openeo_processes_dask.process_implementations"
    load("process_implementations/data_model.py", pi + ".data_model")
    agg = load("process_implementations/cubes/aggregate.py", pi + ".cubes.aggregate")
    return agg, math, arrays


# ---- the xarray surface aggregate_temporal / aggregate_temporal_period touch -------------------------------
class _Coord:
    def __init__(self, values):
        self.values = np.asarray(values)


class _Dims:
    def __init__(self, temporal):
        self.temporal_dims = temporal


class DuckCube:
    def __init__(self, array, dims, coords, attrs=None):
        self.array = np.asarray(array)
        self.dims = tuple(dims)
        self.coords = {k: np.asarray(v) for k, v in coords.items()}
        self.attrs = dict(attrs or {})

    @property
    def openeo(self):
        return _Dims(tuple(d for d in self.dims if d in ("t", "time", "temporal", "DATE")))

    def __getitem__(self, key):
        return _Coord(self.coords[key])

    def __setitem__(self, key, values):
        values = np.asarray(values)
        if key in self.dims and len(values) != self.array.shape[self.dims.index(key)]:
            raise ValueError(f"conflicting sizes for dimension {key!r}: length {len(values)} on the coordinate "
                             f"and length {self.array.shape[self.dims.index(key)]} on the data")
        self.coords[key] = values

    def groupby_bins(self, dim, bins):
        codes = np.asarray(pd.cut(self.coords[dim], bins).codes)  # xarray: pd.cut(group, bins) with defaults
        groups = [np.nonzero(codes == b)[0] for b in range(len(bins) - 1)]
        return _Grouped(self, dim, groups, dim + "_bins", np.arange(len(bins) - 1))

    def resample(self, indexer):
        (dim, freq), = indexer.items()
        idx = pd.DatetimeIndex(self.coords[dim])
        series = pd.Series(np.arange(len(idx)), index=idx)
        groups, labels = [], []
        for label, members in series.resample(freq):
            groups.append(np.asarray(members.values, dtype=np.int64))
            labels.append(label.to_datetime64())
        return _Grouped(self, dim, groups, dim, np.asarray(labels, dtype="datetime64[ns]"))

    def sel(self, indexers):
        (dim, labels), = indexers.items()
        have = [str(v) for v in self.coords[dim]]
        pos = [have.index(str(lab)) for lab in labels]  # KeyError-equivalent: ValueError if absent
        ax = self.dims.index(dim)
        out = DuckCube(np.take(self.array, pos, axis=ax), self.dims, self.coords, self.attrs)
        out.coords[dim] = np.asarray([have[p] for p in pos])
        return out

    def rename(self, mapping):
        dims = tuple(mapping.get(d, d) for d in self.dims)
        return DuckCube(self.array, dims, {mapping.get(k, k): v for k, v in self.coords.items()}, self.attrs)


class _Grouped:
    def __init__(self, cube, dim, groups, new_dim, new_labels):
        self.cube, self.dim, self.groups, self.new_dim, self.new_labels = cube, dim, groups, new_dim, new_labels

    def reduce(self, reducer, keep_attrs=False, **kwargs):
        c = self.cube
        ax = c.dims.index(self.dim)
        parts = []
        for members in self.groups:
            if len(members) == 0:
                parts.append(None)
                continue
            parts.append(np.asarray(reducer(np.take(c.array, members, axis=ax), axis=ax, **kwargs)))
        proto = next(p for p in parts if p is not None)
        filled = [np.full(proto.shape, np.nan, dtype=np.result_type(proto.dtype, np.float32)) if p is None else p
                  for p in parts]
        if any(p is None for p in parts):
            filled = [f.astype(np.result_type(*filled)) for f in filled]
        out = np.stack(filled, axis=ax)
        dims = tuple(self.new_dim if d == self.dim else d for d in c.dims)
        coords = {k: v for k, v in c.coords.items() if k != self.dim}
        coords[self.new_dim] = self.new_labels
        return DuckCube(out, dims, coords, c.attrs if keep_attrs else {})


# ---- cases ---------------------------------------------------------------------------------------------------
def interval_cases():
    starts = ["2020-01-01", "2020-01-05", "2020-01-10", "2020-01-11", "2020-01-20", "2020-01-21", "2020-01-25",
              "2020-01-30", "2020-01-31", "2020-02-29", "2019-11-15", "2021-04-30", "2021-06-09", "2021-05-01",
              "2021-10-31", "2021-12-31", "1999-12-31", "2001-01-01"]
    spans = [0, 3, 9, 20, 45, 200, 400, 4000]
    out = []
    for s in starts:
        for d in spans:
            a = pd.Timestamp(s) + pd.Timedelta(hours=12)
            out.append((a.to_datetime64(), (a + pd.Timedelta(days=d)).to_datetime64()))
    # the extents of the reference's own tests (tests/test_aggregate.py:99-150: 2018-05-01 .. 2018-06-01, and a
    # decade-crossing one) and of the bench config (2021 daily)
    out.append((np.datetime64("2018-05-01T00:00:00"), np.datetime64("2018-06-01T00:00:00")))
    out.append((np.datetime64("2017-01-01T00:00:00"), np.datetime64("2023-12-31T00:00:00")))
    out.append((np.datetime64("2021-01-01T12:00:00"), np.datetime64("2021-12-31T12:00:00")))
    return out


def run_get_intervals(agg):
    rows = []
    for period in ("decade", "decade-ad", "tropical-season", "dekad"):
        for start, end in interval_cases():
            if period == "dekad" and (end - start) > np.timedelta64(1000, "D"):
                continue  # hundreds of dekads add bytes, not cases
            t = np.array([start, end], dtype="datetime64[ns]")
            entry = {"period": period, "start": str(t[0]), "end": str(t[1])}
            try:
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    iv, labels = agg.get_intervals({"t": _Coord(t)}, period)
                entry["intervals"] = np.asarray(iv).tolist()
                entry["labels"] = [str(x) for x in labels]
            except Exception as e:  # the reference raises for some extents (e.g. day 30 -> "2020-1-31" is fine, "2021-4-31" is not)
                entry["error"] = type(e).__name__
            rows.append(entry)
    return rows


def temporal_cases():
    """name -> (t labels, intervals, labels or None)"""
    daily = pd.date_range("2021-01-01T12:00:00", "2021-12-31T12:00:00", freq="D").values
    midnight = pd.date_range("2021-01-01", "2021-04-30", freq="D").values
    months = [[f"2021-{m:02d}-01T00:00:00", (f"2021-{m + 1:02d}-01T00:00:00" if m < 12 else "2022-01-01T00:00:00")]
              for m in range(1, 13)]
    fixture_t = pd.date_range("2018-05-01", "2018-06-01", periods=20).values  # like tests/conftest.py's temporal axis
    return {
        # tests/test_aggregate.py:28-34: two intervals with a gap
        "ref_test_gap": (fixture_t, [["2018-05-01T00:00:00", "2018-05-15T00:00:00"],
                                     ["2018-05-18T00:00:00", "2018-06-01T00:00:00"]], None),
        "monthly_noon": (daily, months, None),
        "monthly_midnight_edges": (midnight, months[:4], None),
        "custom_labels": (midnight, months[:3], ["jan", "feb", "mar"]),
        "pair_gap_single": (midnight, [["2021-01-05", "2021-01-20"], ["2021-01-20", "2021-02-10"],
                                       ["2021-03-01", "2021-03-15"]], None),
        "empty_bin": (midnight[:40], months[:4], None),
        "subday": (pd.date_range("2021-03-01", periods=50, freq="7h").values,
                   [["2021-03-01T00:00:00", "2021-03-04T12:00:00"], ["2021-03-04T12:00:00", "2021-03-09T00:00:00"],
                    ["2021-03-10T00:00:00", "2021-03-20T00:00:00"]], None),
    }


def period_cases():
    """name -> (t labels, period)"""
    hourly = pd.date_range("2018-05-01", "2018-05-02", periods=13).values
    month = pd.date_range("2018-05-01", "2018-06-01", periods=20).values
    two_years = pd.date_range("2019-11-20T06:00:00", "2021-12-10T06:00:00", freq="9D").values
    sparse = np.array(["2020-01-03", "2020-01-04", "2020-03-20", "2020-07-01", "2020-07-02", "2021-02-11"],
                      dtype="datetime64[ns]")
    decades = pd.date_range("1998-03-01", "2023-09-01", freq="173D").values
    out = {}
    for p in ("hour",):
        out[f"{p}/hourly"] = (hourly, p)
    for p in ("day", "week", "dekad", "month"):
        out[f"{p}/month"] = (month, p)
    for p in ("week", "dekad", "month", "season", "tropical-season", "year", "decade", "decade-ad"):
        out[f"{p}/two_years"] = (two_years, p)
    for p in ("day", "month", "season", "year"):
        out[f"{p}/sparse"] = (sparse, p)
    for p in ("year", "decade", "decade-ad", "tropical-season"):
        out[f"{p}/decades"] = (decades, p)
    return out


def _cube(t, seed, nan_frac=0.3):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 10000, size=(len(t), 2, 3, 4)).astype(np.float32)
    a[rng.random(a.shape) < nan_frac] = np.nan
    return DuckCube(a, ("t", "bands", "y", "x"), {"t": np.asarray(t), "bands": np.array(["B04", "B08"])},
                    {"crs": "EPSG:32633"})


def main():
    agg, math, arrays = load_reference_aggregate()
    warnings.simplefilter("ignore")
    doc = {"pandas": pd.__version__, "numpy": np.__version__, "get_intervals": run_get_intervals(agg),
           "aggregate_temporal": {}, "aggregate_temporal_period": {}}
    store = {}
    reducers = {
        "mean": lambda data, axis=None, **kw: math.mean(data, axis=axis),
        "max": lambda data, axis=None, **kw: math._max(data, axis=axis),
        "median": lambda data, axis=None, **kw: math.median(data, axis=axis),
        "sd": lambda data, axis=None, **kw: math.sd(data, axis=axis),
        "count": lambda data, axis=None, **kw: arrays.count(data, axis=axis),
        "mean_keepnan": lambda data, axis=None, **kw: math.mean(data, ignore_nodata=False, axis=axis),
    }
    for k, (name, (t, intervals, labels)) in enumerate(temporal_cases().items()):
        cube = _cube(t, 100 + k)
        store[f"at/{name}/in"] = cube.array
        doc["aggregate_temporal"][name] = {"t": [str(x) for x in np.asarray(t).astype("datetime64[ns]")],
                                           "intervals": intervals, "labels_in": labels}
        for rname, red in reducers.items():
            out = agg.aggregate_temporal(data=copy.deepcopy(cube), intervals=intervals, reducer=red,
                                         labels=labels)
            assert out.dims == cube.dims, out.dims
            store[f"at/{name}/{rname}"] = out.array
            doc["aggregate_temporal"][name]["labels_out"] = [str(x) for x in out.coords["t"]]
    # overlapping intervals raise (cubes/aggregate.py:75-78)
    try:
        agg.aggregate_temporal(data=_cube(temporal_cases()["custom_labels"][0], 1),
                               intervals=[["2021-01-01", "2021-02-01"], ["2021-01-15", "2021-03-01"]],
                               reducer=reducers["mean"])
        doc["overlap_error"] = None
    except Exception as e:
        doc["overlap_error"] = type(e).__name__
    for k, (name, (t, period)) in enumerate(period_cases().items()):
        cube = _cube(t, 200 + k)
        key = name.replace("/", "__")
        store[f"atp/{key}/in"] = cube.array
        entry = {"t": [str(x) for x in np.asarray(t).astype("datetime64[ns]")], "period": period}
        try:
            for rname in ("mean", "max", "count", "median"):
                out = agg.aggregate_temporal_period(data=copy.deepcopy(cube), reducer=reducers[rname], period=period)
                store[f"atp/{key}/{rname}"] = out.array
                entry["labels_out"] = [str(x) for x in out.coords["t"]]
        except Exception as e:
            entry["error"] = type(e).__name__
        doc["aggregate_temporal_period"][name] = entry
    with open(os.path.join(HERE, "reference_cubes.json"), "w") as f:
        json.dump(doc, f, indent=0, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "reference_cubes.npz"), **store)
    n_err = sum(1 for r in doc["get_intervals"] if "error" in r)
    print(f"get_intervals: {len(doc['get_intervals'])} cases ({n_err} raise); aggregate_temporal: "
          f"{len(doc['aggregate_temporal'])}; periods: {len(doc['aggregate_temporal_period'])}; arrays: {len(store)}")
    for name, e in doc["aggregate_temporal_period"].items():
        print(" ", name, e.get("error") or len(e["labels_out"]))


if __name__ == "__main__":
    main()
