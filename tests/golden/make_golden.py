"""Generate golden vectors for the oracle by running the REFERENCE's own functions.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference package cannot be imported normally (its ``__init__`` needs
rioxarray/odc/geopandas, and ``math.py`` imports dask, ``cubes/utils.py``
imports xarray).  The array-level reducers are, however, plain numpy
functions, so this script loads the *unmodified* reference source files
``exceptions.py``, ``utils.py``, ``cubes/utils.py``, ``comparison.py``,
``math.py``, ``arrays.py`` and ``logic.py`` one by one under their real module names, with
inert stand-ins for the absent third-party modules (``dask``, ``xarray``,
``openeo_pg_parser_networkx``).  None of the stand-ins is on the numpy code
path the reducers take for ndarray inputs.

Outputs (committed):
  tests/golden/spec_examples.json   every ``examples`` entry of the hot-path process specs
  tests/golden/reference_reducers.npz   reference outputs on seeded inputs
  tests/golden/reference_binary.npz     binary elementwise processes (math.py:98-115, comparison.py:45-127,
                                        logic.py:11-40) on operand pairs incl. broadcasting and scalars
"""

import importlib.util
import json
import os
import sys
import types
import warnings

import numpy as np
import pandas as pd

REF = "/root/reference/openeo_processes_dask"
HERE = os.path.dirname(os.path.abspath(__file__))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load_reference_modules():
    class _Never:  # isinstance(x, _Never) is always False for ndarrays
        pass

    dask = _stub("dask")
    dask.array = _stub("dask.array", Array=_Never, from_delayed=None, map_blocks=None)
    dask.delayed = _stub("dask.delayed", delayed=lambda f: f)
    xr = _stub("xarray", DataArray=_Never)
    xr.core = _stub("xarray.core")
    xr.core.duck_array_ops = _stub("xarray.core.duck_array_ops", isnull=pd.isnull)
    _stub("openeo_pg_parser_networkx")
    _stub("openeo_pg_parser_networkx.pg_schema", DateTime=_Never, ParameterReference=_Never)

    base = "openeo_processes_dask"
    for pkg in (base, base + ".process_implementations", base + ".process_implementations.cubes"):
        m = types.ModuleType(pkg)
        m.__path__ = []
        sys.modules[pkg] = m

    def load(rel, modname):
        spec = importlib.util.spec_from_file_location(modname, os.path.join(REF, rel))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    pi = base + ".process_implementations"
    load("process_implementations/exceptions.py", pi + ".exceptions")
    load("process_implementations/utils.py", pi + ".utils")
    load("process_implementations/cubes/utils.py", pi + ".cubes.utils")
    load("process_implementations/comparison.py", pi + ".comparison")
    math = load("process_implementations/math.py", pi + ".math")
    arrays = load("process_implementations/arrays.py", pi + ".arrays")
    logic = load("process_implementations/logic.py", pi + ".logic")
    return math, arrays, logic


def spec_examples():
    out = {}
    names = ["mean", "median", "sd", "variance", "sum", "min", "max", "count", "quantiles",
             "product", "extrema"]
    for n in names:
        with open(os.path.join(REF, "specs/openeo-processes", n + ".json")) as f:
            out[n] = json.load(f)["examples"]
    return out


def seeded_inputs():
    """Small cubes covering the edge cases of the path (NaN, inf, all-NaN, ties, 1 valid)."""
    rng = np.random.default_rng(42)  # the reference fixture's seed, tests/conftest.py:33
    cases = {}
    a = rng.integers(-100, 100, size=(20, 3, 6, 7)).astype(np.float32)  # ties, like the fixture
    cases["int_valued_f32"] = a
    b = rng.normal(3000, 1000, size=(37, 2, 5, 9)).astype(np.float32)
    b[rng.random(b.shape) < 0.3] = np.nan
    b[:, 0, 0, 0] = np.nan  # all-NaN column
    b[1:, 0, 0, 1] = np.nan  # one valid sample
    b[2:, 0, 0, 2] = np.nan  # two valid samples
    b[3, 1, 2, 3] = np.inf
    b[5, 1, 2, 3] = -np.inf
    b[4, 1, 2, 4] = np.inf
    cases["nan30_f32"] = b
    c = rng.normal(0, 1, size=(16, 4, 4, 4)).astype(np.float64)
    c[rng.random(c.shape) < 0.4] = np.nan
    cases["nan40_f64"] = c
    d = rng.integers(0, 10000, size=(9, 2, 4, 5)).astype(np.float32)
    cases["even_odd_f32"] = d
    # truth values for _any / _all / first / last: zeros, ones, NaN, inf; all-NaN, all-zero and all-one columns
    e = rng.choice(np.array([0.0, 1.0, 2.5, np.nan, np.inf], dtype=np.float32), size=(7, 3, 5, 6),
                   p=[0.35, 0.3, 0.1, 0.2, 0.05])
    e[:, 0, 0, 0] = np.nan
    e[:, 0, 0, 1] = 0.0
    e[:, 0, 0, 2] = 1.0
    e[:, 0, 0, 3] = np.array([np.nan, 0, 0, 0, 0, 0, 0], dtype=np.float32)
    e[:, 0, 0, 4] = np.array([1, 1, 1, np.nan, 1, 1, np.nan], dtype=np.float32)
    cases["truth_f32"] = e
    return cases


def main():
    math, arrays, logic = load_reference_modules()
    with open(os.path.join(HERE, "spec_examples.json"), "w") as f:
        json.dump(spec_examples(), f, indent=1, sort_keys=True)

    fns = {
        "min": math._min, "max": math._max, "median": math.median, "mean": math.mean,
        "sd": math.sd, "variance": math.variance, "sum": math._sum, "product": math.product,
    }
    store = {}
    warnings.simplefilter("ignore", RuntimeWarning)
    for cname, x in seeded_inputs().items():
        store[f"in/{cname}"] = x
        for axis in range(x.ndim):
            for ign in (True, False):
                for rname, fn in fns.items():
                    store[f"out/{cname}/{rname}/axis{axis}/ign{int(ign)}"] = np.asarray(
                        fn(x.copy(), ignore_nodata=ign, axis=axis))
                store[f"out/{cname}/quantiles3/axis{axis}/ign{int(ign)}"] = np.asarray(
                    math.quantiles(x.copy(), probabilities=[0.1, 0.5, 0.9], ignore_nodata=ign, axis=axis))
                store[f"out/{cname}/quantiles_p4/axis{axis}/ign{int(ign)}"] = np.asarray(
                    math.quantiles(x.copy(), probabilities=4, ignore_nodata=ign, axis=axis))
                # the remaining reducer-category processes (arrays.py:461-495, logic.py:58-87)
                store[f"out/{cname}/first/axis{axis}/ign{int(ign)}"] = np.asarray(
                    arrays.first(x.copy(), ignore_nodata=ign, axis=axis))
                store[f"out/{cname}/last/axis{axis}/ign{int(ign)}"] = np.asarray(
                    arrays.last(x.copy(), ignore_nodata=ign, axis=axis))
                store[f"out/{cname}/any/axis{axis}/ign{int(ign)}"] = np.asarray(
                    logic._any(x.copy(), ignore_nodata=ign, axis=axis))
                store[f"out/{cname}/all/axis{axis}/ign{int(ign)}"] = np.asarray(
                    logic._all(x.copy(), ignore_nodata=ign, axis=axis))
                if cname in ("nan30_f32", "nan40_f64", "truth_f32"):   # cumulative processes, math.py:125-174
                    for rname, fn in (("cumsum", math.cumsum), ("cumproduct", math.cumproduct),
                                      ("cummin", math.cummin), ("cummax", math.cummax)):
                        store[f"out/{cname}/{rname}/axis{axis}/ign{int(ign)}"] = np.asarray(
                            fn(x.copy(), ignore_nodata=ign, axis=axis))
            if cname in ("nan30_f32", "truth_f32"):   # sort along a dimension, arrays.py:560-584
                # called the way apply_dimension calls it (cubes/apply.py:66-84): the dimension moved to the
                # LAST axis.  (Called directly with another axis, nodata=False assigns through a flattened
                # boolean mask and mixes the values of different slices.)
                for asc in (True, False):
                    for nd in (True, False):
                        moved = np.ascontiguousarray(np.moveaxis(x, axis, -1))
                        res = np.asarray(arrays.sort(moved, asc=asc, nodata=nd, axis=moved.ndim - 1))
                        store[f"out/{cname}/sort_asc{int(asc)}_nodata{int(nd)}/axis{axis}/ign1"] = np.ascontiguousarray(
                            np.moveaxis(res, -1, axis))
            if cname in ("nan30_f32", "nan40_f64", "truth_f32") and axis >= 1:
                # array_interpolate_linear, arrays.py:405-458 (`if axis:` -- axis 0 of an N-d array is not
                # usable in the reference; apply_dimension always passes the last axis)
                store[f"out/{cname}/interp/axis{axis}/ign1"] = np.asarray(
                    arrays.array_interpolate_linear(x.copy(), axis=axis))
            store[f"out/{cname}/count/axis{axis}/ign1"] = np.asarray(arrays.count(x.copy(), axis=axis))
            store[f"out/{cname}/count_true/axis{axis}/ign1"] = np.asarray(
                arrays.count(x.copy(), condition=True, axis=axis))
            store[f"out/{cname}/element1/axis{axis}/ign1"] = np.asarray(
                arrays.array_element(x.copy(), index=1, axis=axis))
        if cname in ("nan30_f32", "nan40_f64", "int_valued_f32"):   # elementwise: math.py:292-298, 332-338
            store[f"out/{cname}/lsr_0_5000"] = np.asarray(math.linear_scale_range(x.copy(), 0, 5000))
            store[f"out/{cname}/lsr_swapped_255"] = np.asarray(math.linear_scale_range(x.copy(), 5000, 0, 0, 255))
            store[f"out/{cname}/lsr_float"] = np.asarray(math.linear_scale_range(x.copy(), -0.3, 2500.7, 0.1, 0.9))
            store[f"out/{cname}/clip_1000_4000"] = np.asarray(math.clip(x.copy(), 1000, 4000))
            store[f"out/{cname}/clip_float"] = np.asarray(math.clip(x.copy(), -0.25, 0.75))
        store[f"out/{cname}/nd01"] = np.asarray(
            math.normalized_difference(np.take(x, 0, axis=1), np.take(x, 1, axis=1)))
    # reference-test vectors (tests/test_math.py:9-39, 78-87)
    v = np.array([2, 4, 4, 4, 5, 5, 7, 9])
    store["out/test_math/quantiles_1"] = math.quantiles(data=v, probabilities=[0.005, 0.01, 0.02, 0.05, 0.1, 0.5])
    store["out/test_math/quantiles_2"] = math.quantiles(data=v, probabilities=4)
    store["out/test_math/quantiles_3"] = math.quantiles(data=np.array([-1, -0.5, np.nan, 1]), probabilities=[2])
    store["out/test_math/quantiles_4"] = math.quantiles(data=np.array([-1, -0.5, np.nan, 1]), q=4, ignore_nodata=False)
    store["out/test_math/quantiles_5"] = math.quantiles(data=np.array([]), probabilities=[0.1, 0.5])
    store["out/test_math/nd_uint16"] = np.asarray(
        math.normalized_difference(np.array([1000], dtype=np.uint16), np.array([2000], dtype=np.uint16)))
    np.savez_compressed(os.path.join(HERE, "reference_reducers.npz"), **store)
    print("wrote", len(store), "arrays")


def binary_cases():
    """name -> (x, y): equal shapes, a (y,x) plane / a (t,1,1,x) slab broadcast over a cube, scalars, integers."""
    cubes = seeded_inputs()
    b = cubes["nan30_f32"]
    c = cubes["nan40_f64"]
    e = cubes["truth_f32"]
    rng = np.random.default_rng(7)
    u = rng.integers(0, 65536, size=(5, 6, 7)).astype(np.uint16)
    w = rng.integers(0, 65536, size=(5, 6, 7)).astype(np.uint16)
    i = rng.integers(-2**31, 2**31 - 1, size=(4, 33)).astype(np.int32)
    j = rng.integers(-2**31, 2**31 - 1, size=(4, 33)).astype(np.int32)
    z = b[:, 0].copy()
    z[::3] = 0.0
    return {
        "f32_same": (b[:, 0], b[:, 1]),
        "f32_zero_div": (b[:, 1], z),
        "f32_plane": (b, b[3, 1]),
        "f32_slab": (b[:, :, :, :], b[:, :1, :1, :]),
        "f32_scalar_r": (b, 2.5),
        "f32_scalar_l": (1000, b),
        "f64_same": (c[:, 0], c[:, 2]),
        "f64_f32_mixed": (c[:, 0, :, :], b[:16, 0, :4, :4]),
        "truth": (e[:, 0], e[:, 1]),
        "truth_plane": (e, e[2, 2]),
        "u16_same": (u, w),
        "u16_scalar": (u, 3),
        "i32_same": (i, j),
        "view_strided": (b[::2, 0, :, ::2][:18, :, :4], b[1::2, 1, :, 1::2]),
    }


def main_binary():
    math, arrays, logic = load_reference_modules()
    comparison = sys.modules["openeo_processes_dask.process_implementations.comparison"]
    fns = {"add": math.add, "subtract": math.subtract, "multiply": math.multiply, "divide": math.divide,
           "eq": comparison.eq, "neq": comparison.neq, "gt": comparison.gt, "gte": comparison.gte,
           "lt": comparison.lt, "lte": comparison.lte, "and": logic._and, "or": logic._or, "xor": logic.xor}
    store = {}
    warnings.simplefilter("ignore", RuntimeWarning)
    for cname, (x, y) in binary_cases().items():
        store[f"in/{cname}/x"] = np.asarray(x)
        store[f"in/{cname}/y"] = np.asarray(y)
        for fname, fn in fns.items():
            xx = x.copy() if hasattr(x, "copy") else x
            yy = y.copy() if hasattr(y, "copy") else y
            store[f"out/{cname}/{fname}"] = np.asarray(fn(xx, yy))
    # unary math processes with exact results (math.py:212-222, 305-315)
    cubes = seeded_inputs()
    special = np.array([-0.0, 0.0, -0.5, 0.5, -1.5, 2.5, np.nan, np.inf, -np.inf, 1e-40, -1e-40, 3.0e38, 16777217.0])
    rng = np.random.default_rng(11)
    unary_in = {"nan30_f32": cubes["nan30_f32"], "nan40_f64": cubes["nan40_f64"], "truth_f32": cubes["truth_f32"],
                "special_f32": special.astype(np.float32), "special_f64": special,
                "i16": rng.integers(-32767, 32767, size=(7, 9)).astype(np.int16),
                "u16": rng.integers(0, 65535, size=(7, 9)).astype(np.uint16),
                "i32": rng.integers(-2**31 + 1, 2**31 - 1, size=(5, 8)).astype(np.int32)}
    ufns = {"absolute": math.absolute, "sgn": math.sgn, "sqrt": math.sqrt, "floor": math.floor, "ceil": math.ceil,
            "int": math._int}
    for cname, x in unary_in.items():
        store[f"uin/{cname}"] = x
        for fname, fn in ufns.items():
            store[f"uout/{cname}/{fname}"] = np.asarray(fn(x.copy()))
    np.savez_compressed(os.path.join(HERE, "reference_binary.npz"), **store)
    print("wrote", len(store), "arrays (binary)")


if __name__ == "__main__":
    if "--binary" in sys.argv:
        main_binary()
    else:
        main()
