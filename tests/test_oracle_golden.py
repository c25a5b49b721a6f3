"""Pin the oracle against the reference's golden vectors (CPU, no GPU needed).

1. every ``examples`` entry of the hot-path openEO process specs the reference ships
   (``specs/openeo-processes/{mean,median,sd,variance,sum,min,max,count,quantiles,product,extrema}.json``)
2. outputs of the reference's own ``math.py`` / ``arrays.py`` functions on seeded inputs
   (``tests/golden/make_golden.py``)
3. the value vectors of the reference's tests (``tests/test_math.py:9-39,78-87``).
"""

import warnings

import numpy as np
import pytest

from oracle import reducers as R


def _arr(data):
    return np.array([np.nan if v is None else v for v in data], dtype=float)


SPEC_NAMES = ["mean", "median", "sd", "variance", "sum", "min", "max", "product"]


@pytest.mark.parametrize("name", SPEC_NAMES)
def test_spec_examples_reducers(spec_examples, name):
    for ex in spec_examples[name]:
        args = ex["arguments"]
        data = _arr(args["data"])
        if name in ("min", "max") and len(data) == 0:
            # np.nanmin([]) raises ValueError in numpy, hence in the reference (math.py:120)
            with pytest.raises(ValueError):
                R.BY_NAME[name](data)
            continue
        got = R.BY_NAME[name](data, ignore_nodata=args.get("ignore_nodata", True))
        want = ex["returns"]
        if want is None:
            # the reference (numpy) returns 0.0 / 1.0 for nansum/nanprod of all-NaN, NaN otherwise;
            # the spec's "null" for an EMPTY array is NaN in every case.
            if name in ("sum", "product") and len(data) and args.get("ignore_nodata", True):
                continue
            assert np.isnan(got), (name, ex)
        else:
            np.testing.assert_allclose(got, want, rtol=1e-12)


def test_spec_examples_quantiles(spec_examples):
    for ex in spec_examples["quantiles"]:
        args = ex["arguments"]
        got = R.quantiles(_arr(args["data"]), probabilities=args["probabilities"],
                          ignore_nodata=args.get("ignore_nodata", True))
        want = _arr(ex["returns"])
        assert len(got) == len(want)
        np.testing.assert_allclose(np.round(got, 2), want, equal_nan=True)


def test_spec_examples_count(spec_examples):
    for ex in spec_examples["count"]:
        args = ex["arguments"]
        if any(isinstance(v, str) for v in args["data"]):
            continue  # string arrays are outside the raster path
        cond = args.get("condition")
        if isinstance(cond, dict):
            cond = lambda x, y=2: np.asarray(x) > y  # the spec's gt(element, 2) child graph
        data = _arr(args["data"]) if len(args["data"]) else np.array([])
        assert R.count(data, condition=cond) == ex["returns"]


def test_spec_examples_extrema(spec_examples):
    for ex in spec_examples["extrema"]:
        args = ex["arguments"]
        data = _arr(args["data"])
        if len(data) == 0:
            continue  # np.nanmin of an empty array raises in numpy; the reference does too
        got = R.extrema(data, ignore_nodata=args.get("ignore_nodata", True))
        np.testing.assert_allclose(got, _arr(ex["returns"]), equal_nan=True)


def _cases(golden):
    return sorted({k.split("/")[1] for k in golden.files if k.startswith("in/")})


def test_oracle_equals_reference_outputs(golden):
    """Bit-for-bit equality with what the reference's functions returned."""
    checked = 0
    for cname in _cases(golden):
        x = golden[f"in/{cname}"]
        for axis in range(x.ndim):
            for ign in (True, False):
                for rname in ("min", "max", "median", "mean", "sd", "variance", "sum", "product"):
                    want = golden[f"out/{cname}/{rname}/axis{axis}/ign{int(ign)}"]
                    got = R.BY_NAME[rname](x.copy(), ignore_nodata=ign, axis=axis)
                    assert got.dtype == want.dtype
                    np.testing.assert_array_equal(got, want)
                    checked += 1
                want = golden[f"out/{cname}/quantiles3/axis{axis}/ign{int(ign)}"]
                got = R.quantiles(x.copy(), probabilities=[0.1, 0.5, 0.9], ignore_nodata=ign, axis=axis)
                assert got.shape == want.shape and got.shape[axis] == 3
                np.testing.assert_array_equal(got, want)
                want = golden[f"out/{cname}/quantiles_p4/axis{axis}/ign{int(ign)}"]
                got = R.quantiles(x.copy(), probabilities=4, ignore_nodata=ign, axis=axis)
                np.testing.assert_array_equal(got, want)
            want = golden[f"out/{cname}/count/axis{axis}/ign1"]
            got = R.count(x, axis=axis)
            assert got.dtype == want.dtype == np.int64
            np.testing.assert_array_equal(got, want)
            np.testing.assert_array_equal(R.count(x, condition=True, axis=axis),
                                          golden[f"out/{cname}/count_true/axis{axis}/ign1"])
        np.testing.assert_array_equal(
            R.normalized_difference(np.take(x, 0, axis=1), np.take(x, 1, axis=1)),
            golden[f"out/{cname}/nd01"])
    assert checked > 100


def test_reference_test_vectors(golden):
    v = np.array([2, 4, 4, 4, 5, 5, 7, 9])
    q1 = R.quantiles(v, probabilities=[0.005, 0.01, 0.02, 0.05, 0.1, 0.5])
    np.testing.assert_array_equal(q1, golden["out/test_math/quantiles_1"])
    assert [round(float(q), 2) for q in q1] == [2.07, 2.14, 2.28, 2.7, 3.4, 4.5]
    assert [round(float(q), 2) for q in R.quantiles(v, probabilities=4)] == [4, 4.5, 5.5]
    assert [round(float(q), 2) for q in R.quantiles(np.array([-1, -0.5, np.nan, 1]), probabilities=[2])] == [-0.5]
    q4 = R.quantiles(np.array([-1, -0.5, np.nan, 1]), q=4, ignore_nodata=False)
    assert len(q4) == 3 and np.all(np.isnan(q4))
    q5 = R.quantiles(np.array([]), probabilities=[0.1, 0.5])
    assert len(q5) == 2 and np.all(np.isnan(q5))
    assert R._sum([5, 1]) == 6 and R._sum([-2, 4, 2.5]) == 4.5
    assert np.isnan(R._sum([1, np.nan], ignore_nodata=False))
    assert R.normalized_difference(5, 3) == 0.25 and R.normalized_difference(1, 1) == 0
    # the uint16 wrap quirk
    got = R.normalized_difference(np.array([1000], dtype=np.uint16), np.array([2000], dtype=np.uint16))
    np.testing.assert_array_equal(got, golden["out/test_math/nd_uint16"])


def test_quirks():
    # quirk 3: np.arange(1/p, 1, 1/p) has p entries for p=3
    assert len(R.resolve_probabilities(3)) == 3 and len(R.resolve_probabilities(4)) == 3
    # quirk 4: all-NaN sum -> 0, product -> 1, others NaN
    a = np.full((3, 2), np.nan, dtype=np.float32)
    assert (R._sum(a, axis=0) == 0).all() and (R.product(a, axis=0) == 1).all()
    assert np.isnan(R.mean(a, axis=0)).all() and np.isnan(R._max(a, axis=0)).all()
    # quirk 5: one valid sample -> NaN sd
    b = np.array([[1.0], [np.nan]], dtype=np.float32)
    assert np.isnan(R.sd(b, axis=0)).all()
    # quirk 6: count excludes +-inf and NaN, int64
    c = np.array([1, np.inf, -np.inf, np.nan, 2], dtype=np.float32)
    assert R.count(c) == 2 and R.count(c, condition=True) == 5


def test_oracle_first_last_any_all_equal_reference_outputs(golden):
    """first / last (arrays.py:461-495), _any / _all (logic.py:58-87), array_element."""
    for cname in _cases(golden):
        x = golden[f"in/{cname}"]
        for axis in range(x.ndim):
            for ign in (True, False):
                for rname, fn in (("first", R.first), ("last", R.last), ("any", R._any), ("all", R._all)):
                    want = golden[f"out/{cname}/{rname}/axis{axis}/ign{int(ign)}"]
                    got = np.asarray(fn(x.copy(), ignore_nodata=ign, axis=axis))
                    assert got.dtype == want.dtype, (cname, rname, axis, ign)
                    np.testing.assert_array_equal(got, want, err_msg=f"{cname}/{rname}/axis{axis}/ign{ign}")
            np.testing.assert_array_equal(R.array_element(x.copy(), index=1, axis=axis),
                                          golden[f"out/{cname}/element1/axis{axis}/ign1"])
    # spec examples (specs/openeo-processes/first.json, last.json, any.json, all.json)
    assert R.first(np.array([1.0, 0, 3, 2])) == 1 and np.isnan(R.first(np.array([np.nan, 7.0]), ignore_nodata=False))
    assert R.first(np.array([np.nan, 7.0])) == 7 and R.last(np.array([0.0, 1, np.nan])) == 1
    assert np.isnan(R.last(np.array([0.0, 1, np.nan]), ignore_nodata=False))
    assert R._any(np.array([0.0, np.nan])) == False and np.isnan(R._any(np.array([0.0, np.nan]), ignore_nodata=False))  # noqa: E712
    assert R._all(np.array([1.0, np.nan])) == True and np.isnan(R._all(np.array([1.0, np.nan]), ignore_nodata=False))  # noqa: E712
    assert R._all(np.array([0.0, np.nan]), ignore_nodata=False) == 0
    assert np.isnan(R.array_element(np.arange(6.0).reshape(2, 3), index=5, axis=0, return_nodata=True)).all()
    with pytest.raises(IndexError):
        R.array_element(np.arange(3.0), index=5, axis=0)
    assert R.array_element(np.arange(3.0), label="b", axis=0, dim_labels=np.array(["a", "b", "c"])) == 1.0


def test_oracle_cumulative_equal_reference_outputs(golden):
    """cumsum / cumproduct / cummin / cummax (math.py:125-174)."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2] in ("cumsum", "cumproduct", "cummin", "cummax"):
            x = golden["in/" + parts[1]]
            got = getattr(R, parts[2])(x.copy(), ignore_nodata=bool(int(parts[4][3:])), axis=int(parts[3][4:]))
            assert got.dtype == golden[key].dtype
            np.testing.assert_array_equal(got, golden[key], err_msg=key)
            n += 1
    assert n == 96
    # spec examples (specs/openeo-processes/cumsum.json etc.)
    np.testing.assert_array_equal(R.cumsum(np.array([1.0, 3, 5, 3, 1])), [1, 4, 9, 12, 13])
    np.testing.assert_array_equal(R.cumsum(np.array([1.0, 3, np.nan, 3, 1])), [1, 4, np.nan, 7, 8])
    np.testing.assert_array_equal(R.cumsum(np.array([1.0, 3, np.nan, 3, 1]), ignore_nodata=False),
                                  [1, 4, np.nan, np.nan, np.nan])
    np.testing.assert_array_equal(R.cumproduct(np.array([1.0, 2, 3, np.nan, 3, 1])), [1, 2, 6, np.nan, 18, 18])
    np.testing.assert_array_equal(R.cummin(np.array([5.0, 3, np.nan, 1, 5])), [5, 3, np.nan, 1, 1])
    np.testing.assert_array_equal(R.cummax(np.array([1.0, 3, np.nan, 5, 1]), ignore_nodata=False),
                                  [1, 3, np.nan, np.nan, np.nan])


def test_oracle_sort_equals_reference_outputs(golden):
    """sort along a dimension (arrays.py:560-584), generated the way apply_dimension calls it."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2].startswith("sort_"):
            x = golden["in/" + parts[1]]
            got = R.sort(x.copy(), asc=parts[2][8] == "1", nodata=parts[2][-1] == "1", axis=int(parts[3][4:]))
            np.testing.assert_array_equal(got, golden[key], err_msg=key)
            n += 1
    assert n == 32
    # spec examples (specs/openeo-processes/sort.json)
    v = np.array([6, -1, 2, np.nan, 7, 4, np.nan, 8, 3, 9, 9])
    np.testing.assert_array_equal(R.sort(v), [-1, 2, 3, 4, 6, 7, 8, 9, 9])
    np.testing.assert_array_equal(R.sort(v, asc=False, nodata=True), [9, 9, 8, 7, 6, 4, 3, 2, -1, np.nan, np.nan])
    np.testing.assert_array_equal(R.sort(v, nodata=False), [np.nan, np.nan, -1, 2, 3, 4, 6, 7, 8, 9, 9])


def test_oracle_interpolate_equals_reference_outputs(golden):
    """array_interpolate_linear (arrays.py:405-458) along axes >= 1 (the reference's `if axis:`)."""
    n = 0
    for key in golden.files:
        parts = key.split("/")
        if len(parts) == 5 and parts[2] == "interp":
            x = golden["in/" + parts[1]]
            got = R.array_interpolate_linear(x.copy(), axis=int(parts[3][4:]))
            assert got.dtype == golden[key].dtype
            np.testing.assert_array_equal(got, golden[key], err_msg=key)
            n += 1
    assert n == 9
    # spec examples (specs/openeo-processes/array_interpolate_linear.json)
    np.testing.assert_array_equal(R.array_interpolate_linear(np.array([np.nan, 1, np.nan, 6, np.nan, -8])),
                                  [np.nan, 1, 3.5, 6, -1, -8])
    np.testing.assert_array_equal(R.array_interpolate_linear(np.array([np.nan, 1, np.nan, np.nan])),
                                  [np.nan, 1, np.nan, np.nan])


def test_oracle_elementwise_equal_reference_outputs(golden):
    """clip / linear_scale_range (math.py:292-298, 332-338) incl. the reference-test vectors tests/test_math.py:90-109."""
    cases = (("lsr_0_5000", lambda x: R.linear_scale_range(x, 0, 5000)),
             ("lsr_swapped_255", lambda x: R.linear_scale_range(x, 5000, 0, 0, 255)),
             ("lsr_float", lambda x: R.linear_scale_range(x, -0.3, 2500.7, 0.1, 0.9)),
             ("clip_1000_4000", lambda x: R.clip(x, 1000, 4000)), ("clip_float", lambda x: R.clip(x, -0.25, 0.75)))
    for cname in ("nan30_f32", "nan40_f64", "int_valued_f32"):
        x = golden["in/" + cname]
        for key, fn in cases:
            want = golden[f"out/{cname}/{key}"]
            got = fn(x.copy())
            assert got.dtype == want.dtype
            np.testing.assert_array_equal(got, want, err_msg=f"{cname}/{key}")
    np.testing.assert_array_equal(R.linear_scale_range(np.array([5, 0]), 0, 5), [1, 0])
    np.testing.assert_array_equal(R.linear_scale_range(np.array([5, 0]), 5, 0), [0, 1])
    np.testing.assert_array_equal(R.clip(np.array([5, 0]), 0, 2), [2, 0])


def _scalar_or_array(a):
    # the generator stored python scalars as 0-d arrays; weak-scalar promotion needs the python value back
    return a.item() if a.ndim == 0 else a


def test_oracle_binary_equal_reference_outputs(golden_binary):
    """add / subtract / multiply / divide (math.py:98-115), eq .. lte (comparison.py:45-127), and / or / xor
    (logic.py:11-40): outputs of the reference's own functions on equal shapes, broadcast planes, scalars, integers."""
    g = golden_binary
    cases = sorted({k.split("/")[1] for k in g.files if k.startswith("in/")})
    assert len(cases) == 14
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", RuntimeWarning)
        for c in cases:
            x, y = _scalar_or_array(g[f"in/{c}/x"]), _scalar_or_array(g[f"in/{c}/y"])
            for name, fn in R.BINARY.items():
                want = g[f"out/{c}/{name}"]
                got = np.asarray(fn(x, y))
                assert got.dtype == want.dtype, (c, name, got.dtype, want.dtype)
                np.testing.assert_array_equal(got, want, err_msg=f"{c}/{name}")
    # three-valued logic (specs/openeo-processes/and.json, or.json, xor.json examples)
    nan = np.nan
    np.testing.assert_array_equal(R._and(np.array([1.0, 1, 0, 0, nan, nan]), np.array([1.0, 0, 0, nan, 1, 0])), [1, 0, 0, 0, nan, 0])
    np.testing.assert_array_equal(R._or(np.array([1.0, 0, 0, nan, nan]), np.array([0.0, 0, nan, 1, nan])), [1, 0, nan, 1, nan])
    np.testing.assert_array_equal(R.xor(np.array([1.0, 1, 0, nan]), np.array([1.0, 0, 0, 1])), [0, 1, 0, nan])


def _bits_equal(got, want, msg):
    """values, NaN masks AND the sign of zeros (np.sign(-0.0) is +0.0, ceil(-0.5) is -0.0)"""
    assert got.dtype == want.dtype and got.shape == want.shape, (msg, got.dtype, want.dtype)
    np.testing.assert_array_equal(got, want, err_msg=msg)
    if got.dtype.kind == "f":
        np.testing.assert_array_equal(np.signbit(got) & ~np.isnan(got), np.signbit(want) & ~np.isnan(want), err_msg=msg + " (sign)")


def test_oracle_unary_equal_reference_outputs(golden_binary):
    """absolute / sgn / sqrt / floor / ceil / int (math.py:212-222, 305-315) vs the reference's own outputs."""
    g = golden_binary
    cases = sorted(k.split("/")[1] for k in g.files if k.startswith("uin/"))
    assert len(cases) == 8
    for c in cases:
        for name, fn in R.UNARY.items():
            _bits_equal(np.asarray(fn(g[f"uin/{c}"].copy())), g[f"uout/{c}/{name}"], f"{c}/{name}")
