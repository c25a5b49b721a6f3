"""Array-level reducers of the reference, restated on numpy (TEST INFRASTRUCTURE).

Every function follows one reference one-liner; the numpy call is the same
one the reference makes, so eager-numpy results are identical by construction
(and are additionally pinned by ``tests/golden``).  Citations are into
``/root/reference/openeo_processes_dask/process_implementations/``.
"""

from __future__ import annotations

import warnings

import numpy as np

# name -> (nan-aware numpy function, plain numpy function, extra kwargs)
#   _min      math.py:118-122      np.nanmin / np.min
#   _max      math.py:177-181      np.nanmax / np.max
#   median    math.py:184-188      np.nanmedian / np.median
#   mean      math.py:191-195      np.nanmean / np.mean
#   sd        math.py:198-202      np.nanstd / np.std, ddof=1
#   variance  math.py:205-209      np.nanvar / np.var, ddof=1
#   _sum      math.py:390-398      np.nansum / np.sum   (len(data)==0 -> nan)
#   product   math.py:401-409      np.nanprod / np.prod (len(data)==0 -> nan)
_TABLE = {
    "min": (np.nanmin, np.min, {}),
    "max": (np.nanmax, np.max, {}),
    "median": (np.nanmedian, np.median, {}),
    "mean": (np.nanmean, np.mean, {}),
    "sd": (np.nanstd, np.std, {"ddof": 1}),
    "variance": (np.nanvar, np.var, {"ddof": 1}),
    "sum": (np.nansum, np.sum, {}),
    "product": (np.nanprod, np.prod, {}),
}

REDUCER_NAMES = tuple(_TABLE) + ("count", "quantiles")


def reduce(name, data, ignore_nodata=True, axis=None, keepdims=False):
    """Apply reference reducer ``name`` the way ``math.py`` does."""
    nan_fn, plain_fn, extra = _TABLE[name]
    data = np.asarray(data)
    if name in ("sum", "product") and len(data) == 0:  # math.py:391-392, 402-403
        return np.nan
    fn = nan_fn if ignore_nodata else plain_fn
    with warnings.catch_warnings():
        # numpy emits (does not raise) "Mean of empty slice", "All-NaN slice",
        # "Degrees of freedom <= 0"; the reference lets them through.
        warnings.simplefilter("ignore", RuntimeWarning)
        return fn(data, axis=axis, keepdims=keepdims, **extra)


def _min(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("min", data, ignore_nodata, axis, keepdims)


def _max(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("max", data, ignore_nodata, axis, keepdims)


def median(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("median", data, ignore_nodata, axis, keepdims)


def mean(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("mean", data, ignore_nodata, axis, keepdims)


def sd(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("sd", data, ignore_nodata, axis, keepdims)


def variance(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("variance", data, ignore_nodata, axis, keepdims)


def _sum(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("sum", data, ignore_nodata, axis, keepdims)


def product(data, ignore_nodata=True, axis=None, keepdims=False):
    return reduce("product", data, ignore_nodata, axis, keepdims)


def resolve_probabilities(probabilities=None, q=None):
    """math.py:344-371 -- validation and int -> arange expansion.

    Raises ValueError subclasses by *name* only in the product; here plain
    ValueError keeps the oracle dependency-free.  Note the reference's
    ``np.arange(1/p, 1, 1/p)`` yields p entries (last one 1.0) for
    p in {3, 6, 7, 14, ...} because of floating-point accumulation; evaluating
    the same expression reproduces that.
    """
    if probabilities is not None and q is not None:
        raise ValueError("QuantilesParameterConflict")
    if probabilities is None and q is None:
        raise ValueError("QuantilesParameterMissing")
    if isinstance(probabilities, int) and not isinstance(probabilities, bool):
        probabilities = np.arange(1.0 / probabilities, 1, 1.0 / probabilities)
    elif (
        isinstance(probabilities, list)
        and len(probabilities) == 1
        and isinstance(probabilities[0], int)
    ):
        probabilities = np.arange(1.0 / probabilities[0], 1, 1.0 / probabilities[0])
    elif isinstance(probabilities, list):
        probabilities = np.array(probabilities)
    if q is not None:
        probabilities = np.arange(1.0 / q, 1, 1.0 / q)
    return probabilities


def quantiles(data, probabilities=None, q=None, ignore_nodata=True, axis=None, keepdims=False):
    """math.py:341-387.  The q-axis ends at position ``axis`` (``if axis:`` quirk)."""
    probabilities = resolve_probabilities(probabilities, q)
    data = np.asarray(data)
    if data.size == 0:  # math.py:373-374
        return np.array([np.nan] * len(probabilities))
    fn = np.nanquantile if ignore_nodata else np.quantile
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", RuntimeWarning)
        result = fn(data, q=probabilities, method="linear", axis=axis, keepdims=keepdims)
    if axis:  # math.py:384-385 (falsy for axis 0, where the q-axis already leads)
        result = np.moveaxis(result, 0, axis)
    return result


def is_infinite(x):
    """comparison.py:24-28."""
    x = np.asarray(x)
    if np.issubdtype(x.dtype, np.number):
        return np.isinf(x)
    return np.zeros(x.shape, dtype=bool)


def is_valid(x):
    """comparison.py:31-33 with cubes/utils.py:17-25 (notnull == ~isnan for numeric arrays)."""
    import pandas as pd

    x = np.asarray(x)
    return np.logical_and(~pd.isnull(x), np.logical_not(is_infinite(x)))


def count(data, condition=None, context=None, axis=None, keepdims=False):
    """arrays.py:587-606."""
    data = np.asarray(data)
    if condition is None:
        return np.nansum(is_valid(data), axis=axis, keepdims=keepdims)
    if condition is True:
        return np.nansum(np.ones_like(data), axis=axis, keepdims=keepdims)
    if callable(condition):
        context = dict(context or {})
        context.pop("x", None)
        return np.nansum(condition(x=data, **context), axis=axis, keepdims=keepdims)
    raise Exception("Could not count values as condition is not callable. ")


def normalized_difference(x, y):
    """math.py:412-414.  dtype-preserving for floats; unsigned ints wrap."""
    with np.errstate(divide="ignore", invalid="ignore"):
        return (x - y) / (x + y)


def extrema(data, ignore_nodata=True, axis=None, keepdims=False):
    """math.py:322-329 without the dask.delayed wrapper: array([min, max])."""
    data = np.asarray(data)
    return np.array(
        [_min(data, ignore_nodata, axis, keepdims), _max(data, ignore_nodata, axis, keepdims)]
    )


BY_NAME = {
    "min": _min,
    "max": _max,
    "median": median,
    "mean": mean,
    "sd": sd,
    "variance": variance,
    "sum": _sum,
    "product": product,
    "count": count,
    "quantiles": quantiles,
}


# ---- remaining reducer-category processes ----------------------------------------
def first(data, ignore_nodata=True, axis=None):
    """arrays.py:461-485: the first not-null element along ``axis`` (NaN when the slice has none);
    ``ignore_nodata=False``: element 0.  ``axis=None`` flattens first (arrays.py:471-473).

    Restated as one gather: index k of the first valid sample per slice (0 when there is none, which then
    reads the NaN at position 0 -- the same value the reference's fill loop leaves behind).

    QUIRK reproduced on purpose: the reference fills the gaps with ``np.nan_to_num(first_elem, True, data[i])``
    for i = 0..max(k) (arrays.py:478-480), and nan_to_num also turns +-inf into +-finfo.max.  So whenever
    plane 0 holds a NaN anywhere (the loop runs), an infinite result is clamped to +-max if it was already
    in place before the last iteration: k == 0, or k < max(k) over the whole array."""
    data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    if axis is None:
        data, axis = data.reshape(-1), 0
    if not ignore_nodata or data.dtype.kind not in "fc":
        return np.take(data, 0, axis=axis)
    where = np.argmax(~np.isnan(data), axis=axis)
    out = np.squeeze(np.take_along_axis(data, np.expand_dims(where, axis), axis), axis)
    if np.isnan(np.take(data, 0, axis=axis)).any():
        clamp = np.isinf(out) & ((where == 0) | (where < where.max()))
        out = np.where(clamp, np.copysign(np.finfo(data.dtype).max, out), out).astype(data.dtype)
    return out


def last(data, ignore_nodata=True, axis=None):
    """arrays.py:488-495: ``first`` of the array flipped along ``axis``."""
    data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    return first(np.flip(data, axis=axis), ignore_nodata=ignore_nodata, axis=axis)


def _any(data, ignore_nodata=True, axis=None):
    """logic.py:58-71: ``np.any(np.nan_to_num(data))``; with ``ignore_nodata=False`` NaN where the slice has
    a NaN and no true element (the result then is a float array: np.where(bool, bool, nan))."""
    data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    hit = np.any(np.nan_to_num(data), axis=axis)
    if ignore_nodata:
        return hit
    undecided = np.any(np.isnan(data), axis=axis) & ~hit
    return np.where(undecided, np.nan, hit)


def _all(data, ignore_nodata=True, axis=None):
    """logic.py:74-87: ``np.all(data)`` (NaN is truthy); with ``ignore_nodata=False`` NaN where the slice has
    a NaN and no false element."""
    data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    every = np.all(data, axis=axis)
    if ignore_nodata:
        return every
    undecided = np.any(np.isnan(data), axis=axis) & every
    return np.where(undecided, np.nan, every)


def array_element(data, index=None, label=None, return_nodata=False, axis=None, dim_labels=None):
    """arrays.py:66-122 for ndarray input: ``np.take(data, index, axis)``; labels resolved in ``dim_labels``;
    a missing element gives a NaN-filled float64 array with ``return_nodata`` or raises IndexError here
    (ArrayElementNotAvailable in the product)."""
    if (index is None) == (label is None):
        raise ValueError("exactly one of index / label")
    data = np.asarray(data)
    if label is not None:
        if dim_labels is None or len(dim_labels) == 0:
            raise ValueError("array is not labeled")
        (hits,) = np.where(np.asarray(dim_labels) == label)
        index = int(hits[0]) if len(hits) else None
    n = data.shape[0 if axis is None else axis] if axis is not None else data.size
    if index is not None and -n <= index < n:
        return np.take(data, index, axis=axis)
    if not return_nodata:
        raise IndexError(f"no element {index if index is not None else label}")
    ax = 0 if axis is None else axis
    return np.full(data.shape[:ax] + data.shape[ax + 1:], np.nan)


# ---- cumulative processes (what apply_dimension runs along a dimension) ---------------
def _cumulative(data, ignore_nodata, axis, nan_fn, plain_fn, fill=None, accumulate=None):
    data = np.array(data)
    gaps = np.isnan(data) if data.dtype.kind == "f" else np.zeros(data.shape, bool)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", RuntimeWarning)
        if accumulate is not None:   # cummin / cummax: math.py:149-174
            src = np.where(gaps, fill, data) if ignore_nodata else data
            out = accumulate(src, axis=axis)
        else:                        # cumsum / cumproduct: math.py:125-146
            out = (nan_fn if ignore_nodata else plain_fn)(data, axis=axis)
    if gaps.any():   # with axis=None numpy's cumulative functions return the flattened array
        out[gaps if axis is not None else gaps.reshape(-1)] = np.nan
    return out


def cumsum(data, ignore_nodata=True, axis=None):
    """math.py:125-134: np.nancumsum / np.cumsum, then result[nan_mask] = nan."""
    return _cumulative(data, ignore_nodata, axis, np.nancumsum, np.cumsum)


def cumproduct(data, ignore_nodata=True, axis=None):
    """math.py:137-146: np.nancumprod / np.cumprod, then result[nan_mask] = nan."""
    return _cumulative(data, ignore_nodata, axis, np.nancumprod, np.cumprod)


def cummin(data, ignore_nodata=True, axis=None):
    """math.py:149-160: np.minimum.accumulate over the +inf-filled (or raw) data, then result[nan_mask] = nan."""
    return _cumulative(data, ignore_nodata, axis, None, None, np.inf, np.minimum.accumulate)


def cummax(data, ignore_nodata=True, axis=None):
    """math.py:163-174: np.maximum.accumulate over the -inf-filled (or raw) data, then result[nan_mask] = nan."""
    return _cumulative(data, ignore_nodata, axis, None, None, -np.inf, np.maximum.accumulate)


def sort(data, asc=True, nodata=None, axis=None):
    """arrays.py:560-584: np.sort / -np.sort(-data); nodata None drops the NaNs (flattening N-d input),
    False moves them to the front, True leaves numpy's order (NaNs last).

    nodata=False follows what the reference computes THROUGH apply_dimension, where the dimension always
    sits on the last axis (cubes/apply.py:66): each slice keeps its own samples.  (Called directly with
    another axis the reference's flattened boolean assignment mixes slices; not reproduced.)"""
    data = np.asarray(data)
    if len(data) == 0:
        return data
    ordered = np.sort(data, axis=axis) if asc else -np.sort(-data, axis=axis)
    if nodata is True:
        return ordered
    gaps = np.isnan(ordered) if ordered.dtype.kind == "f" else np.zeros(ordered.shape, bool)
    if nodata is None:
        return ordered[~gaps]
    ax = 0 if axis is None else axis
    moved = np.ascontiguousarray(np.moveaxis(ordered, ax, -1))
    mgaps = np.isnan(moved) if moved.dtype.kind == "f" else np.zeros(moved.shape, bool)
    out = np.flip(moved, axis=-1).copy()      # NaNs in front; refill the valid slots in sorted order
    out[~np.flip(mgaps, axis=-1)] = moved[~mgaps]
    return np.ascontiguousarray(np.moveaxis(out, -1, ax))


def array_interpolate_linear(data, axis=None, positions=None):
    """arrays.py:405-458 for ndarray input: along ``axis`` every slice is passed to the reference's ``interp``:
    valid = isfinite; all valid or fewer than two valid -> unchanged; else the invalid samples become
    np.interp(x[~valid], x[valid], data[valid], left=nan, right=nan).  ``positions`` = x (default 0..n-1,
    arrays.py:421-425)."""
    data = np.array(data)
    n = data.shape[axis] if axis is not None else len(data)
    x = np.arange(n) if positions is None else np.asarray(positions)

    def fill(col):
        col = np.asarray(col)
        ok = np.isfinite(col)
        if ok.all() or ok.sum() < 2:
            return col
        col = col.copy()
        col[~ok] = np.interp(x[~ok], x[ok], col[ok], left=np.nan, right=np.nan)
        return col

    if data.dtype.kind != "f":
        return data
    if axis is None:
        return fill(data)
    return np.apply_along_axis(fill, axis, data)


def clip(x, min, max):
    """math.py:332-338: np.clip(x, min, max) (min > max raises MinMaxSwapped in the product)."""
    if min > max:
        raise ValueError("min > max")
    return np.clip(np.asarray(x), min, max)


def linear_scale_range(x, inputMin, inputMax, outputMin=0.0, outputMax=1.0):
    """math.py:292-298, the same python expression."""
    x = np.asarray(x)
    x = clip(x, inputMax, inputMin) if inputMax < inputMin else clip(x, inputMin, inputMax)
    return ((x - inputMin) / (inputMax - inputMin)) * (outputMax - outputMin) + outputMin


# ---- binary elementwise processes (merge_cubes overlap resolvers, apply chains) -------------------------------
def _notnull(x):
    import pandas as pd

    return ~np.asarray(pd.isnull(x))  # cubes/utils.py:16-25 (xarray's isnull: an array even for a scalar)


def add(x, y):
    """math.py:113-115"""
    return x + y


def subtract(x, y):
    """math.py:103-105"""
    return x - y


def multiply(x, y):
    """math.py:108-110"""
    return x * y


def divide(x, y):
    """math.py:98-100"""
    with np.errstate(all="ignore"):
        return x / y


def _cmp(x, y, res):
    return np.where(np.logical_and(_notnull(x), _notnull(y)), res, np.nan)


def eq(x, y):
    """comparison.py:45-87 (numeric operands, no delta)"""
    return _cmp(x, y, x == y)


def neq(x, y):
    """comparison.py:90-100"""
    return _cmp(x, y, np.logical_not(eq(x, y)))


def gt(x, y):
    """comparison.py:103-105"""
    return _cmp(x, y, x > y)


def gte(x, y):
    """comparison.py:117-119: (x - y) >= 0"""
    with np.errstate(all="ignore"):
        return _cmp(x, y, (x - y) >= 0)


def lt(x, y):
    """comparison.py:108-110"""
    return _cmp(x, y, x < y)


def lte(x, y):
    """comparison.py:122-124"""
    return _cmp(x, y, x <= y)


def _and(x, y):
    """logic.py:11-19"""
    nan_x, nan_y = ~_notnull(x), ~_notnull(y)
    xy = np.logical_and(x, y)
    xy = np.where(~np.logical_and(nan_x, xy), xy, np.nan)
    return np.where(~np.logical_and(nan_y, xy), xy, np.nan)


def _or(x, y):
    """logic.py:22-32"""
    nan_x, nan_y = ~_notnull(x), ~_notnull(y)
    xy = np.logical_or(np.nan_to_num(x), np.nan_to_num(y))
    xy = np.where(~np.logical_and(nan_x, np.logical_not(xy)), xy, np.nan)
    return np.where(~np.logical_and(nan_y, np.logical_not(xy)), xy, np.nan)


def xor(x, y):
    """logic.py:35-40"""
    nan_x, nan_y = ~_notnull(x), ~_notnull(y)
    xy = np.logical_xor(x, y)
    xy = np.where(~nan_x, xy, np.nan)
    return np.where(~nan_y, xy, np.nan)


BINARY = {"add": add, "subtract": subtract, "multiply": multiply, "divide": divide, "eq": eq, "neq": neq, "gt": gt,
          "gte": gte, "lt": lt, "lte": lte, "and": _and, "or": _or, "xor": xor}


# ---- unary math processes with exact results -----------------------------------------------------------------
def absolute(x):
    """math.py:305-306"""
    return np.abs(x)


def sgn(x):
    """math.py:309-310"""
    return np.sign(x)


def sqrt(x):
    """math.py:313-314"""
    with np.errstate(all="ignore"):
        return np.sqrt(x)


def floor(x):
    """math.py:212-213"""
    return np.floor(x)


def ceil(x):
    """math.py:216-217"""
    return np.ceil(x)


def _int(x):
    """math.py:220-221"""
    return np.trunc(x)


UNARY = {"absolute": absolute, "sgn": sgn, "sqrt": sqrt, "floor": floor, "ceil": ceil, "int": _int}


def order(data, asc=True, nodata=None, axis=None):
    """arrays.py:498-534 restated line by line (stable argsort, np.flip over all axes for asc=False, NaN handling)."""
    data = np.asarray(data)
    if len(data) == 0:
        return data
    perm = np.argsort(data, kind="mergesort", axis=axis)
    if not asc:
        perm = np.flip(perm)
    null = (lambda a: np.isnan(a) if a.dtype.kind == "f" else np.zeros(a.shape, bool))
    if nodata is None:
        if data.ndim > 1:
            raise ValueError("order with nodata=None is not supported for arrays with more than one dimension")
        sorted_data = np.take_along_axis(data, perm, axis=axis)
        return perm[~null(sorted_data)]
    if nodata is False:
        sorted_data = data[perm]
        return np.append(perm[null(sorted_data)], perm[~null(sorted_data)])
    return perm


def rearrange(data, order, axis=None):
    """arrays.py:537-555."""
    data = np.asarray(data)
    if len(data) == 0:
        return data
    if data.ndim == 1 and axis is None:
        axis = 0
    order = np.asarray(order)
    if order.ndim != 1:
        raise ValueError("rearrange: order must be one-dimensional")
    return np.take(data, indices=order, axis=axis)
