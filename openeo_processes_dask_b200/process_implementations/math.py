"""Array-level reducers with the reference's signatures, executed by the CUDA library.

Each function is the B200 counterpart of one numpy one-liner of the reference's
``process_implementations/math.py`` (line numbers below are the reference's).  ``data`` may be

* a ``numpy.ndarray`` (or list)     -> strip pipeline through ``eoc_*_host``; numpy result
* a CUDA ``torch.Tensor``           -> device kernels on the current stream; cuda tensor result
* a ``DeviceArray``                 -> as above, DeviceArray result (what xarray holds in ``.data``)
* a ``dask.array.Array``            -> lazy ``map_blocks`` over spatial chunks (reduced axis in one
                                       chunk), each block going through the numpy path

Result dtypes follow numpy: float32 stays float32, float64 stays float64, integer input gives
float64 for mean/median/sd/variance, the input dtype for min/max, (u)int64 for sum/product;
``quantiles`` is float64, ``count`` int64.
"""

from __future__ import annotations

import logging

import numpy as np

from .. import _lib as L
from .. import engine
from ..device_array import DeviceArray, LazyMasked, LazyNormalizedDifference
from ..lazy_elementwise import LazyElementwise
from ..views import normalize_axes
from .exceptions import MinMaxSwapped, QuantilesParameterConflict, QuantilesParameterMissing

__all__ = ["_min", "_max", "median", "mean", "sd", "variance", "_sum", "product", "quantiles", "extrema",
           "normalized_difference", "cumsum", "cumproduct", "cummin", "cummax", "clip", "linear_scale_range",
           "add", "subtract", "multiply", "divide", "absolute", "sgn", "sqrt", "floor", "ceil", "_int", "_round",
           "exp", "log", "ln", "cos", "arccos", "cosh", "arcosh", "sin", "arcsin", "sinh", "arsinh", "tan", "arctan",
           "tanh", "artanh", "arctan2", "mod", "power"]

logger = logging.getLogger(__name__)


# ---- input kinds ----------------------------------------------------------------------------------
def _is_dask(x) -> bool:
    mod = type(x).__module__ or ""
    return mod.startswith("dask.") and hasattr(x, "map_blocks")


def _unwrap(data):
    """-> (raw array, rewrap function)."""
    if isinstance(data, (LazyNormalizedDifference, LazyMasked)):
        data = data.materialize()
    if isinstance(data, DeviceArray):
        return data.tensor, DeviceArray
    if engine.is_torch(data):
        return data, (lambda t: t)
    return np.asarray(data), (lambda a: a)


def _keepdims(res, shape, axis, keepdims):
    if not keepdims:
        return res
    shape = list(shape)
    for ax in normalize_axes(axis, len(shape)):
        shape[ax] = 1
    return res.reshape(shape)


def _int_result_dtype(op, in_dtype: np.dtype):
    """numpy's result dtype for integer input (None = keep the float64 the kernel produced)."""
    if in_dtype.kind not in "iub":
        return None
    if op in (L.OP_MIN, L.OP_MAX):
        return in_dtype if in_dtype.kind != "b" else np.dtype(np.bool_)
    if op in (L.OP_SUM, L.OP_PRODUCT):
        return np.dtype(np.uint64) if in_dtype.kind == "u" else np.dtype(np.int64)
    return None


def _dask_blocks(data, fn, axis, new_len=None, dtype=None):
    axes = normalize_axes(axis, data.ndim)
    data = data.rechunk({ax: -1 for ax in axes})
    kwargs = dict(drop_axis=list(axes), dtype=dtype)
    if new_len is not None:
        kwargs.update(new_axis=[axes[0]], chunks=None)
        chunks = list(c for i, c in enumerate(data.chunks) if i not in axes)
        chunks.insert(axes[0], (new_len,))
        kwargs["chunks"] = tuple(chunks)
    return data.map_blocks(fn, **kwargs)


def _reduce(op, data, ignore_nodata, axis, keepdims, ddof=1):
    if _is_dask(data):
        probe = np.dtype(data.dtype)
        out_dt = _int_result_dtype(op, probe) or engine.acc_dtype(probe)
        out = _dask_blocks(data, lambda b: _reduce(op, b, ignore_nodata, axis, False, ddof), axis, dtype=out_dt)
        if keepdims:
            import dask.array as da

            out = da.expand_dims(out, normalize_axes(axis, data.ndim))
        return out
    if isinstance(data, LazyNormalizedDifference):  # ndvi(..., fuse=True) -> fused nd + reduce kernel
        return _keepdims(data.reduce(op, axis, bool(ignore_nodata), ddof), data.shape, axis, keepdims)
    if isinstance(data, LazyMasked):  # mask(..., fuse=True) -> mask consumed in the load stage of the reduce
        res = data.reduce(op, axis, bool(ignore_nodata), ddof)
        cast = _int_result_dtype(op, data.dtype)
        if cast is not None:
            res = DeviceArray(engine._astype(res.tensor, cast)) if isinstance(res, DeviceArray) else engine._astype(res, cast)
        return _keepdims(res, data.shape, axis, keepdims)
    raw, rewrap = _unwrap(data)
    in_dt = engine.np_dtype_of(raw)
    shape = tuple(raw.shape)
    res = engine.reduce(raw, [op], axis, ignore_nodata=bool(ignore_nodata), ddof=ddof)[op]
    cast = _int_result_dtype(op, in_dt)
    if cast is not None:
        res = engine._astype(res, cast)
    return rewrap(_keepdims(res, shape if shape else (1,), axis, keepdims))


# ---- the reducers ---------------------------------------------------------------------------------
def _min(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:118-122 -- np.nanmin / np.min."""
    _require_nonempty(data, axis, "minimum")
    return _reduce(L.OP_MIN, data, ignore_nodata, axis, keepdims)


def _max(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:177-181 -- np.nanmax / np.max."""
    _require_nonempty(data, axis, "maximum")
    return _reduce(L.OP_MAX, data, ignore_nodata, axis, keepdims)


def mean(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:191-195 -- np.nanmean / np.mean."""
    return _reduce(L.OP_MEAN, data, ignore_nodata, axis, keepdims)


def sd(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:198-202 -- np.nanstd / np.std with ddof=1."""
    return _reduce(L.OP_SD, data, ignore_nodata, axis, keepdims, ddof=1)


def variance(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:205-209 -- np.nanvar / np.var with ddof=1."""
    return _reduce(L.OP_VARIANCE, data, ignore_nodata, axis, keepdims, ddof=1)


def _sum(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:390-398 -- np.nansum / np.sum; ``len(data) == 0`` -> nan."""
    if len(data) == 0:
        return np.nan
    return _reduce(L.OP_SUM, data, ignore_nodata, axis, keepdims)


def product(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:401-409 -- np.nanprod / np.prod; ``len(data) == 0`` -> nan."""
    if len(data) == 0:
        return np.nan
    return _reduce(L.OP_PRODUCT, data, ignore_nodata, axis, keepdims)


def _require_nonempty(data, axis, what):
    shape = tuple(np.shape(data)) if not hasattr(data, "shape") else tuple(data.shape)
    if not shape:
        return
    if any(shape[ax] == 0 for ax in normalize_axes(axis, len(shape))):
        # numpy: "zero-size array to reduction operation fmin which has no identity"
        raise ValueError(f"zero-size array to reduction operation {what} which has no identity")


def median(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:184-188 -- np.nanmedian / np.median."""
    if _is_dask(data):
        return _dask_blocks(data, lambda b: median(b, ignore_nodata, axis, False), axis,
                            dtype=engine.acc_dtype(np.dtype(data.dtype)))
    if isinstance(data, LazyMasked) and _selectable(data, axis):  # mask(..., fuse=True) -> one pass
        res, _ = data.select(axis, bool(ignore_nodata), True, None)
        return _keepdims(res, data.shape, axis, keepdims)
    raw, rewrap = _unwrap(data)
    shape = tuple(raw.shape)
    res, _ = engine.select(raw, axis, bool(ignore_nodata), True, None)
    return rewrap(_keepdims(res, shape if shape else (1,), axis, keepdims))


def _selectable(lazy, axis) -> bool:
    """The fused mask + selection path: a single reduced axis of a non-empty cube."""
    return lazy.ndim >= 1 and all(n > 0 for n in lazy.shape) and len(normalize_axes(axis, lazy.ndim)) == 1


def _resolve_probabilities(probabilities, q):
    """reference math.py:344-371, including ``np.arange(1/p, 1, 1/p)`` evaluated verbatim (it has p
    entries, the last being 1.0, for p in {3, 6, 7, 14, ...})."""
    if probabilities is not None and q is not None:
        raise QuantilesParameterConflict(
            "The process `quantiles` requires either the `probabilities` or `q` parameter to be set.")
    if probabilities is None and q is None:
        raise QuantilesParameterMissing(
            "The process `quantiles` only allows that either the `probabilities` or the `q` parameter is set.")
    if isinstance(probabilities, int) and not isinstance(probabilities, bool):
        probabilities = np.arange(1.0 / probabilities, 1, 1.0 / probabilities)
    elif isinstance(probabilities, list) and len(probabilities) == 1 and isinstance(probabilities[0], int):
        probabilities = np.arange(1.0 / probabilities[0], 1, 1.0 / probabilities[0])
    elif isinstance(probabilities, list):
        probabilities = np.array(probabilities)
    if q is not None:
        logger.warning("The 'q' parameter has been **deprecated**. Please use the parameter `probabilities` instead.")
        probabilities = np.arange(1.0 / q, 1, 1.0 / q)
    return np.asarray(probabilities, dtype=np.float64)


def quantiles(data, probabilities=None, q=None, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:341-387 -- np.nanquantile / np.quantile(method="linear").

    The quantile axis ends up where the reduced axis was (the reference's ``if axis: moveaxis``), the
    result is float64 (numpy's promotion with float64 probabilities).
    """
    probs = _resolve_probabilities(probabilities, q)
    if _is_dask(data):
        return _dask_blocks(data, lambda b: quantiles(b, probabilities=list(probs), ignore_nodata=ignore_nodata,
                                                      axis=axis), axis, new_len=len(probs), dtype=np.float64)
    if (isinstance(data, LazyMasked) and _selectable(data, axis) and 0 < len(probs) <= L.SELECT_MAX_Q
            and not keepdims):
        _, res = data.select(axis, bool(ignore_nodata), False, probs)
        return res
    raw, rewrap = _unwrap(data)
    if int(np.prod(raw.shape)) == 0:  # math.py:373-374
        return np.array([np.nan] * len(probs))
    shape = tuple(raw.shape)
    res = None
    for lo in range(0, max(len(probs), 1), L.SELECT_MAX_Q):
        _, part = engine.select(raw, axis, bool(ignore_nodata), False, probs[lo:lo + L.SELECT_MAX_Q])
        if part is None:  # no probabilities at all
            pos = normalize_axes(axis, len(shape))[0]
            kept = [n for i, n in enumerate(shape) if i not in normalize_axes(axis, len(shape))]
            kept.insert(pos, 0)
            return rewrap(engine._empty_like_kind(raw, kept, np.dtype(np.float64)))
        pos = normalize_axes(axis, len(shape))[0]
        res = part if res is None else _concat(res, part, pos)
    if keepdims:  # numpy keeps the reduced axes as length 1 right after the quantile axis
        ax = normalize_axes(axis, len(shape))
        res = res.reshape(list(res.shape[:ax[0] + 1]) + [1] * len(ax) + list(res.shape[ax[0] + 1:]))
    return rewrap(res)


def _concat(a, b, axis):
    if engine.is_torch(a):
        import torch

        return torch.cat([a, b], dim=axis)
    return np.concatenate([a, b], axis=axis)


def extrema(data, ignore_nodata=True, axis=None, keepdims=False):
    """reference math.py:322-329: [min, max].  The reference builds a lazy ``(2,)`` dask array
    (``da.from_delayed``) whatever the input; here dask input stays lazy (same array kind out as in), numpy /
    device input is evaluated."""
    if isinstance(data, list):
        data = np.array(data)
    if _is_dask(data):
        import dask.array as da

        lo = _reduce(L.OP_MIN, data, ignore_nodata, axis, False)
        hi = _reduce(L.OP_MAX, data, ignore_nodata, axis, False)
        return da.stack([lo, hi])
    raw, rewrap = _unwrap(data)
    res = engine.reduce(raw, [L.OP_MIN, L.OP_MAX], axis, ignore_nodata=bool(ignore_nodata))
    lo, hi = res[L.OP_MIN], res[L.OP_MAX]
    if engine.is_torch(lo):
        import torch

        return rewrap(torch.stack([lo, hi]))
    return np.array([lo, hi])


def normalized_difference(x, y):
    """reference math.py:412-414 -- (x - y) / (x + y), elementwise.

    Same-shape, same-dtype arrays take ONE pass of ``eoc_nd_reduce`` with no reducer selected (integer cubes with
    numpy's semantics: difference and sum wrap in the sample dtype, float64 quotient).  Everything else --
    scalars, broadcasting, mixed dtypes, dask -- is the same expression through ``eoc_binary``: ``subtract``,
    ``add``, ``divide`` (numpy's promotion, broadcasting and integer wrap-around), three launches.
    There is no host arithmetic on either route.
    """
    if not (_is_dask(x) or _is_dask(y) or engine._is_host_scalar(x) or engine._is_host_scalar(y)
            or isinstance(x, list) or isinstance(y, list)):
        xr_, rewrap = _unwrap(x)
        yr_, _ = _unwrap(y)
        dt = engine.np_dtype_of(xr_)
        if (tuple(xr_.shape) == tuple(yr_.shape) and dt == engine.np_dtype_of(yr_) and len(xr_.shape) > 0
                and engine.is_torch(xr_) == engine.is_torch(yr_) and dt in L.DTYPE_CODE):
            _, nd = engine.nd_reduce(xr_, yr_, None, [], want_nd=True, integer_wrap=dt.kind in "iu")
            return rewrap(nd)
    return divide(subtract(x, y), add(x, y))


# ---- extension: several reducers of the same array in one pass ---------------------------------------
_STREAM_OPS = {
    "sum": L.OP_SUM, "mean": L.OP_MEAN, "min": L.OP_MIN, "max": L.OP_MAX, "sd": L.OP_SD,
    "variance": L.OP_VARIANCE, "product": L.OP_PRODUCT,
}


def reduce_many(data, reducers, ignore_nodata=True, axis=None):
    """EXTENSION (not a reference process): evaluate several reducers of ``data`` over ``axis`` with one
    read of the cube.  ``reducers`` is a list of process names or ``(name, kwargs)`` pairs out of
    ``sum mean min max sd variance product median quantiles``; the result is a list of arrays in the same
    order, each identical to what the single process returns.

    In a process graph these are sibling ``reduce_dimension`` nodes over the same cube (BASELINE config 3:
    ``quantiles([0.1, 0.5, 0.9])`` next to ``sd``); the reference runs one full numpy pass per node
    (math.py:198-202, 341-387).  A backend that spots the pattern can call this instead: order statistics
    and streaming reducers come out of the same register-resident column (``eoc_select_fused``).
    """
    specs = [(r, {}) if isinstance(r, str) else (r[0], dict(r[1])) for r in reducers]
    names = [n.lstrip("_") for n, _ in specs]
    unknown = [n for n in names if n not in _STREAM_OPS and n not in ("median", "quantiles")]
    if unknown:
        raise ValueError(f"reduce_many: unsupported reducer(s) {unknown}")
    if _is_dask(data) or isinstance(data, LazyNormalizedDifference) or sum(n == "quantiles" for n in names) > 1:
        single = {"sum": _sum, "min": _min, "max": _max}
        return [(single.get(n) or globals()[n])(data, ignore_nodata=ignore_nodata, axis=axis, **kw)
                for n, (_, kw) in zip(names, specs)]
    raw, rewrap = _unwrap(data)
    in_dt = engine.np_dtype_of(raw)
    ops = sorted({_STREAM_OPS[n] for n in names if n in _STREAM_OPS})
    want_median = "median" in names
    probs = None
    for n, (_, kw) in zip(names, specs):
        if n == "quantiles":
            probs = _resolve_probabilities(kw.get("probabilities"), kw.get("q"))
    if probs is not None and (len(probs) == 0 or len(probs) > L.SELECT_MAX_Q or int(np.prod(raw.shape)) == 0):
        return [(({"sum": _sum, "min": _min, "max": _max}).get(n) or globals()[n])(
            data, ignore_nodata=ignore_nodata, axis=axis, **kw) for n, (_, kw) in zip(names, specs)]
    if want_median or probs is not None:
        med, qs, outs = engine.select(raw, axis, bool(ignore_nodata), want_median, probs, ops=ops)
    else:
        med, qs, outs = None, None, engine.reduce(raw, ops, axis, ignore_nodata=bool(ignore_nodata))
    res = []
    for n in names:
        if n == "median":
            res.append(rewrap(med))
        elif n == "quantiles":
            res.append(rewrap(qs))
        else:
            op = _STREAM_OPS[n]
            r = outs[op]
            cast = _int_result_dtype(op, in_dt)
            res.append(rewrap(engine._astype(r, cast) if cast is not None else r))
    return res


# ---- cumulative processes (what apply_dimension runs along a dimension) -----------------------------------
def _scan(op, data, ignore_nodata, axis):
    if isinstance(data, list):
        data = np.asarray(data)
    if _is_dask(data):
        ax = normalize_axes(axis, data.ndim)[0] if axis is not None else None
        src = data.rechunk({ax: -1}) if ax is not None else data.rechunk(data.shape)
        return src.map_blocks(lambda b: _scan(op, b, ignore_nodata, axis), dtype=engine.acc_dtype(np.dtype(data.dtype)))
    raw, rewrap = _unwrap(data)
    in_dt = engine.np_dtype_of(raw)
    if axis is None:  # numpy: cumulative functions flatten without an axis
        raw = raw.reshape(-1)
        axis = 0
    res = engine.scan(raw, op, axis, bool(ignore_nodata))
    if in_dt.kind in "iu" and op in (L.SCAN_SUM, L.SCAN_PRODUCT):   # np.cumsum of integers: (u)int64
        res = engine._astype(res, np.dtype(np.uint64) if in_dt.kind == "u" else np.dtype(np.int64))
    elif in_dt.kind in "iu" and not ignore_nodata:                    # np.minimum.accumulate keeps the dtype;
        res = engine._astype(res, in_dt)                              # the inf-filled copy (ignore) is float64
    return rewrap(res)


def cumsum(data, ignore_nodata=True, axis=None):
    """reference math.py:125-134 -- np.nancumsum / np.cumsum, NaN slots stay NaN."""
    return _scan(L.SCAN_SUM, data, ignore_nodata, axis)


def cumproduct(data, ignore_nodata=True, axis=None):
    """reference math.py:137-146 -- np.nancumprod / np.cumprod, NaN slots stay NaN."""
    return _scan(L.SCAN_PRODUCT, data, ignore_nodata, axis)


def cummin(data, ignore_nodata=True, axis=None):
    """reference math.py:149-160 -- np.minimum.accumulate over the +inf-filled (or raw) data, NaN slots stay NaN."""
    return _scan(L.SCAN_MIN, data, ignore_nodata, axis)


def cummax(data, ignore_nodata=True, axis=None):
    """reference math.py:163-174 -- np.maximum.accumulate over the -inf-filled (or raw) data, NaN slots stay NaN."""
    return _scan(L.SCAN_MAX, data, ignore_nodata, axis)


# ---- elementwise processes (what apply runs over a cube) ----------------------------------------------------
def _elementwise(op, x, params, integral_result):
    if isinstance(x, (list, int, float)):
        x = np.asarray(x)
    if _is_dask(x):
        return x.map_blocks(lambda b: _elementwise(op, b, params, integral_result),
                            dtype=x.dtype if integral_result and np.dtype(x.dtype).kind in "iu" else engine.acc_dtype(np.dtype(x.dtype)))
    raw, rewrap = _unwrap(x)
    in_dt = engine.np_dtype_of(raw)
    res = engine.elementwise(raw, op, params)
    if integral_result and in_dt.kind in "iu":
        res = engine._astype(res, in_dt)
    return rewrap(res)


def clip(x, min, max):
    """reference math.py:332-338 -- np.clip(x, min, max); ``min > max`` raises MinMaxSwapped."""
    if min > max:
        raise MinMaxSwapped("The minimum value should be lower than or equal to the maximum value.")
    if isinstance(x, LazyElementwise):  # np.clip == minimum(maximum(x, min), max), NaN stays NaN
        return x.binary(L.TAPE_MAXNUM, min).binary(L.TAPE_MINNUM, max)
    integral = float(min).is_integer() and float(max).is_integer() and not isinstance(min, float) and not isinstance(max, float)
    return _elementwise(L.ELEM_CLIP, x, [min, max], integral)


def linear_scale_range(x, inputMin, inputMax, outputMin=0.0, outputMax=1.0):
    """reference math.py:292-298 -- clip to the input range, then
    ``((x - inputMin) / (inputMax - inputMin)) * (outputMax - outputMin) + outputMin``."""
    if isinstance(x, LazyElementwise):  # the reference's own sequence of operations, one rounding each
        lo, hi = (inputMax, inputMin) if inputMax < inputMin else (inputMin, inputMax)
        return (clip(x, lo, hi).binary(L.TAPE_SUB, inputMin).binary(L.TAPE_DIV, inputMax - inputMin)
                .binary(L.TAPE_MUL, outputMax - outputMin).binary(L.TAPE_ADD, outputMin))
    return _elementwise(L.ELEM_LINEAR_SCALE_RANGE, x, [inputMin, inputMax, outputMin, outputMax], False)


# ---- binary elementwise processes (apply chains, merge_cubes overlap resolvers) -----------------------------
def _binary(op, x, y):
    """x OP y through ``eoc_binary`` (numpy broadcasting and promotion); the result has the operands' array kind
    (DeviceArray if either side is one, CUDA tensor, numpy; dask stays lazy)."""
    if isinstance(x, list):
        x = np.asarray(x)
    if isinstance(y, list):
        y = np.asarray(y)
    if _is_dask(x) or _is_dask(y):
        import dask.array as da

        d = x if _is_dask(x) else y
        other = y if d is x else x
        # python scalars are weak (NEP 50): float32 cube + 2.5 stays float32
        common = np.result_type(np.dtype(d.dtype), other if engine._is_host_scalar(other) and not hasattr(other, "dtype")
                                else np.dtype(other.dtype))
        return da.map_blocks(lambda a, b: _binary(op, a, b), x, y, dtype=engine.binary_result_dtype(op, common))
    if op in _TAPE_OF_BIN and (isinstance(x, LazyElementwise) or isinstance(y, LazyElementwise)):
        fused = (x.binary(_TAPE_OF_BIN[op], y) if isinstance(x, LazyElementwise)
                 else y.binary(_TAPE_OF_BIN[op], x, reverse=True))
        if fused is not None:
            return fused
    wrap = isinstance(x, DeviceArray) or isinstance(y, DeviceArray)
    rx = _unwrap(x)[0] if not engine._is_host_scalar(x) else x
    ry = _unwrap(y)[0] if not engine._is_host_scalar(y) else y
    res = engine.binary(rx, ry, op)
    return DeviceArray(res) if wrap else res


_TAPE_OF_BIN = {L.BIN_ADD: L.TAPE_ADD, L.BIN_SUBTRACT: L.TAPE_SUB, L.BIN_MULTIPLY: L.TAPE_MUL, L.BIN_DIVIDE: L.TAPE_DIV}


def add(x, y):
    """reference math.py:113-115 -- ``x + y``."""
    return _binary(L.BIN_ADD, x, y)


def subtract(x, y):
    """reference math.py:103-105 -- ``x - y``."""
    return _binary(L.BIN_SUBTRACT, x, y)


def multiply(x, y):
    """reference math.py:108-110 -- ``x * y``."""
    return _binary(L.BIN_MULTIPLY, x, y)


def divide(x, y):
    """reference math.py:98-100 -- ``x / y`` (true division: integers give float64; x/0 is +-inf, 0/0 NaN)."""
    return _binary(L.BIN_DIVIDE, x, y)


# ---- unary math processes with exact results (apply chains) ---------------------------------------------------
def _is_integer_array(x) -> bool:
    dt = getattr(x, "dtype", None)
    if dt is None:
        return isinstance(x, (int, np.integer)) and not isinstance(x, bool)
    return (engine.np_dtype_of(_unwrap(x)[0]) if not isinstance(x, np.ndarray) else x.dtype).kind in "iu"


def absolute(x):
    """reference math.py:305-306 -- np.abs.  Integer input keeps its dtype (computed through float64: exact for
    the <= 32-bit integers the kernels read, except that abs(iinfo.min) saturates instead of wrapping)."""
    if isinstance(x, LazyElementwise):
        return x.unary(L.TAPE_ABS)
    return _elementwise(L.ELEM_ABSOLUTE, x, [], True)


def sgn(x):
    """reference math.py:309-310 -- np.sign: -1 / 0 / +1, NaN stays NaN, sign(-0.0) is +0.0."""
    if isinstance(x, LazyElementwise):
        return x.unary(L.TAPE_SGN)
    return _elementwise(L.ELEM_SGN, x, [], True)


def sqrt(x):
    """reference math.py:313-314 -- np.sqrt (IEEE, exactly rounded); negative input gives NaN.  Integer input gives
    numpy's result dtype (float32 for 16-bit, float64 for 32-bit integers; float64 -> float32 rounding of an
    exactly rounded float64 square root is itself exactly rounded, since 53 >= 2 * 24 + 2)."""
    if isinstance(x, LazyElementwise):
        return x.unary(L.TAPE_SQRT)
    res = _elementwise(L.ELEM_SQRT, x, [], False)
    if _is_integer_array(x):
        in_dt = np.asarray(x).dtype if isinstance(x, (list, int, np.ndarray, np.integer)) else engine.np_dtype_of(_unwrap(x)[0])
        want = np.sqrt(np.zeros(1, dtype=in_dt)).dtype
        raw, rewrap = _unwrap(res)
        res = rewrap(engine._astype(raw, np.dtype(want)))
    return res


_TAPE_OF_ROUNDING = {L.ELEM_FLOOR: L.TAPE_FLOOR, L.ELEM_CEIL: L.TAPE_CEIL, L.ELEM_TRUNC: L.TAPE_TRUNC}


def _rounding(op, x):
    if isinstance(x, LazyElementwise):
        return x.unary(_TAPE_OF_ROUNDING[op])
    if _is_integer_array(x):  # numpy: floor / ceil / trunc of integers is the input itself
        return x.clone() if engine.is_torch(x) else (x if isinstance(x, DeviceArray) else np.array(x))
    return _elementwise(op, x, [], False)


def floor(x):
    """reference math.py:212-213 -- np.floor."""
    return _rounding(L.ELEM_FLOOR, x)


def ceil(x):
    """reference math.py:216-217 -- np.ceil (ceil(-0.5) is -0.0)."""
    return _rounding(L.ELEM_CEIL, x)


def _int(x):
    """reference math.py:220-221 -- np.trunc."""
    return _rounding(L.ELEM_TRUNC, x)


# ---- unary / binary math processes through the expression kernel (eoc_apply_tape) ----------------------------------
def _numpy_result_dtype(np_func, *dtypes_or_scalars):
    """dtype numpy gives ``np_func`` for operands of these dtypes (python scalars stay weak)"""
    args = [np.ones(1, dtype=a) if isinstance(a, np.dtype) else a for a in dtypes_or_scalars]
    with np.errstate(all="ignore"):
        return np.asarray(np_func(*args)).dtype


def _tape_unary(code, np_func, x, imm=0.0):
    """one-step expression ``f(x)``: extends a LazyElementwise, else one launch of ``eoc_apply_tape``.  Computed in
    float32 for float32 input, in float64 otherwise, and cast to numpy's result dtype for the input dtype."""
    if isinstance(x, (list, int, float)):
        x = np.asarray(x)
    if isinstance(x, LazyElementwise):
        return x.unary(code, imm)
    if _is_dask(x):
        return x.map_blocks(lambda b: _tape_unary(code, np_func, b, imm), dtype=_numpy_result_dtype(np_func, np.dtype(x.dtype)))
    raw, rewrap = _unwrap(x)
    in_dt = engine.np_dtype_of(raw)
    want = _numpy_result_dtype(np_func, in_dt)
    raw = engine._native(raw)
    compute = np.dtype(np.float32) if engine.np_dtype_of(raw) == np.float32 else np.dtype(np.float64)
    res = engine.apply_tape(raw, None, [(code, 0, imm)], compute)
    return rewrap(engine._astype(res, want))


def _round(x, p=0):
    """reference math.py:224-225 -- np.around(x, decimals=p): round half to even at 10^-p, computed as numpy does
    (scale by the power of ten in the array's dtype, rint, scale back).  Integers with p >= 0 are returned as they
    are."""
    p = int(p)
    if _is_integer_array(x) and p >= 0:
        return x.clone() if engine.is_torch(x) else (x if isinstance(x, DeviceArray) else np.array(x))
    if p >= 0:
        return _tape_unary(L.TAPE_ROUND, lambda a: np.around(a, decimals=p), x, 10.0 ** p)
    return _tape_unary(L.TAPE_ROUND_NEG, lambda a: np.around(a, decimals=p), x, 10.0 ** (-p))


def exp(p):
    """reference math.py:228-229 -- np.exp."""
    return _tape_unary(L.TAPE_EXP, np.exp, p)


def ln(x):
    """reference math.py:236-237 -- np.log (ln(0) = -inf, ln(x < 0) = NaN)."""
    return _tape_unary(L.TAPE_LN, np.log, x)


def log(x, base):
    """reference math.py:232-233 -- ``np.log(x) / np.log(base)``.  ``np.log(base)`` of a python scalar is a
    numpy float64 scalar, which is NOT weak: a float32 cube gives a float64 result, as in the reference."""
    denominator = np.log(base) if engine._is_host_scalar(base) and not isinstance(base, DeviceArray) else ln(base)
    return divide(ln(x), denominator)


def cos(x):
    """reference math.py:240-241"""
    return _tape_unary(L.TAPE_COS, np.cos, x)


def arccos(x):
    """reference math.py:244-245"""
    return _tape_unary(L.TAPE_ACOS, np.arccos, x)


def cosh(x):
    """reference math.py:248-249"""
    return _tape_unary(L.TAPE_COSH, np.cosh, x)


def arcosh(x):
    """reference math.py:252-253"""
    return _tape_unary(L.TAPE_ACOSH, np.arccosh, x)


def sin(x):
    """reference math.py:256-257"""
    return _tape_unary(L.TAPE_SIN, np.sin, x)


def arcsin(x):
    """reference math.py:260-261"""
    return _tape_unary(L.TAPE_ASIN, np.arcsin, x)


def sinh(x):
    """reference math.py:264-265"""
    return _tape_unary(L.TAPE_SINH, np.sinh, x)


def arsinh(x):
    """reference math.py:268-269"""
    return _tape_unary(L.TAPE_ASINH, np.arcsinh, x)


def tan(x):
    """reference math.py:272-273"""
    return _tape_unary(L.TAPE_TAN, np.tan, x)


def arctan(x):
    """reference math.py:276-277"""
    return _tape_unary(L.TAPE_ATAN, np.arctan, x)


def tanh(x):
    """reference math.py:280-281"""
    return _tape_unary(L.TAPE_TANH, np.tanh, x)


def artanh(x):
    """reference math.py:284-285"""
    return _tape_unary(L.TAPE_ATANH, np.arctanh, x)


def _tape_binary(code, rcode, np_func, a, b):
    """``np_func(a, b)`` for two arrays (numpy broadcasting) or an array and a scalar, in one launch; LazyElementwise
    operands extend their expression instead.  Result dtype is numpy's; integer results (``mod`` / ``power`` of
    integers) are computed in float64 -- exact while they fit 2^53, they do not wrap."""
    if isinstance(a, list):
        a = np.asarray(a)
    if isinstance(b, list):
        b = np.asarray(b)
    if isinstance(a, LazyElementwise) or isinstance(b, LazyElementwise):
        fused = a.binary(code, b) if isinstance(a, LazyElementwise) else b.binary(code, a, reverse=True)
        if fused is not None:
            return fused
    sa = engine._is_host_scalar(a) and not isinstance(a, DeviceArray)
    sb = engine._is_host_scalar(b) and not isinstance(b, DeviceArray)
    if _is_dask(a) or _is_dask(b):
        import dask.array as da

        kinds = [v if s and not hasattr(v, "dtype") else np.dtype(v.dtype) for v, s in ((a, sa), (b, sb))]
        return da.map_blocks(lambda p, q: _tape_binary(code, rcode, np_func, p, q), a, b,
                             dtype=_numpy_result_dtype(np_func, *kinds))
    if sa and sb:
        a, sa = np.asarray(a).reshape(1), False
        scalar_result = True
    else:
        scalar_result = False
    wrap = isinstance(a, DeviceArray) or isinstance(b, DeviceArray)
    ra = a if sa else _unwrap(a)[0]
    rb = b if sb else _unwrap(b)[0]
    kinds = [(v if not hasattr(v, "dtype") else np.dtype(v.dtype)) if s else engine.np_dtype_of(v)
             for v, s in ((ra, sa), (rb, sb))]
    want = _numpy_result_dtype(np_func, *kinds)
    compute = np.dtype(np.float32) if want == np.float32 and all(
        (not isinstance(k, np.dtype)) or k == np.float32 for k in kinds) else np.dtype(np.float64)
    any_torch = engine.is_torch(ra) or engine.is_torch(rb)

    device = next((v.device for v in (ra, rb) if engine.is_torch(v)), "cuda")

    def dev(v):
        if not engine.is_torch(v):
            import torch

            v = np.asarray(v)
            v = torch.from_numpy(np.ascontiguousarray(engine._native(v.view(np.uint8) if v.dtype == np.bool_ else v))).to(device)
        return engine._astype(engine._native(v), compute)

    if sa or sb:
        arr, scalar = (rb, ra) if sa else (ra, rb)
        res = engine.apply_tape(dev(arr), None, [(rcode if sa else code, 0, float(scalar))], compute)
    else:
        import torch

        ta, tb = dev(ra), dev(rb)
        if tb.device != ta.device:
            tb = tb.to(ta.device)
        ta, tb = torch.broadcast_tensors(ta, tb)
        res = engine.apply_tape(ta, tb, [(code, 2, 0.0)], compute)
    res = engine._astype(res, want)
    if scalar_result:
        res = res.reshape(())
    if not any_torch:
        return res.cpu().numpy()
    return DeviceArray(res) if wrap else res


def arctan2(y, x):
    """reference math.py:288-289 -- np.arctan2(y, x)."""
    return _tape_binary(L.TAPE_ATAN2, L.TAPE_RATAN2, np.arctan2, y, x)


def mod(x, y):
    """reference math.py:301-302 -- np.mod: the remainder has the sign of the divisor; x mod 0 is NaN for floats."""
    return _tape_binary(L.TAPE_MOD, L.TAPE_RMOD, np.mod, x, y)


def power(base, p):
    """reference math.py:317-319 -- ``base ** p``."""
    return _tape_binary(L.TAPE_POW, L.TAPE_RPOW, np.power, base, p)
