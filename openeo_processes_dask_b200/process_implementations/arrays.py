"""Array processes of the reducer category with the reference's signatures.

``count`` (reference ``arrays.py:587-606``):

* ``condition=None``  valid samples = not NaN and not +-inf, int64   -> EOC_OP_COUNT
* ``condition=True``  every sample, in the data's value dtype        -> EOC_OP_COUNT_ALL
* callable            ``np.nansum(condition(x=data, **context))``    -> condition evaluated
  elementwise by the caller's function, then an EOC_OP_SUM reduction of its 1/0/NaN result
"""

from __future__ import annotations

import numpy as np

from .. import _lib as L
from .. import engine
from . import math as _math
from .exceptions import (ArrayElementNotAvailable, ArrayElementParameterConflict, ArrayElementParameterMissing,
                         ArrayNotLabeled)

__all__ = ["count", "first", "last", "array_element", "sort", "order", "rearrange", "array_interpolate_linear"]


def _take(raw, index, axis):
    """np.take(data, index, axis) for numpy arrays and CUDA tensors (a strided view copy, no kernel of ours)."""
    if engine.is_torch(raw):
        n = raw.shape[axis]
        if not -n <= index < n:
            raise IndexError(f"index {index} is out of bounds for axis {axis} with size {n}")
        return raw.select(axis, index).clone()
    return np.take(raw, index, axis=axis)


def _first_last(op, data, ignore_nodata, axis):
    if isinstance(data, list):
        data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    if _math._is_dask(data):
        ax = 0 if axis is None else axis
        src = data.flatten() if axis is None else data
        return _math._dask_blocks(src, lambda b: _first_last(op, b, ignore_nodata, ax), ax, dtype=data.dtype)
    raw, rewrap = _math._unwrap(data)
    if axis is None:  # reference: data.flatten(), axis 0
        raw = raw.reshape(-1)
        axis = 0
    in_dt = engine.np_dtype_of(raw)
    if in_dt.kind in "iub" or not ignore_nodata:
        # no NaN to skip (integers), or NaN is not skipped: a plain element access like the reference's np.take
        return rewrap(_take(raw, 0 if op == L.OP_FIRST else -1, axis))
    res = engine.reduce(raw, [op], axis, ignore_nodata=True)[op]
    return rewrap(_clamp_like_reference(op, raw, res, axis))


def _clamp_like_reference(op, raw, res, axis):
    """Reference quirk, reproduced for parity: ``first`` fills gaps with ``np.nan_to_num(first_elem, True,
    data[i])`` for i = 0..max(k) (arrays.py:478-480; k = position of a slice's first valid sample), and
    nan_to_num also maps +-inf to +-finfo.max.  So when plane 0 holds a NaN anywhere, an INFINITE result is
    clamped if it was in place before the last iteration (k == 0 or k < max k).  Only results that contain an
    infinity take this path (one flag read back from the device); ``last`` is the same on the flipped axis."""
    torch_in = engine.is_torch(res)
    xp_isinf = (lambda a: a.isinf()) if torch_in else np.isinf
    inf_mask = xp_isinf(res)
    if not bool(inf_mask.any()):
        return res
    edge = _take(raw, 0 if op == L.OP_FIRST else -1, axis)
    if not bool((edge.isnan() if torch_in else np.isnan(edge)).any()):
        return res
    fmax = float(np.finfo(engine.np_dtype_of(res)).max)
    import torch

    if not torch_in:  # host arrays: the (rare) fix-up still runs on the device
        raw, res = torch.from_numpy(np.ascontiguousarray(raw)).cuda(), torch.from_numpy(np.ascontiguousarray(res)).cuda()
        inf_mask = res.isinf()
    valid = (~raw.isnan()).to(torch.uint8)
    if op == L.OP_LAST:
        valid = valid.flip(axis)
    k = valid.argmax(dim=axis)
    clamp = inf_mask & ((k == 0) | (k < k.max()))
    out = torch.where(clamp, torch.copysign(torch.full_like(res, fmax), res), res)
    return out if torch_in else out.cpu().numpy()


def first(data, ignore_nodata=True, axis=None):
    """reference arrays.py:461-485: first not-NaN element along ``axis`` (NaN when there is none);
    ``ignore_nodata=False``: element 0 whatever it holds."""
    return _first_last(L.OP_FIRST, data, ignore_nodata, axis)


def last(data, ignore_nodata=True, axis=None):
    """reference arrays.py:488-495: ``first`` of the flipped array."""
    return _first_last(L.OP_LAST, data, ignore_nodata, axis)


def array_element(data, index=None, label=None, return_nodata=False, axis=None, context=None, dim_labels=None):
    """reference arrays.py:66-122: ``np.take(data, index, axis)``; a ``label`` is looked up in ``dim_labels``
    (what ``reduce_dimension`` passes along).  Missing element: NaN-filled result with ``return_nodata`` or
    ``ArrayElementNotAvailable``."""
    if index is None and label is None:
        raise ArrayElementParameterMissing(
            "The process `array_element` requires either the `index` or `labels` parameter to be set.")
    if index is not None and label is not None:
        raise ArrayElementParameterConflict(
            "The process `array_element` only allows that either the `index` or the `labels` parameter is set.")
    if isinstance(data, list):
        data = np.asarray(data)
    labels = [] if dim_labels is None else dim_labels
    if label is not None:
        if len(labels) == 0:
            raise ArrayNotLabeled(
                "The array is not a labeled array, but the `label` parameter is set. Use the `index` instead.")
        if hasattr(label, "to_numpy"):
            label = label.to_numpy()
        (hits,) = np.where(np.asarray(labels) == label)
        index = int(hits[0]) if len(hits) else None
    if _math._is_dask(data):
        shape = data.shape
        n = shape[0 if axis is None else axis]
        if index is not None and -n <= index < n:
            import dask.array as da

            return da.take(data, index, axis=axis)
        raw, rewrap = None, None
    else:
        raw, rewrap = _math._unwrap(data)
        shape = tuple(raw.shape)
        try:
            if index is None:
                raise IndexError
            return rewrap(_take(raw, index, axis))
        except IndexError:
            pass
    if not return_nodata:
        raise ArrayElementNotAvailable(
            f"The array has no element with the specified index or label: {index if index is not None else label}")
    ax = 0 if axis is None else axis
    out_shape = tuple(shape[:ax]) + tuple(shape[ax + 1:])
    if raw is not None and engine.is_torch(raw):
        import torch

        return rewrap(torch.full(out_shape, float("nan"), dtype=torch.float64, device=raw.device))
    if raw is None:
        import dask.array as da

        return da.full(out_shape, np.nan)
    return np.full(out_shape, np.nan)


def count(data, condition=None, context=None, axis=None, keepdims=False):
    if isinstance(data, list):
        data = np.asarray(data)
    if condition is None:
        if _math._is_dask(data):
            return _math._dask_blocks(data, lambda b: count(b, axis=axis), axis, dtype=np.int64)
        raw, rewrap = _math._unwrap(data)
        res = engine.reduce(raw, [L.OP_COUNT], axis)[L.OP_COUNT]
        return rewrap(_math._keepdims(res, tuple(raw.shape) or (1,), axis, keepdims))
    if condition is True:
        if _math._is_dask(data):
            return _math._dask_blocks(data, lambda b: count(b, condition=True, axis=axis), axis, dtype=data.dtype)
        raw, rewrap = _math._unwrap(data)
        in_dt = engine.np_dtype_of(raw)
        res = engine.reduce(raw, [L.OP_COUNT_ALL], axis)[L.OP_COUNT_ALL]
        if in_dt.kind in "iub":  # np.nansum(np.ones_like(int data)) is an integer sum
            res = engine._astype(res, np.dtype(np.uint64) if in_dt.kind == "u" else np.dtype(np.int64))
        return rewrap(_math._keepdims(res, tuple(raw.shape) or (1,), axis, keepdims))
    if callable(condition):
        context = dict(context) if context else {}
        context.pop("x", None)
        flags = condition(x=data, **context)
        return _math._sum(flags, ignore_nodata=True, axis=axis, keepdims=keepdims)
    raise Exception("Could not count values as condition is not callable. ")


def sort(data, asc=True, nodata=None, axis=None):
    """reference arrays.py:560-584: ``np.sort`` (ascending) or ``-np.sort(-data)`` (descending) along ``axis``;
    ``nodata=True`` leaves the NaNs last, ``False`` puts them first, ``None`` drops them (which flattens
    anything that is not one-dimensional, exactly like the reference's boolean indexing)."""
    if isinstance(data, list):
        data = np.asarray(data)
    if len(data) == 0:
        return data
    if _math._is_dask(data):
        data = data.compute()  # the reference loads the array as well (dask has no lazy sort)
    raw, rewrap = _math._unwrap(data)
    in_dt = engine.np_dtype_of(raw)
    if axis is None:  # np.sort(axis=None) sorts the flattened array
        raw = raw.reshape(-1)
        axis = 0
    res = engine.sort(raw, axis, descending=not asc, nan_first=(nodata is False))
    if in_dt.kind in "iub":
        res = engine._astype(res, in_dt)
    if in_dt.kind in "iu" and not asc:
        res = _wrapped_negation_first(res, in_dt, axis)
    if nodata is None:
        res = res[~res.isnan()] if engine.is_torch(res) else res[~np.isnan(res)]
    return rewrap(res)


def _wrapped_negation_first(res, in_dt, axis):
    """The reference sorts descending as ``-np.sort(-data)`` (arrays.py:571-574).  For integers the negation WRAPS:
    -0 is 0 for unsigned samples and -INT_MIN is INT_MIN for signed ones, so those values sort FIRST and come out in
    front of the descending rest ([0, 5, 3] uint16 -> [0, 5, 3]).  Reproduced by rotating every descending column by
    its number of such values (they sit at its end)."""
    wrapv = 0 if in_dt.kind == "u" else int(np.iinfo(in_dt).min)
    hit = res == wrapv
    if not bool(hit.any()):
        return res
    n = res.shape[axis]
    if engine.is_torch(res):
        nz = hit.sum(dim=axis, keepdim=True)
        shape = [1] * res.ndim
        shape[axis] = n
        j = engine.torch.arange(n, device=res.device).reshape(shape)
        return engine.torch.gather(res, axis, (j - nz) % n)
    nz = hit.sum(axis=axis, keepdims=True)
    shape = [1] * res.ndim
    shape[axis] = n
    return np.take_along_axis(res, (np.arange(n).reshape(shape) - nz) % n, axis=axis)


def _flip_all(a):
    if engine.is_torch(a):
        return a.flip(tuple(range(a.ndim)))
    return np.flip(a)


def _isnull(a):
    if engine.is_torch(a):
        return a.isnan() if a.is_floating_point() else a.new_zeros(a.shape, dtype=engine.torch.bool)
    return np.isnan(a) if a.dtype.kind == "f" else np.zeros(a.shape, dtype=bool)


def order(data, asc=True, nodata=None, axis=None):
    """reference arrays.py:498-534: ``np.argsort(data, kind="mergesort", axis)`` -- stable, NaNs last -- computed by
    ``eoc_order``; then the reference's post-processing, quirks included: ``asc=False`` is ``np.flip`` of the
    permutation over ALL axes (so for a 1-D array the NaNs come first and ties appear in reverse index order),
    ``nodata=None`` drops the positions of the NaNs (1-D only, same ValueError), ``nodata=False`` moves them to the
    front (1-D only: the reference's ``data[permutation]`` cannot index anything else), ``nodata=True`` returns the
    permutation.  A dask array is loaded first, as in the reference."""
    if isinstance(data, list):
        data = np.asarray(data)
    if len(data) == 0:
        return data
    if _math._is_dask(data):
        data = data.compute()
    raw, rewrap = _math._unwrap(data)
    if axis is None:  # np.argsort(axis=None) works on the flattened array
        perm = engine.order(raw.reshape(-1), 0)
    else:
        perm = engine.order(raw, axis)
    if not asc:
        perm = _flip_all(perm)
    if nodata is True:
        return rewrap(perm)
    if nodata is None and raw.ndim > 1:
        raise ValueError("order with nodata=None is not supported for arrays with more than one dimension, as this "
                         "would result in sparse multi-dimensional arrays.")
    if raw.ndim > 1:
        raise IndexError("too many indices for array: order with nodata=False indexes the data with the whole "
                         "permutation and only works on one dimension (reference arrays.py:525-529)")
    null = _isnull(engine.take(raw, perm.cpu().numpy() if engine.is_torch(perm) else perm, 0))
    if nodata is None:
        return rewrap(perm[~null])
    both = [perm[null], perm[~null]]
    return rewrap(engine.torch.cat(both) if engine.is_torch(perm) else np.append(*both))


def rearrange(data, order, axis=None, source_transposed_axis=None):
    """reference arrays.py:537-555: ``np.take(data, indices=order, axis=axis)`` for a one-dimensional ``order``
    (``eoc_take``: a gather of planes; ``axis=None`` on more than one dimension takes from the flattened array)."""
    if len(data) == 0:
        return data
    if isinstance(data, list):
        data = np.asarray(data)
    if len(data.shape) == 1 and axis is None:
        axis = 0
    if isinstance(order, list):
        order = np.asarray(order)
    if len(order.shape) != 1:
        raise ValueError(f"rearrange: order must be one-dimensional, but has {len(order.shape)} dimensions. ")
    if _math._is_dask(data):
        ax = axis
        src = data.rechunk({ax: -1}) if ax is not None else data.rechunk(data.shape)
        chunks = None
        if ax is not None:
            chunks = list(src.chunks)
            chunks[ax % data.ndim] = (len(order),)
        idx = np.asarray(order)
        return src.map_blocks(lambda b: rearrange(b, idx, axis=axis), dtype=data.dtype,
                              chunks=tuple(chunks) if chunks else ((len(order),),),
                              **({} if ax is not None else {"drop_axis": list(range(1, data.ndim))}))
    raw, rewrap = _math._unwrap(data)
    idx = order.tensor if hasattr(order, "tensor") else order
    idx = idx.cpu().numpy() if engine.is_torch(idx) else np.asarray(idx)
    if axis is None:
        return rewrap(engine.take(raw.reshape(-1), idx, 0))
    return rewrap(engine.take(raw, idx, axis))


def array_interpolate_linear(data, axis=None, dim_labels=None):
    """reference arrays.py:405-458: gaps (NaN / +-inf) inside each 1-D slice along ``axis`` are filled by
    ``np.interp`` between the neighbouring finite samples, gaps at the ends become NaN, slices with fewer than
    two finite samples stay as they are.  Sample positions: ``dim_labels`` when given (strings are read as
    datetimes, then as numbers, like the reference), else 0..n-1 -- which is what the reference uses under
    ``apply_dimension`` (it passes no labels).  Integer data has no gaps and is returned unchanged."""
    if isinstance(data, list):
        data = np.asarray(data, dtype=float)
    if _math._is_dask(data):
        ax = _math.normalize_axes(axis, data.ndim)[0] if axis is not None else 0
        return data.rechunk({ax: -1}).map_blocks(lambda b: array_interpolate_linear(b, axis=axis, dim_labels=dim_labels),
                                                 dtype=data.dtype)
    raw, rewrap = _math._unwrap(data)
    if engine.np_dtype_of(raw).kind != "f":
        return data
    positions = None
    if dim_labels is not None and len(dim_labels) > 0:
        x = np.array(dim_labels)
        if x.dtype.type is np.str_:
            try:
                x = np.array(x, dtype="datetime64").astype(float)
            except Exception:
                try:
                    x = np.array(x, dtype=float)
                except Exception:
                    x = np.arange(len(x))
        elif x.dtype.kind == "M":
            x = x.astype(float)
        positions = np.asarray(x, dtype=np.float64)
    if axis is None:
        if raw.ndim != 1:
            raise ValueError("array_interpolate_linear without an axis needs a one-dimensional array")
        axis = 0
    return rewrap(engine.interpolate_linear(raw, axis, positions))
