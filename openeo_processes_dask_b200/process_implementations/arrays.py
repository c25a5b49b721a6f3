"""``count`` with the reference's signature (reference ``arrays.py:587-606``).

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

__all__ = ["count"]


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
