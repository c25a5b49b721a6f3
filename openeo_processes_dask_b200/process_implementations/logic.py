"""``_any`` / ``_all`` with the reference's signatures (reference ``logic.py:58-87``).

* ``_any``: ``np.any(np.nan_to_num(data), axis)`` -- NaN counts as False, +-inf as True.
* ``_all``: ``np.all(data, axis)`` -- NaN is truthy (numpy's rule), only a zero makes it False.
* ``ignore_nodata=False``: the result becomes a float array (``np.where(mask, bool, nan)``) that is NaN where
  the column holds a NaN which could still change the outcome (any: no True seen; all: no False seen).

One launch of the column kernel (EOC_OP_ANY / EOC_OP_ALL) produces 0 / 1 / NaN; this layer only restores the
result dtype numpy would give: bool with ``ignore_nodata=True``, float64 otherwise.
"""

from __future__ import annotations

import numpy as np

from .. import _lib as L
from .. import engine
from . import math as _math

__all__ = ["_any", "_all", "_and", "_or", "xor"]


def _logic(op, data, ignore_nodata, axis):
    if isinstance(data, list):
        data = np.asarray(data)
    if len(data) == 0:
        return np.nan
    out_dt = np.dtype(np.bool_) if ignore_nodata else np.dtype(np.float64)
    if _math._is_dask(data):
        return _math._dask_blocks(data, lambda b: _logic(op, b, ignore_nodata, axis), axis, dtype=out_dt)
    raw, rewrap = _math._unwrap(data)
    if engine.np_dtype_of(raw) == np.bool_:
        raw = engine._astype(raw, np.dtype(np.uint8))
    res = engine.reduce(raw, [op], axis, ignore_nodata=bool(ignore_nodata))[op]
    return rewrap(engine._astype(res, out_dt))


def _any(data, ignore_nodata=True, axis=None):
    """reference logic.py:58-71."""
    return _logic(L.OP_ANY, data, ignore_nodata, axis)


def _all(data, ignore_nodata=True, axis=None):
    """reference logic.py:74-87."""
    return _logic(L.OP_ALL, data, ignore_nodata, axis)


# ---- three-valued and / or / xor (reference logic.py:11-40), the usual merge_cubes overlap resolvers on masks ----
def _and(x, y):
    """reference logic.py:11-19 -- NaN only where the other operand is true: and(NaN, False) is False."""
    return _math._binary(L.BIN_AND, x, y)


def _or(x, y):
    """reference logic.py:22-32 -- NaN only where the other operand is false: or(NaN, True) is True."""
    return _math._binary(L.BIN_OR, x, y)


def xor(x, y):
    """reference logic.py:35-40 -- NaN wherever an operand is NaN."""
    return _math._binary(L.BIN_XOR, x, y)
