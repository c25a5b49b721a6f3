"""Validity predicates and the order comparisons ``count(condition=...)`` is used with.

Counterparts of the reference's ``comparison.py:24-33`` (is_infinite / is_valid) and ``:103-121``
(eq / neq / gt / gte / lt / lte): comparisons return float64 arrays holding 1.0 / 0.0 and NaN wherever an
operand is nodata, exactly what ``np.where(notnull(x) & notnull(y), x > y, np.nan)`` yields there; one launch
of ``eoc_binary`` each.  The validity predicates are elementwise helpers evaluated with torch (plumbing); the
reduction that follows (``count``) runs in the CUDA library.
"""

from __future__ import annotations

import numpy as np

from .. import _lib as L
from .. import engine
from ..device_array import DeviceArray

__all__ = ["is_infinite", "is_valid", "is_nan", "eq", "neq", "gt", "gte", "lt", "lte"]


def _raw(x):
    return x.tensor if isinstance(x, DeviceArray) else x


def _rewrap(like, res):
    return DeviceArray(res) if isinstance(like, DeviceArray) else res


def _isnull(x):
    if engine.is_torch(x):
        import torch

        return torch.isnan(x) if x.is_floating_point() else torch.zeros_like(x, dtype=torch.bool)
    import pandas as pd

    return pd.isnull(x)


def is_nan(x):
    if x is None:
        return True
    r = _raw(x)
    return _rewrap(x, _isnull(r)) if engine.is_torch(r) else np.isnan(r)


def is_infinite(x):
    r = _raw(x)
    if engine.is_torch(r):
        import torch

        return _rewrap(x, torch.isinf(r) if r.is_floating_point() else torch.zeros_like(r, dtype=torch.bool))
    if np.isscalar(r) or hasattr(r, "dtype"):
        kind = np.dtype(type(r)).type if np.isscalar(r) else np.dtype(r.dtype).type
        if np.issubdtype(kind, np.number):
            return np.isinf(r)
    return False


def is_valid(x):
    r = _raw(x)
    inf = _raw(is_infinite(x))
    if engine.is_torch(r):
        return _rewrap(x, ~_isnull(r) & ~inf)
    return np.logical_and(~_isnull(r), np.logical_not(inf))


def _compare(x, y, op):
    """np.where(notnull(x) & notnull(y), x OP y, nan) in one launch of ``eoc_binary`` (float64 1 / 0 / NaN)."""
    from .math import _binary

    return _binary(op, x, y)


def eq(x, y, delta=None, case_sensitive=True):
    """reference comparison.py:45-87, numeric / boolean operands.  ``delta`` (np.isclose with atol) and string
    operands are host-side metadata comparisons in the reference and are not part of the device path."""
    if delta:
        raise NotImplementedError("eq with delta is not part of the B200 path")
    return _compare(x, y, L.BIN_EQ)


def neq(x, y, delta=None, case_sensitive=True):
    """reference comparison.py:90-100."""
    if delta:
        raise NotImplementedError("neq with delta is not part of the B200 path")
    return _compare(x, y, L.BIN_NEQ)


def gt(x, y):
    return _compare(x, y, L.BIN_GT)


def gte(x, y):
    """reference comparison.py:117-119 -- ``(x - y) >= 0``."""
    return _compare(x, y, L.BIN_GTE)


def lt(x, y):
    return _compare(x, y, L.BIN_LT)


def lte(x, y):
    return _compare(x, y, L.BIN_LTE)
