"""Validity predicates and the order comparisons ``count(condition=...)`` is used with.

Counterparts of the reference's ``comparison.py:24-33`` (is_infinite / is_valid) and ``:103-121``
(gt / gte / lt / lte): comparisons return float arrays holding 1.0 / 0.0 and NaN wherever an operand
is nodata, exactly what ``np.where(notnull(x) & notnull(y), x > y, np.nan)`` yields there.  These are
elementwise helpers, not reductions; on device arrays they are evaluated with torch (plumbing), the
reduction that follows (``count``) runs in the CUDA library.
"""

from __future__ import annotations

import numpy as np

from .. import engine
from ..device_array import DeviceArray

__all__ = ["is_infinite", "is_valid", "is_nan", "gt", "gte", "lt", "lte"]


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
    rx, ry = _raw(x), _raw(y)
    if engine.is_torch(rx) or engine.is_torch(ry):
        import torch

        t = rx if engine.is_torch(rx) else ry
        rx = rx if engine.is_torch(rx) else torch.as_tensor(rx, device=t.device)
        ry = ry if engine.is_torch(ry) else torch.as_tensor(ry, device=t.device)
        ok = ~_isnull(rx) & ~_isnull(ry)
        res = torch.where(ok, op(rx, ry).to(torch.float64), torch.full((), float("nan"), dtype=torch.float64,
                                                                        device=t.device))
        like = x if isinstance(x, DeviceArray) else y
        return _rewrap(like, res)
    ok = np.logical_and(~_isnull(rx), ~_isnull(ry))
    return np.where(ok, op(rx, ry), np.nan)


def gt(x, y):
    return _compare(x, y, lambda a, b: a > b)


def gte(x, y):
    return _compare(x, y, lambda a, b: (a - b) >= 0)


def lt(x, y):
    return _compare(x, y, lambda a, b: a < b)


def lte(x, y):
    return _compare(x, y, lambda a, b: a <= b)
