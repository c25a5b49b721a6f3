"""``reduce_dimension`` / ``reduce_spatial`` with the reference's signatures (``cubes/reduce.py``).

The work is ``data.reduce(reducer, dim=...)``: the cube hands ``reducer`` its array and the integer
axis; with the reducers of this package that is one CUDA launch over an [outer, R, inner] view.
"""

from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from ._xr_interop import dims_of
from ..exceptions import DimensionNotAvailable

__all__ = ["reduce_dimension", "reduce_spatial"]


def reduce_dimension(data, reducer: Callable, dimension: str, context: Optional[dict] = None):
    """reference cubes/reduce.py:13-46."""
    if dimension not in data.dims:
        raise DimensionNotAvailable(f"Provided dimension ({dimension}) not found in data.dims: {data.dims}")
    labels = data[dimension].values
    reduced = data.reduce(reducer, dim=dimension, keep_attrs=True, positional_parameters={"data": 0},
                          context=context, dim_labels=labels)
    # the reference records the smallest label of the dimension it removed (cubes/reduce.py:36-44)
    if "reduced_dimensions_min_values" not in data.attrs:
        reduced.attrs["reduced_dimensions_min_values"] = {}
    try:
        smallest = np.asarray(data.coords[dimension].values).min()
    except (TypeError, KeyError):
        smallest = 0
    reduced.attrs["reduced_dimensions_min_values"][dimension] = smallest
    return reduced


def reduce_spatial(data, reducer: Callable, context: Optional[dict] = None):
    """reference cubes/reduce.py:49-62: the reducer receives the tuple of spatial axes."""
    spatial = dims_of(data).spatial_dims or None
    return data.reduce(reducer, dim=spatial, keep_attrs=True, positional_parameters={"data": 0},
                       named_parameters={"context": context})


def reduce_dimension_many(data, reducers, dimension: str, ignore_nodata: bool = True):
    """EXTENSION (not a reference process): several reducers over the same dimension with ONE read of the
    cube -- what sibling ``reduce_dimension`` nodes over the same input collapse to.  ``reducers`` maps a
    result name to a process name or ``(name, kwargs)``; returns ``{result name: cube}``, every cube equal
    to ``reduce_dimension(data, <that reducer>, dimension)`` (a ``quantiles`` entry keeps its quantile
    axis at the position of ``dimension``, labelled by the probabilities)."""
    from ..math import _resolve_probabilities, reduce_many

    if dimension not in data.dims:
        raise DimensionNotAvailable(f"Provided dimension ({dimension}) not found in data.dims: {data.dims}")
    axis = list(data.dims).index(dimension)
    keys = list(reducers)
    arrays = reduce_many(data.data, [reducers[k] for k in keys], ignore_nodata=ignore_nodata, axis=axis)
    out = {}
    for k, arr in zip(keys, arrays):
        spec = reducers[k]
        name = (spec if isinstance(spec, str) else spec[0]).lstrip("_")
        if name == "quantiles":
            kw = {} if isinstance(spec, str) else spec[1]
            probs = _resolve_probabilities(kw.get("probabilities"), kw.get("q"))
            coords = {c: data.coords[c] for c in data.coords if c != dimension
                      and dimension not in getattr(data.coords[c], "dims", (getattr(data.coords[c], "dim", None),))}
            coords[dimension] = list(probs)
            out[k] = type(data)(arr, dims=data.dims, coords=coords, attrs=dict(data.attrs))
        else:
            out[k] = data.reduce(lambda block, axis=None, _a=arr, **kwargs: _a, dim=dimension, keep_attrs=True)
    return out
