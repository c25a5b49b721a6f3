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
