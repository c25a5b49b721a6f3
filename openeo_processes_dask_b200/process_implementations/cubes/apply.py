"""``apply`` / ``apply_dimension`` with the reference's signatures (``cubes/apply.py:24-128``).

``apply_dimension`` hands the cube's array and the integer axis of ``dimension`` to ``process`` (the
reference goes through ``xr.apply_ufunc``, which moves the core dimension to the end first and back
afterwards -- the array processes of this package take any axis, so nothing is transposed) and rebuilds the
cube around the result: the dimension may come back with a different length (``quantiles`` with several
probabilities, ``sort`` with nodata dropped), in which case it is relabelled 0..n-1 exactly like the
reference; ``target_dimension`` renames it.
"""

from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from ...lazy_elementwise import LazyElementwise
from ..exceptions import DimensionNotAvailable
from .cube import Coordinate, DataCube

__all__ = ["apply", "apply_dimension"]


def apply(data, process: Callable, context: Optional[dict] = None):
    """reference cubes/apply.py:24-39: ``process`` applied to the whole array (elementwise processes).  On a
    DeviceArray-backed float cube (or a ``LazyCast``) the chain of elementwise processes in ``process`` is fused into
    one launch of ``eoc_apply_tape``; the values are those of the node-by-node evaluation."""
    x = data.data
    if LazyElementwise.wraps(x):
        # the child process graph extends ONE expression; it runs as a single kernel below (lazy_elementwise.py)
        x = LazyElementwise(x)
    out = process(x, positional_parameters={"x": 0}, named_parameters={"context": context})
    if isinstance(out, LazyElementwise):
        out = out.materialize()
    if tuple(out.shape) != data.shape:
        raise ValueError(f"apply: the process changed the shape {data.shape} -> {tuple(out.shape)}")
    return DataCube(out, data.dims, data.coords, data.attrs, data.name)


def apply_dimension(data, process: Callable, dimension: str, target_dimension: Optional[str] = None,
                    context: Optional[dict] = None):
    """reference cubes/apply.py:42-128."""
    if context is None:
        context = {}
    if dimension not in data.dims:
        raise DimensionNotAvailable(f"Provided dimension ({dimension}) not found in data.dims: {data.dims}")
    keepdims = target_dimension is not None
    if target_dimension is None:
        target_dimension = dimension
    axis = data.dims.index(dimension)
    labels = data[dimension].values
    out = process(data.data, positional_parameters={"data": 0}, named_parameters={"context": context}, axis=axis,
                  keepdims=keepdims, source_transposed_axis=axis, context=context)  # no dim_labels, like the reference
    if len(tuple(out.shape)) == data.ndim - 1:  # the process reduced the dimension: one value per pixel
        out = out.reshape(tuple(out.shape[:axis]) + (1,) + tuple(out.shape[axis:]))
    if len(tuple(out.shape)) != data.ndim:
        raise ValueError(f"apply_dimension: the process returned {len(tuple(out.shape))} dimensions for a "
                         f"{data.ndim}-dimensional cube")
    n = int(out.shape[axis])
    coords = {k: c for k, c in data.coords.items() if c.dim != dimension}
    dims = list(data.dims)
    if dimension == target_dimension:
        # labels preserved when the number of values is unchanged, else 0..n-1 (cubes/apply.py:96-103)
        new_labels = labels if n == len(labels) else np.arange(n)
        coords[dimension] = Coordinate(dimension, dimension, np.asarray(new_labels))
    elif target_dimension in data.dims:
        # the target exists: allowed only with a single label, which is dropped (cubes/apply.py:104-117)
        t_ax = data.dims.index(target_dimension)
        if data.shape[t_ax] != 1:
            raise Exception(
                f"Cannot rename dimension {dimension} to {target_dimension} as {target_dimension} already exists in "
                f"dataset and contains more than one label: {data[target_dimension].values}. See process definition. ")
        idx = [slice(None)] * out.ndim
        idx[t_ax] = 0
        out = out[tuple(idx)]
        dims.pop(t_ax)
        coords = {k: c for k, c in coords.items() if c.dim != target_dimension}
        dims[dims.index(dimension)] = target_dimension
        coords[target_dimension] = Coordinate(target_dimension, target_dimension, np.arange(n))
    else:
        dims[axis] = target_dimension
        coords[target_dimension] = Coordinate(target_dimension, target_dimension, np.arange(n))
    return DataCube(out, tuple(dims), coords, data.attrs, data.name)
