"""Cube-level glue of the reference restated on plain arrays (TEST INFRASTRUCTURE).

A cube here is ``(array, dims, coords)``: a numpy array, a tuple of dimension
names and a dict ``dim -> 1-D label array``.  The reference delegates this
orchestration to xarray, which is not installable here, so the xarray
behaviours it relies on are restated and marked [U] (upstream behaviour):

* ``DataArray.reduce(f, dim=d)`` calls ``f(data, axis=dims.index(d))`` and
  drops ``d``                                          [U]  cubes/reduce.py:26-34
* ``groupby_bins(t, bins=edges)`` == ``pandas.cut(t, edges)``: right-closed
  ``(a, b]`` bins, values outside every bin dropped, groups reduced in original
  index order, empty bins re-indexed to NaN            [U]  cubes/aggregate.py:92-96
"""

from __future__ import annotations

import numpy as np
import pandas as pd

from . import reducers


def reduce_dimension(array, dims, coords, reducer, dimension, **reducer_kwargs):
    """cubes/reduce.py:13-46.  Returns (array, dims, attrs_update)."""
    if dimension not in dims:
        raise KeyError("DimensionNotAvailable")
    axis = dims.index(dimension)
    out = reducer(array, axis=axis, **reducer_kwargs)
    out_dims = tuple(d for d in dims if d != dimension)
    attrs = {"reduced_dimensions_min_values": {dimension: np.asarray(coords[dimension]).min()}}
    return out, out_dims, attrs


def reduce_spatial(array, dims, reducer, spatial_dims=("x", "y"), **reducer_kwargs):
    """cubes/reduce.py:49-62: reducer gets a tuple axis over the spatial dims."""
    axes = tuple(dims.index(d) for d in spatial_dims if d in dims)
    out = reducer(array, axis=axes, **reducer_kwargs)
    return out, tuple(d for d in dims if d not in spatial_dims)


def temporal_bins(intervals, labels=None):
    """cubes/aggregate.py:66-88: edges (float seconds), all-bin labels, picked labels.

    Returns (edges, labels_all_bins, labels) where ``labels_all_bins[i]`` names bin
    ``(edges[i], edges[i+1]]`` (filler names for the gaps between non-adjacent
    intervals) and ``labels`` are the ones selected into the output, in order.
    """
    iv = np.array(intervals, dtype=np.datetime64).astype("datetime64[s]").astype(float)
    flat = iv.reshape(iv.shape[0] * iv.shape[1])
    if not labels:
        labels = np.array(intervals, dtype="datetime64[s]").astype(str)[:, 0]
    if (iv[1:, 0] < iv[:-1, 1]).any():
        raise NotImplementedError("overlapping time ranges")
    mask = np.zeros((len(labels) * 2) - 2).astype(bool)
    mask[1::2] = np.isin(iv[1:, 0], iv[:-1, 1])
    mask = np.append(mask, np.array([False, True]))
    labels_nans = np.arange(len(labels) * 2).astype(str)
    labels_nans[::2] = labels
    labels_nans = labels_nans[~mask]
    edges = np.unique(flat)
    return edges, labels_nans, np.asarray(labels)


def bin_codes(t_values, edges):
    """Right-closed bin index per timestamp (-1 = in no bin); pandas.cut semantics [U]."""
    t = np.array(np.asarray(t_values).astype(str), dtype="datetime64[s]").astype(float)
    return np.asarray(pd.cut(t, edges).codes, dtype=np.int64)


def aggregate_temporal(array, dims, coords, intervals, reducer, labels=None, dimension="t",
                       **reducer_kwargs):
    """cubes/aggregate.py:29-101 on plain arrays.  Output keeps the t axis position."""
    axis = dims.index(dimension)
    edges, labels_all, labels_sel = temporal_bins(intervals, labels)
    codes = bin_codes(coords[dimension], edges)
    moved = np.moveaxis(np.asarray(array), axis, 0)
    per_bin = []
    for b in range(len(edges) - 1):
        members = np.nonzero(codes == b)[0]
        if len(members) == 0:  # xarray re-indexes missing groups to NaN [U]
            probe = reducer(moved[:1], axis=0, **reducer_kwargs)
            per_bin.append(np.full(probe.shape, np.nan, dtype=np.result_type(probe.dtype, np.float32)))
        else:
            per_bin.append(np.asarray(reducer(moved[members], axis=0, **reducer_kwargs)))
    label_to_bin = {lab: i for i, lab in enumerate(labels_all)}
    picked = [per_bin[label_to_bin[str(lab)]] for lab in labels_sel]
    out = np.moveaxis(np.stack(picked, axis=0), 0, axis)
    out_coords = dict(coords)
    out_coords[dimension] = np.asarray([str(l) for l in labels_sel])
    return out, dims, out_coords


def ndvi(array, dims, coords, nir="nir", red="red", band_dim="bands", common_names=None):
    """cubes/indices.py:15-59: pick bands by label (else by common_name), (nir-red)/(nir+red)."""
    bands = list(coords[band_dim])
    axis = dims.index(band_dim)

    def pick(name):
        if name in bands:
            return bands.index(name)
        if common_names is not None and name in list(common_names):
            return list(common_names).index(name)
        raise KeyError(name)

    a = np.take(array, pick(nir), axis=axis)
    b = np.take(array, pick(red), axis=axis)
    return reducers.normalized_difference(a, b), tuple(d for d in dims if d != band_dim)


def mask(array, dims, mask_array, mask_dims, replacement=None):
    """cubes/mask.py:21-118 restated on numpy: ``data.where(_not(mask), replacement)``.

    ``_not`` (logic.py:51-54) is ``where(notnull(x), logical_not(x), nan)``; ``DataArray.where(cond, other)``
    keeps data where ``cond`` is truthy -- and NaN is truthy -- so a sample is replaced exactly where the mask
    is truthy and not NaN.  The mask broadcasts by dimension NAME over the dims it lacks.  Result dims: the
    data's, except that a mask without the temporal dim (and not the bands case) leaves the cube in
    (t, bands, y, x) order -- the reference transposes back only in the bands case (cubes/mask.py:91-116).
    dtype as xarray's where: floats keep theirs, integers are promoted to float64 by a NaN replacement.
    PARITY UNPINNED by reference tests (tests/test_mask.py:40-87 checks NaN counts and shapes only).
    Returns (array, dims).
    """
    array, mask_array = np.asarray(array), np.asarray(mask_array)
    dims, mask_dims = tuple(dims), tuple(mask_dims)
    if replacement is None:
        replacement = np.nan
    has_t, has_b = "t" in dims, "bands" in dims
    if not (has_b and "bands" not in mask_dims) and (has_t and "t" not in mask_dims):
        order = tuple(d for d in ("t", "bands", "y", "x") if d in dims)
        if set(order) == set(dims):
            array = np.transpose(array, [dims.index(d) for d in order])
            dims = order
    m = np.transpose(mask_array, [mask_dims.index(d) for d in dims if d in mask_dims])
    m = m.reshape([array.shape[i] if d in mask_dims else 1 for i, d in enumerate(dims)])
    with np.errstate(invalid="ignore"):
        replace = (m != 0) & ~(np.isnan(m) if m.dtype.kind == "f" else np.zeros(m.shape, bool))
    out = np.where(replace, replacement, array)
    if array.dtype.kind == "f":
        out = out.astype(array.dtype)
    elif replacement == replacement and float(replacement).is_integer():
        out = out.astype(array.dtype)
    return out, dims
