"""``mask`` with the reference's signature (``cubes/mask.py:21-118``), fused into the reduce that follows.

The reference computes ``data.where(_not(mask), replacement)``: a sample is replaced where the mask is truthy
and not NaN (``_not`` keeps NaN, which ``where`` treats as "keep"), after checking that the temporal labels
agree and the spatial dims match.  Here the mask becomes a canonical uint8 cube (``eoc_mask_canonical``) that
broadcasts over the dims it lacks with stride 0, and

* for a device cube the result holds a ``LazyMasked``: a following ``reduce_dimension`` / ``aggregate`` with a
  reducer of this package consumes mask and data in ONE pass (cloud mask -> NaN -> temporal median: the
  Sentinel-2 composite); anything else that touches it materialises it by one elementwise launch
  (``eoc_masked_reduce`` with no reducers) -- the same bits either way;
* numpy in -> numpy out: the masked cube is materialised (``fuse=False`` forces that for device cubes too).

Not implemented (out of scope): implicit ``resample_cube_spatial`` of a mask on a different grid
(``cubes/mask.py:84-86``) -- differing spatial labels raise ``NotImplementedError``.
"""

from __future__ import annotations

import numpy as np

from ... import engine
from ...device_array import DeviceArray, LazyMasked
from ..exceptions import DimensionLabelCountMismatch, DimensionMismatch, LabelMismatch
from ._xr_interop import dims_of
from .cube import DataCube

__all__ = ["mask"]


def _labels_equal(a, b) -> bool:
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and bool(np.all(a == b))


def mask(data, mask, replacement=None, fuse=None):
    if replacement is None:
        replacement = np.nan
    d, m = dims_of(data), dims_of(mask)
    # temporal dims must agree (names and labels) unless the mask has none (cubes/mask.py:31-55)
    check_labels = set(d.temporal_dims) == set(m.temporal_dims)
    if not check_labels and len(m.temporal_dims) != 0:
        raise DimensionMismatch(
            f"data and mask temporal dimensions do no match: data has temporal dimensions ({d.temporal_dims}) "
            f"and mask {m.temporal_dims}.")
    if check_labels:
        for n in d.temporal_dims:
            a, b = data[n].values, mask[n].values
            if len(a) != len(b):
                raise DimensionLabelCountMismatch(
                    f"data and mask temporal dimensions do no match: data has {len(a)} temporal dimensions labels "
                    f"and mask {len(b)}.")
            if not _labels_equal(a, b):
                raise LabelMismatch(f"data and mask temporal dimension labels don't match for dimension {n}.")
    if set(d.spatial_dims) != set(m.spatial_dims):
        raise DimensionMismatch(
            f"data and mask spatial dimensions do no match: data has spatial dimensions ({d.spatial_dims}) "
            f"and mask {m.spatial_dims}")
    for n in d.spatial_dims:
        if not _labels_equal(data[n].values, mask[n].values):
            raise NotImplementedError(
                f"mask on a different spatial grid (dimension {n}): resample_cube_spatial is out of scope here")
    extra = [n for n in mask.dims if n not in data.dims]
    if extra:
        raise DimensionMismatch(f"mask has dimensions the data lacks: {extra}")

    # output dim order: the reference moves (bands | t, t | bands, y, x) to the front for the broadcast and
    # moves them back only in the bands case (cubes/mask.py:91-116) -- reproduced
    out_dims = data.dims
    if not (len(d.band_dims) != 0 and len(m.band_dims) == 0) and (len(d.temporal_dims) != 0 and len(m.temporal_dims) == 0):
        order = [x for x in (d.temporal_dims[:1] + d.band_dims[:1] + (d.y_dim, d.x_dim)) if x in data.dims]
        if set(order) == set(data.dims):
            out_dims = tuple(order)
    if out_dims != data.dims:
        data = data.transpose(*out_dims)

    # the mask as a canonical uint8 device cube in the data's dim order, size-1 where it lacks a dim
    mraw = mask.data.tensor if isinstance(mask.data, DeviceArray) else mask.data
    canon = engine.mask_canonical(mraw)
    perm = [mask.dims.index(n) for n in data.dims if n in mask.dims]
    canon = canon.permute(*perm) if perm != list(range(len(perm))) else canon
    shape = [data.shape[i] if n in mask.dims else 1 for i, n in enumerate(data.dims)]
    canon = canon.reshape(shape)

    raw = data.data
    if hasattr(raw, "materialize"):  # the result of a previous lazy ndvi / mask
        raw = raw.materialize()
    was_numpy = not (isinstance(raw, DeviceArray) or engine.is_torch(raw))
    if was_numpy:
        import torch

        raw = torch.from_numpy(np.ascontiguousarray(engine._native(np.asarray(raw)))).cuda()
    lazy = LazyMasked(raw, canon, replacement)
    if fuse is None:
        fuse = True
    if fuse and not was_numpy:
        values = lazy
    else:
        values = lazy.materialize()
        if was_numpy:
            values = values.cpu().numpy()
    return DataCube(values, data.dims, data.coords, data.attrs, data.name)
