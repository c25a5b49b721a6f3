"""``ndvi`` with the reference's signature (``cubes/indices.py:15-59``).

Band resolution follows the reference: ``nir`` / ``red`` are looked up among the labels of the bands
dimension first and among a ``common_name`` coordinate second; failures raise the same exception
types.  The arithmetic ``(nir - red) / (nir + red)`` is ``normalized_difference`` (one elementwise
kernel pass).  For device cubes the result holds a ``LazyNormalizedDifference``: a following
``reduce_dimension`` / ``aggregate_temporal`` with a streaming reducer runs the fused kernel ``eoc_nd_reduce``
(read two band planes, write one image) instead of materialising the NDVI cube first; anything else that
touches the lazy array materialises it with the same kernel -- same bits either way (``fuse=False`` evaluates
eagerly; tests compare the two).  Integer cubes follow the reference (numpy wrap-around, float64); a cube cast
with ``cube.astype("float32")`` (``LazyCast``) is converted in the kernel's load stage.
"""

from __future__ import annotations

import numpy as np

from ...device_array import LazyNormalizedDifference
from ..exceptions import BandExists, DimensionAmbiguous, NirBandAmbiguous, RedBandAmbiguous
from ..math import normalized_difference
from ._xr_interop import dims_of
from .cube import DataCube

__all__ = ["ndvi"]


def _has(coord, label) -> bool:
    return bool(np.any(np.asarray(coord.values) == label))


def ndvi(data, nir="nir", red="red", target_band=None, fuse=None):
    band_dims = dims_of(data).band_dims
    if len(band_dims) == 0:
        raise DimensionAmbiguous("Dimension of type `bands` is not available or is ambiguous.")
    band_dim = band_dims[0]
    bands = data.coords[band_dim]
    common = data.coords["common_name"] if "common_name" in data.coords else None

    def where(label, error):
        if _has(bands, label):
            return band_dim
        if common is None:
            raise KeyError(f"{label!r} is neither a label of {band_dim!r} nor a common_name")
        if not _has(common, label):
            raise error
        return "common_name"

    nir_dim = where(nir, NirBandAmbiguous(
        "The NIR band can't be resolved, please specify the specific NIR band name."))
    red_dim = where(red, RedBandAmbiguous(
        "The Red band can't be resolved, please specify the specific Red band name."))
    nir_band = data.sel({nir_dim: nir})
    red_band = data.sel({red_dim: red})

    if fuse is None:
        fuse = True
    if fuse and target_band is None and LazyNormalizedDifference.supports(nir_band.data, red_band.data):
        values = LazyNormalizedDifference(nir_band.data, red_band.data)
    else:
        values = normalized_difference(nir_band.data, red_band.data)

    if isinstance(data, DataCube):
        nd = DataCube(values, nir_band.dims, {k: c for k, c in nir_band.coords.items() if c.dim != band_dim},
                      {}, data.name)
    else:  # pragma: no cover - real xarray objects
        nd = nir_band.copy(data=values)
    if target_band is not None:
        if target_band in data.coords or target_band in data.dims:
            raise BandExists("A band with the specified target name exists.")
        nd = _append_band(data, nd, band_dim, target_band)
    nd.attrs = data.attrs
    return nd


def _append_band(data, nd, band_dim, target_band):
    """The reference merges the index as an extra band labelled ``target_band`` (indices.py:53-57)."""
    if not isinstance(data, DataCube):  # pragma: no cover
        import xarray as xr

        extra = nd.expand_dims(band_dim).assign_coords({band_dim: [target_band]})
        return xr.concat([data.astype(extra.dtype), extra.transpose(*data.dims)], dim=band_dim)
    axis = data.dims.index(band_dim)
    raw, new = data.data, nd.data
    unwrap = getattr(raw, "tensor", None)
    if unwrap is not None:
        raw, new = raw.tensor, new.tensor
    if hasattr(raw, "unsqueeze"):
        import torch

        merged = torch.cat([raw.to(new.dtype), new.unsqueeze(axis)], dim=axis)
        if unwrap is not None:
            from ...device_array import DeviceArray

            merged = DeviceArray(merged)
    else:
        merged = np.concatenate([np.asarray(raw, dtype=new.dtype), np.expand_dims(new, axis)], axis=axis)
    coords = dict(data.coords)
    coords[band_dim] = np.append(np.asarray(data.coords[band_dim].values, dtype=object), target_band)
    coords = {k: c for k, c in coords.items() if k == band_dim or getattr(c, "dim", k) != band_dim}
    return DataCube(merged, data.dims, coords, {}, data.name)
