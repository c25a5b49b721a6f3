"""DataCube: the small part of ``xarray.DataArray`` the reducer path touches.

xarray is not installable in the build environment, so the cube-level processes are written against
this duck type and accept a real ``xarray.DataArray`` wherever one is available (same attribute
names and call conventions): ``dims``, ``data``, ``values``, ``shape``, ``coords``, ``attrs``,
``cube[dim].values``, ``cube.reduce(func, dim=..., keep_attrs=True, **kwargs)`` which calls
``func(data, axis=<int or tuple>, **kwargs)`` like ``Variable.reduce`` does, ``sel({dim: label})``,
``rename``, ``assign_coords``, ``expand_dims`` and the ``.openeo`` dimension accessor.

``data`` may be a numpy array, a CUDA torch tensor or a ``DeviceArray``; it is never copied or moved
by the cube itself.
"""

from __future__ import annotations

import copy as _copy

import numpy as np

from ._xr_interop import OpenEODims


class Coordinate:
    """``cube[name]`` / ``cube.coords[name]``: labels along one dimension."""

    def __init__(self, name, dim, values):
        self.name = name
        self.dim = dim
        self.values = np.asarray(values)

    @property
    def data(self):
        return self.values

    def __len__(self):
        return len(self.values)

    def __iter__(self):
        return iter(self.values)

    def __contains__(self, label):
        return bool(np.any(self.values == label))

    def __array__(self, dtype=None, copy=None):
        return self.values if dtype is None else self.values.astype(dtype)

    def min(self):
        return self.values.min()

    def __repr__(self):
        return f"Coordinate({self.name!r} on {self.dim!r}, {self.values!r})"


class DataCube:
    def __init__(self, data, dims, coords=None, attrs=None, name=None):
        self.data = data
        self.dims = tuple(dims)
        if len(self.dims) != len(tuple(data.shape)):
            raise ValueError(f"dims {self.dims} do not match data of shape {tuple(data.shape)}")
        self.coords = {}
        for key, val in (coords or {}).items():
            if isinstance(val, Coordinate):
                self.coords[key] = Coordinate(key, val.dim, val.values)
            elif isinstance(val, tuple) and len(val) == 2 and isinstance(val[0], str):
                self.coords[key] = Coordinate(key, val[0], val[1])  # non-index coord: (dim, labels)
            else:
                self.coords[key] = Coordinate(key, key, val)
        for key, c in self.coords.items():
            if c.dim in self.dims and len(c) != self.shape[self.dims.index(c.dim)]:
                raise ValueError(f"coordinate {key!r} has {len(c)} labels for a dimension of {self.shape}")
        self.attrs = dict(attrs or {})
        self.name = name

    # -- basics -----------------------------------------------------------------------------------
    @property
    def shape(self):
        return tuple(int(s) for s in self.data.shape)

    @property
    def ndim(self):
        return len(self.dims)

    @property
    def dtype(self):
        return self.data.dtype

    @property
    def values(self):
        return np.asarray(self.data.cpu() if hasattr(self.data, "cpu") else self.data)

    @property
    def openeo(self):
        return OpenEODims(self.dims)

    def __getitem__(self, key):
        if isinstance(key, str):
            if key in self.coords:
                return self.coords[key]
            if key in self.dims:  # a dimension without labels: positional index
                return Coordinate(key, key, np.arange(self.shape[self.dims.index(key)]))
            raise KeyError(key)
        raise TypeError("DataCube supports cube['dim'] only; use .sel / .isel for data access")

    def __repr__(self):
        return f"DataCube(dims={self.dims}, shape={self.shape}, data={type(self.data).__name__})"

    def copy(self, data=None):
        return DataCube(self.data if data is None else data, self.dims, self.coords, _copy.copy(self.attrs),
                        self.name)

    # -- what the processes call ---------------------------------------------------------------------
    def reduce(self, func, dim=None, keep_attrs=False, **kwargs):
        """``func(self.data, axis=..., **kwargs)`` over ``dim`` (a name, a sequence of names or None)."""
        if dim is None:
            names = self.dims
        elif isinstance(dim, str):
            names = (dim,)
        else:
            names = tuple(dim)
        for n in names:
            if n not in self.dims:
                raise ValueError(f"{n!r} not found in array dimensions {self.dims}")
        axes = tuple(self.dims.index(n) for n in names)
        axis = axes[0] if isinstance(dim, str) else axes
        out = func(self.data, axis=axis, **kwargs)
        kept = tuple(d for d in self.dims if d not in names)
        if tuple(out.shape) != tuple(s for d, s in zip(self.dims, self.shape) if d not in names):
            raise ValueError(
                f"reducer returned shape {tuple(out.shape)}; expected the cube without {names} "
                "(a reducer that adds an axis, e.g. quantiles with several probabilities, must be "
                "followed by array_element or used through apply_dimension, as in the reference)")
        coords = {k: c for k, c in self.coords.items() if c.dim in kept}
        return DataCube(out, kept, coords, dict(self.attrs) if keep_attrs else {}, self.name)

    def isel(self, indexers):
        data, dims, coords = self.data, list(self.dims), dict(self.coords)
        for dim, idx in indexers.items():
            ax = dims.index(dim)
            key = [slice(None)] * len(dims)
            key[ax] = idx
            data = data[tuple(key)]
            if np.ndim(idx) == 0 and not isinstance(idx, slice):
                coords = {k: c for k, c in coords.items() if c.dim != dim}
                dims.pop(ax)
            else:
                coords = {k: (Coordinate(k, c.dim, c.values[idx]) if c.dim == dim else c) for k, c in coords.items()}
        return DataCube(data, dims, coords, self.attrs, self.name)

    def sel(self, indexers):
        """Label-based selection of single labels; ``{coord_name: label}`` may name a non-index coord."""
        out = self
        for key, label in indexers.items():
            coord = out[key]
            hits = np.nonzero(coord.values == label)[0]
            if len(hits) == 0:
                raise KeyError(f"{label!r} not found on {key!r}")
            out = out.isel({coord.dim: int(hits[0])})
        return out

    def rename(self, mapping):
        if not isinstance(mapping, dict):
            return DataCube(self.data, self.dims, self.coords, self.attrs, mapping)
        dims = tuple(mapping.get(d, d) for d in self.dims)
        coords = {mapping.get(k, k): Coordinate(mapping.get(k, k), mapping.get(c.dim, c.dim), c.values)
                  for k, c in self.coords.items()}
        return DataCube(self.data, dims, coords, self.attrs, self.name)

    def assign_coords(self, coords=None, **kw):
        merged = dict(self.coords)
        for key, val in {**(coords or {}), **kw}.items():
            if isinstance(val, tuple) and len(val) == 2 and isinstance(val[0], str):
                merged[key] = Coordinate(key, val[0], val[1])
            else:
                merged[key] = Coordinate(key, key, val)
        return DataCube(self.data, self.dims, merged, self.attrs, self.name)

    def drop_vars(self, name):
        return DataCube(self.data, self.dims, {k: c for k, c in self.coords.items() if k != name}, self.attrs,
                        self.name)

    def astype(self, dtype):
        """``DataArray.astype``.  A uint8 / int16 / uint16 device cube cast to float32 becomes a ``LazyCast``: the
        conversion happens in the load stage of the kernel that consumes it (e.g. ``ndvi`` -> reduce)."""
        data = self.data
        if hasattr(data, "materialize"):
            data = data.materialize()
        if hasattr(data, "astype"):
            data = data.astype(dtype)
        else:  # a bare CUDA tensor
            from ...device_array import DeviceArray

            data = DeviceArray(data).astype(dtype)
            data = data if type(data) is not DeviceArray else data.tensor
        return DataCube(data, self.dims, self.coords, self.attrs, self.name)

    def transpose(self, *dims):
        perm = [self.dims.index(d) for d in dims]
        data = self.data.permute(*perm) if hasattr(self.data, "permute") else self.data.transpose(*perm)
        return DataCube(data, dims, self.coords, self.attrs, self.name)
