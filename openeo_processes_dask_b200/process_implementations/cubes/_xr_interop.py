"""Dimension typing by name: the ``.openeo`` accessor the cube processes consult.

Same vocabulary and precedence as the reference's accessor (``cubes/_xr_interop.py:7-21``): a
dimension is temporal / x / y / bands if its case-folded name is one of the guesses below, matches
are reported in guess order, everything else is "other".  ``OpenEODims`` works on a plain tuple of
dimension names, so it serves both the built-in DataCube and (registered below when xarray is
importable and the reference has not registered its own) ``xarray.DataArray.openeo``.
"""

from __future__ import annotations

GUESSES = {
    "temporal": ("DATE", "time", "t", "year", "quarter", "month", "week", "day", "hour", "second"),
    "x": ("x", "lon", "longitude"),
    "y": ("y", "lat", "latitude"),
    "bands": ("b", "bands", "band"),
}


class OpenEODims:
    def __init__(self, dims):
        self._dims = tuple(dims)
        folded = [str(d).casefold() for d in self._dims]
        self._found = {
            kind: [self._dims[folded.index(g)] for g in guesses if g in folded]
            for kind, guesses in GUESSES.items()
        }
        self._extra = {"spatial": [], "temporal": [], "bands": [], "other": []}

    def _existing(self, names):
        return tuple(d for d in names if d in self._dims)

    @property
    def spatial_dims(self):
        return self._existing(self._found["x"] + self._found["y"] + self._extra["spatial"])

    @property
    def temporal_dims(self):
        return self._existing(self._found["temporal"] + self._extra["temporal"])

    @property
    def band_dims(self):
        return self._existing(self._found["bands"] + self._extra["bands"])

    @property
    def other_dims(self):
        typed = set(self.spatial_dims) | set(self.temporal_dims) | set(self.band_dims)
        return tuple(d for d in self._dims if d not in typed)

    @property
    def x_dim(self):
        return next(iter(self._existing(self._found["x"])), None)

    @property
    def y_dim(self):
        return next(iter(self._existing(self._found["y"])), None)

    def add_dim_type(self, name, type):
        if name not in self._dims:
            raise ValueError("Trying to add a dimension that doesn't exist")
        if type not in self._extra:
            raise ValueError(f"Type {type} is not understood")
        self._extra[type].append(name)


def dims_of(cube) -> OpenEODims:
    """The accessor for any cube-like object (uses ``cube.openeo`` when it exists)."""
    acc = getattr(cube, "openeo", None)
    return acc if acc is not None else OpenEODims(cube.dims)


try:  # pragma: no cover - xarray is absent in the build environment
    import xarray as _xr

    if not hasattr(_xr.DataArray, "openeo"):

        @_xr.register_dataarray_accessor("openeo")
        class _OpenEOAccessor(OpenEODims):
            def __init__(self, obj):
                super().__init__(obj.dims)
except Exception:
    pass
