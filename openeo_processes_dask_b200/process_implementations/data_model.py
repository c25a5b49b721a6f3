"""Type aliases at the boundary (reference data_model.py:9): a raster cube is an xarray.DataArray
when xarray is installed, else the built-in stand-in with the same duck type."""

try:  # pragma: no cover - xarray is not installable in the build environment
    import xarray as _xr

    RasterCube = _xr.DataArray
except Exception:
    from .cubes.cube import DataCube as RasterCube  # noqa: F401
