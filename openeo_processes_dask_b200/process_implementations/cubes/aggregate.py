"""``aggregate_temporal`` / ``aggregate_temporal_period`` with the reference's signatures.

The reference (``cubes/aggregate.py:29-101``) turns the intervals into bin edges and lets xarray's
``groupby_bins(t, bins).reduce(reducer)`` loop over the groups in Python, gathering a copy of every
group.  Here the same bins become index groups of the t axis and -- when the reducer is one of this
package's streaming reducers -- ONE segmented CUDA reduction produces all bins
(``eoc_segmented_reduce``); any other reducer is applied per group.

Bin semantics follow the reference, not the openEO spec text: ``groupby_bins`` uses ``pandas.cut``
defaults, i.e. RIGHT-closed bins ``(start, end]``; timestamps are truncated to seconds; labels default
to the interval starts as strings; bins between non-adjacent intervals are computed and dropped; an
empty bin yields NaN; overlapping intervals raise ``NotImplementedError``.
"""

from __future__ import annotations

import functools
from typing import Callable, Optional

import numpy as np

from ... import _lib as L
from ... import engine
from ...device_array import DeviceArray
from .. import arrays as _arrays
from .. import math as _math
from ..exceptions import DimensionNotAvailable, TooManyDimensions
from ._xr_interop import dims_of
from .cube import DataCube

__all__ = ["aggregate_temporal", "aggregate_temporal_period", "aggregate_spatial"]


# ---- which reducers can be fused into the segmented kernel ---------------------------------------------
_FUSABLE = {
    _math.mean: L.OP_MEAN, _math._sum: L.OP_SUM, _math._min: L.OP_MIN, _math._max: L.OP_MAX,
    _math.sd: L.OP_SD, _math.variance: L.OP_VARIANCE, _math.product: L.OP_PRODUCT,
}


def _fusable(reducer):
    """(op, ignore_nodata) if ``reducer`` is a (partial of a @process-wrapped) streaming reducer."""
    kwargs = {}
    f = reducer
    while isinstance(f, functools.partial):
        kwargs = {**f.keywords, **kwargs}
        if f.args:
            return None
        f = f.func
    while hasattr(f, "__wrapped__"):
        f = f.__wrapped__
    extra = set(kwargs) - {"data", "ignore_nodata", "axis", "keepdims", "context", "dim_labels"}
    if extra:
        return None
    if f is _arrays.count:
        return None if (set(kwargs) - {"data", "context"}) else (L.OP_COUNT, True)
    if f in _FUSABLE:
        return _FUSABLE[f], bool(kwargs.get("ignore_nodata", True))
    return None


def _temporal_dim(data, dimension):
    if dimension is not None:
        if dimension not in data.dims:
            raise DimensionNotAvailable(f"A dimension with the specified name: {dimension} does not exist.")
        return dimension
    found = dims_of(data).temporal_dims
    if not found:
        raise DimensionNotAvailable(
            f"No temporal dimension detected on dataset. Available dimensions: {data.dims}")
    if len(found) > 1:
        raise TooManyDimensions(
            f"The data cube contains multiple temporal dimensions: {found}. The parameter `dimension` must be specified.")
    return found[0]


def _interval_strings(intervals):
    """pg-parser TemporalInterval(s) objects -> [[start, end], ...] strings (reference :56-64)."""
    out = []
    for iv in intervals:
        if hasattr(iv, "root") and not isinstance(iv, (list, tuple, str)):
            iv = iv.root
        pair = []
        for bound in iv:
            pair.append(str(bound.root) if hasattr(bound, "root") else bound)
        out.append(pair)
    return out


def plan_bins(t_values, intervals, labels=None):
    """Host logic of aggregate_temporal: index groups of the t axis for the selected labels.

    Returns (groups, labels) where groups[k] lists the positions along t (ascending) that fall in the
    bin named labels[k]; bins are right-closed.
    """
    if hasattr(intervals, "root"):
        intervals = intervals.root
    intervals = _interval_strings(intervals)
    iv = np.array(intervals, dtype=np.datetime64).astype("datetime64[s]").astype(float)
    if not labels:
        labels = np.array(intervals, dtype="datetime64[s]").astype(str)[:, 0]
    labels = [str(l) for l in labels]
    if (iv[1:, 0] < iv[:-1, 1]).any():
        raise NotImplementedError("Aggregating data for overlapping time ranges is not implemented. ")
    edges = np.unique(iv.reshape(-1))
    # name every bin between consecutive edges: interval k starts at edge e_k; a gap after interval k
    # (next start != this end) is an unnamed filler bin
    bin_of_label = {}
    for k, lab in enumerate(labels):
        bin_of_label.setdefault(lab, int(np.searchsorted(edges, iv[k, 0])))
    t = np.array(np.asarray(t_values).astype(str), dtype="datetime64[s]").astype(float)
    code = np.searchsorted(edges, t, side="left") - 1  # (e_j, e_j+1]  <=>  pandas.cut(right=True)
    code[(code < 0) | (code >= len(edges) - 1)] = -1
    groups = [np.nonzero(code == bin_of_label[lab])[0].tolist() for lab in labels]
    return groups, labels


def _apply_groups(data, t, groups, labels, reducer, reducer_kwargs):
    """Reduce each index group of dimension ``t`` and assemble the output cube."""
    axis = data.dims.index(t)
    raw = data.data
    if hasattr(raw, "materialize"):  # lazy ndvi / mask results: aggregate over the evaluated cube
        raw = raw.materialize()
        data = data.copy(data=raw)
    fused = _fusable(reducer)
    rewrap = (lambda x: x)
    if isinstance(raw, DeviceArray):
        raw, rewrap = raw.tensor, DeviceArray
    if fused is not None and not _math._is_dask(raw):
        op, ignore = fused
        res = engine.segmented_reduce(raw, axis, groups, [op], ignore_nodata=ignore, ddof=1)[op]
        in_dt = engine.np_dtype_of(raw)
        empty = [k for k, g in enumerate(groups) if len(g) == 0]
        if op == L.OP_COUNT and empty:
            res = engine._astype(res, np.dtype(np.float64))  # xarray re-indexes missing groups to NaN
            key = [slice(None)] * res.ndim
            key[axis] = empty
            res[tuple(key)] = float("nan")
        cast = _math._int_result_dtype(op, in_dt)
        if cast is not None and not empty:
            res = engine._astype(res, cast)
        out = rewrap(res)
    else:
        parts = []
        for g in groups:
            if len(g) == 0:
                parts.append(None)
                continue
            key = [slice(None)] * data.ndim
            key[axis] = g if not (g == list(range(g[0], g[-1] + 1))) else slice(g[0], g[-1] + 1)
            parts.append(reducer(data.data[tuple(key)], axis=axis, **reducer_kwargs))
        proto = next((p for p in parts if p is not None), None)
        if proto is None:
            shape = [s for i, s in enumerate(data.shape) if i != axis]
            proto = np.full(shape, np.nan)
        filled = []
        for p in parts:
            if p is None:
                p = _nan_like(proto)
            filled.append(p)
        out = _stack(filled, axis)
    coords = {k: c for k, c in data.coords.items() if getattr(c, "dim", k) != t} if isinstance(data, DataCube) else None
    return _rebuild(data, out, t, np.asarray(labels), coords)


def _nan_like(proto):
    if isinstance(proto, DeviceArray):
        return DeviceArray(_nan_like(proto.tensor))
    if engine.is_torch(proto):
        import torch

        return torch.full_like(proto if proto.is_floating_point() else proto.double(), float("nan"))
    return np.full(np.shape(proto), np.nan, dtype=np.result_type(np.asarray(proto).dtype, np.float32))


def _stack(parts, axis):
    if isinstance(parts[0], DeviceArray):
        return DeviceArray(_stack([p.tensor for p in parts], axis))
    if engine.is_torch(parts[0]):
        import torch

        dt = torch.result_type(parts[0], parts[-1])
        return torch.stack([p.to(dt) for p in parts], dim=axis)
    return np.stack([np.asarray(p) for p in parts], axis=axis)


def _rebuild(template, array, t, t_labels, coords):
    if isinstance(template, DataCube):
        coords = dict(coords)
        coords[t] = t_labels
        return DataCube(array, template.dims, coords, dict(template.attrs), template.name)
    import xarray as xr  # pragma: no cover - only reachable with a real DataArray

    new_coords = {k: v for k, v in template.coords.items() if t not in v.dims}
    new_coords[t] = t_labels
    return xr.DataArray(array, dims=template.dims, coords=new_coords, attrs=template.attrs, name=template.name)


def aggregate_temporal(data, intervals, reducer: Callable, labels: Optional[list] = None,
                       dimension: Optional[str] = None, context: Optional[dict] = None, **kwargs):
    """reference cubes/aggregate.py:29-101 (``context`` is accepted and, as there, not forwarded)."""
    t = _temporal_dim(data, dimension)
    groups, labels = plan_bins(data[t].values, intervals, labels)
    return _apply_groups(data, t, groups, labels, reducer, {"positional_parameters": {"data": 0}})


# ---- aggregate_temporal_period --------------------------------------------------------------------------
_PANDAS_FREQ = {"hour": "h", "day": "D", "week": "W", "month": "ME", "season": "QS-DEC", "year": "YS"}


def aggregate_temporal_period(data, reducer: Callable, period: str, dimension: Optional[str] = None):
    """reference cubes/aggregate.py:206-253: calendar bins instead of explicit intervals."""
    from ._periods import calendar_groups, period_intervals

    t = _temporal_dim(data, dimension)
    if period in _PANDAS_FREQ:
        groups, labels = calendar_groups(data[t].values, _PANDAS_FREQ[period])
        return _apply_groups(data, t, groups, labels, reducer, {"positional_parameters": {"data": 0}})
    intervals, labels = period_intervals(data[t].values, period)
    return aggregate_temporal(data=data, intervals=intervals, reducer=reducer, labels=labels, dimension=t)


# ---- aggregate_spatial ------------------------------------------------------------------------------------------
def _geometry_list(geometries):
    """GeoJSON in the shapes the reference accepts (cubes/aggregate.py:275-305) -> (list of geometry dicts, crs name
    or None).  Vector cubes / GeoDataFrames need geopandas, which is not part of this path."""
    if isinstance(geometries, (list, tuple)):
        return [g.get("geometry", g) if isinstance(g, dict) else g for g in geometries], None
    if not isinstance(geometries, dict):
        raise NotImplementedError("aggregate_spatial: geometries must be GeoJSON (dict / list of dicts) here; "
                                  "GeoDataFrame / vector-cube inputs need geopandas")
    crs = geometries.get("crs")
    if isinstance(crs, dict):
        crs = crs.get("properties", {}).get("name")
    kind = geometries.get("type")
    if kind == "FeatureCollection":
        return [f["geometry"] for f in geometries["features"]], crs
    if kind == "Feature":
        return [geometries["geometry"]], crs
    if kind in ("Polygon", "MultiPolygon"):
        return [geometries], crs
    if kind == "GeometryCollection":
        return list(geometries["geometries"]), crs
    raise NotImplementedError(f"aggregate_spatial: unsupported GeoJSON type {kind!r}")


def _rings(geometry):
    kind = geometry["type"]
    if kind == "Polygon":
        return [np.asarray(r, dtype=np.float64)[:, :2] for r in geometry["coordinates"]]
    if kind == "MultiPolygon":
        return [np.asarray(r, dtype=np.float64)[:, :2] for poly in geometry["coordinates"] for r in poly]
    raise NotImplementedError(f"aggregate_spatial: {kind} geometries (only polygons have an interior to aggregate)")


def _index_window(centres, lo, hi):
    """[i0, i1) of the pixel centres inside [lo, hi] for an ascending or descending coordinate."""
    c = np.asarray(centres, dtype=np.float64)
    hit = np.nonzero((c >= lo) & (c <= hi))[0]
    return (0, 0) if len(hit) == 0 else (int(hit[0]), int(hit[-1]) + 1)


def aggregate_spatial(data, geometries, reducer: Callable, chunk_size: int = 2):
    """reference cubes/aggregate.py:256-323: zonal statistics -- ``xvec.zonal_stats(geometries, method="iterate",
    stats=reducer)``: per geometry the cube is masked to the pixels whose CENTRE lies inside it (rasterio
    ``geometry_mask(all_touched=False)``) and reduced over (y, x).  Here the mask of a geometry's bounding window is
    rasterised on the device (``eoc_rasterize_polygon``, even-odd rule) and a streaming reducer of this package runs
    as ONE masked window reduction over all (t, bands) planes (``eoc_zonal_reduce``); any other reducer gets the
    where()-masked window.  Result: a cube with a leading ``geometry`` dimension and the non-spatial dimensions of
    ``data``.  Not done here: reprojection of the geometries (needs pyproj): their CRS must be the cube's."""
    d = dims_of(data)
    x_dim, y_dim = d.x_dim, d.y_dim
    if x_dim is None or y_dim is None:
        raise DimensionNotAvailable(f"aggregate_spatial needs spatial dimensions; found {data.dims}")
    geoms, crs = _geometry_list(geometries)
    cube_crs = data.attrs.get("crs") if hasattr(data, "attrs") else None
    if crs is not None and cube_crs is not None and str(crs).upper().replace("URN:OGC:DEF:CRS:", "") != str(cube_crs).upper():
        raise NotImplementedError(f"aggregate_spatial: geometries in {crs} on a cube in {cube_crs}: reprojection is "
                                  "out of scope here, pass geometries in the cube's CRS")
    other = tuple(n for n in data.dims if n not in (x_dim, y_dim))
    cube = data if tuple(data.dims) == other + (y_dim, x_dim) else data.transpose(*(other + (y_dim, x_dim)))
    raw = cube.data
    if hasattr(raw, "materialize"):
        raw = raw.materialize()
    rewrap = (lambda t: t)
    was_numpy = False
    if isinstance(raw, DeviceArray):
        raw, rewrap = raw.tensor, DeviceArray
    elif _math._is_dask(raw):
        raw = raw.compute()
    if not engine.is_torch(raw):
        was_numpy = True
        raw = engine.torch.from_numpy(np.ascontiguousarray(engine._native(np.asarray(raw)))).cuda()
    xs, ys = np.asarray(data[x_dim].values, dtype=np.float64), np.asarray(data[y_dim].values, dtype=np.float64)
    fused = _fusable(reducer)
    results = []
    for g in geoms:
        rings = _rings(g)
        allv = np.concatenate(rings, axis=0)
        x0, x1 = _index_window(xs, allv[:, 0].min(), allv[:, 0].max())
        y0, y1 = _index_window(ys, allv[:, 1].min(), allv[:, 1].max())
        mask = engine.rasterize_polygon(xs[x0:x1], ys[y0:y1], rings, device=raw.device)
        if fused is not None:
            op, ignore = fused
            res = engine.zonal_reduce(raw, mask, (y0, y1, x0, x1), [op], ignore_nodata=ignore, ddof=1)[op]
            cast = _math._int_result_dtype(op, engine.np_dtype_of(raw))
            if cast is not None:
                res = engine._astype(res, cast)
        else:
            win = raw[..., y0:y1, x0:x1]
            if win.shape[-1] * win.shape[-2] == 0:
                win = engine.torch.full(tuple(raw.shape[:-2]) + (1, 1), float("nan"), device=raw.device)
                mask = engine.torch.zeros((1, 1), dtype=engine.torch.uint8, device=raw.device)
            masked = engine.torch.where(mask.bool(), win.to(engine.torch.float64 if not win.is_floating_point() else win.dtype),
                                        float("nan"))
            res = reducer(masked.contiguous(), axis=(masked.ndim - 2, masked.ndim - 1), positional_parameters={"data": 0})
            res = res.tensor if isinstance(res, DeviceArray) else res
        results.append(res)
    out = engine.torch.stack([r if engine.is_torch(r) else engine.torch.as_tensor(np.asarray(r), device=raw.device)
                              for r in results], dim=0)
    out = out.cpu().numpy() if was_numpy else rewrap(out)
    coords = {k: c for k, c in data.coords.items() if getattr(c, "dim", k) in other}
    coords["geometry"] = np.arange(len(geoms))
    return DataCube(out, ("geometry",) + other, coords, dict(data.attrs), data.name)
