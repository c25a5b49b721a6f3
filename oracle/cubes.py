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


def merge_cubes(c1, c2, resolver=None, band_dims=("bands",)):
    """cubes/merge.py:54-295 on ``(array, dims, coords)`` cubes.  ``resolver(x, y)`` is the overlap resolver as a
    plain function of two arrays.  The xarray behaviours relied on, restated [U]:

    * ``xr.concat([a, b], "__cubes__")`` of equally labelled cubes stacks them on a new leading axis, b aligned to
      a's dimension and label order; reducing that axis with the resolver is ``resolver(x=a, y=b)``   :87-121
    * ``combine_by_coords`` of cubes with disjoint labels along one dimension concatenates them and sorts the
      labels ascending; for a band dimension the reference restores "cube1's bands, then cube2's new ones"  :137-160
    * partial overlap along one dimension: common labels resolved, joined with the rest of both cubes, labels
      ascending                                                                                               :162-235
    * ``broadcast_like``: the lower-dimensional cube is repeated along the dimensions it lacks            :241-291
    Returns (array, dims, coords)."""
    a1, d1, k1 = c1
    a2, d2, k2 = c2
    differing = set(d1) ^ set(d2)
    if not differing:
        a2 = np.transpose(a2, [d2.index(d) for d in d1])
        k2 = dict(k2)
        same = all(set(np.asarray(k1[d]).tolist()) == set(np.asarray(k2[d]).tolist()) for d in d1)
        if same:
            for ax, d in enumerate(d1):
                order = {lab: i for i, lab in enumerate(np.asarray(k2[d]).tolist())}
                a2 = np.take(a2, [order[lab] for lab in np.asarray(k1[d]).tolist()], axis=ax)
            if resolver is None:
                coords = dict(k1)
                coords["__cubes__"] = np.asarray(["cube1", "cube2"])
                return np.stack([a1, a2]), ("__cubes__",) + tuple(d1), coords
            return resolver(a1, a2), tuple(d1), dict(k1)
        differ = [d for d in d1 if set(np.asarray(k1[d]).tolist()) != set(np.asarray(k2[d]).tolist())]
        partial = [d for d in differ if len(np.intersect1d(k1[d], k2[d])) > 0]
        if len(partial) > 1:
            raise ValueError("More than one overlapping dimension, merge not possible.")
        assert len(differ) == 1, "oracle covers one differing dimension"
        d = differ[0]
        ax = d1.index(d)
        l1, l2 = np.asarray(k1[d]), np.asarray(k2[d])
        both = np.intersect1d(l1, l2)
        if len(both) and resolver is None:
            raise KeyError("OverlapResolverMissing")
        pieces, labels = [], []
        if len(both):
            x = np.take(a1, [l1.tolist().index(v) for v in both.tolist()], axis=ax)
            y = np.take(a2, [l2.tolist().index(v) for v in both.tolist()], axis=ax)
            pieces.append(resolver(x, y))
            labels += both.tolist()
        for arr, lab, other in ((a1, l1, l2), (a2, l2, l1)):
            only = [v for v in lab.tolist() if v not in set(other.tolist())]
            pieces.append(np.take(arr, [lab.tolist().index(v) for v in only], axis=ax))
            labels += only
        out = np.concatenate([p.astype(np.result_type(*pieces)) for p in pieces], axis=ax)
        if d in band_dims and not len(both):
            order = l1.tolist() + [v for v in l2.tolist() if v not in set(l1.tolist())]
        else:
            order = sorted(labels)
        out = np.take(out, [labels.index(v) for v in order], axis=ax)
        coords = dict(k1)
        coords[d] = np.asarray(order)
        return out, tuple(d1), coords
    if len(differing) <= 2:
        if resolver is None:
            raise KeyError("OverlapResolverMissing")
        lower_is_1 = len(d1) < len(d2)
        (al, dl, kl), (ah, dh, kh) = ((c1, c2) if lower_is_1 else (c2, c1))
        al = np.transpose(al, [dl.index(d) for d in dh if d in dl])
        al = al.reshape([ah.shape[i] if d in dl else 1 for i, d in enumerate(dh)])
        bl = np.broadcast_to(al, ah.shape)
        out = resolver(bl, ah) if lower_is_1 else resolver(ah, bl)
        return out, tuple(dh), dict(kh)
    raise ValueError("Number of differing dimensions is >2, merge not possible.")


# ---- aggregate_temporal_period (cubes/aggregate.py:104-253) --------------------------------------------------
_PANDAS_FREQ = {"hour": "h", "day": "D", "week": "W", "month": "ME", "season": "QS-DEC", "year": "YS"}


def get_intervals(t_values, period):
    """cubes/aggregate.py:104-203 restated with the same pandas calls (decade / decade-ad / tropical-season /
    dekad bounds and labels).  PINNED: equal to the outputs of the reference's own function on 569 extents
    (tests/golden/reference_cubes.json, tests/test_cubes_golden.py); pandas-version dependent like the original.
    Returns (interval_matrix of strings [n, 2], labels)."""
    fmt = "%Y-%m-%dT%H:%M:%S"
    start, end = pd.to_datetime(t_values[0]), pd.to_datetime(t_values[-1])
    ys, ye, ms, me = start.year, end.year, start.month, end.month
    if period in ("decade", "decade-ad"):
        shift = 1 if period == "decade-ad" else 0
        a = np.datetime64(str(int(np.floor(ys / 10) * 10 + shift)))
        b = np.datetime64(str(int(np.ceil(ye / 10) * 10 + shift)))
        rng = pd.date_range(start=a, end=b, freq="10YS")
        bounds, labels = rng.strftime(fmt), list(rng.strftime("%Y")[:-1])
    elif period == "tropical-season":
        if 5 <= ms < 10:
            a = np.datetime64(f"{ys}-05-01")
        elif ms < 5:
            a = np.datetime64(f"{ys - 1}-11-01")
        else:
            a = np.datetime64(f"{ys}-11-01")
        if 5 <= me < 10:
            b = np.datetime64(f"{ye}-11-01")
        elif me < 5:
            b = np.datetime64(f"{ye}-05-01")
        else:
            b = np.datetime64(f"{ye + 1}-05-01")
        bounds = pd.period_range(start=a, end=b, freq="6M").strftime(fmt)
        labels = []
        for s in bounds[:-1]:
            if "-11-" in s:
                labels.append(s[:5] + "ndjfma")
            if "-05-" in s:
                labels.append(s[:5] + "mjjaso")
    elif period == "dekad":
        d0 = int(np.floor(start.day / 10) * 10 + 1)
        bounds = pd.date_range(start=f"{ys}-{ms}-{d0}", end=f"{ys}-{ms}-22", freq="10D").strftime(fmt)
        for m0 in pd.date_range(start=f"{ys}-{ms}-22", end=t_values[-1], freq="1MS")[:-1]:
            bounds = bounds.append(pd.date_range(start=m0, freq="10D", periods=3).strftime(fmt))
        n = int(np.ceil((end.day - 1) / 10))
        first = f"{ye}-{me}-01"
        if end.day > 21:
            tail = pd.date_range(start=first, freq="10D", periods=n)
            k = 7
            nxt = tail[-1] + pd.DateOffset(days=k)
            while nxt.day != 1:
                k += 1
                nxt = tail[-1] + pd.DateOffset(days=k)
            tail = tail.append(pd.DatetimeIndex(data=[str(nxt)])).strftime(fmt)
        else:
            tail = pd.date_range(start=first, freq="10D", periods=n + 1).strftime(fmt)
        bounds = bounds.append(tail)
        labels = []
        for s in bounds[:-1]:
            ts = pd.Timestamp(s)
            k = int(ts.day_of_year / 10)
            labels.append(f"{ts.year}-{k}" if k > 9 else f"{ts.year}-0{k}")
    else:
        raise NotImplementedError(period)
    b = np.array(bounds, dtype=str)
    matrix = np.zeros((len(b) - 1, 2)).astype(str)
    matrix[:, 0] = b[:-1]
    matrix[:, 1] = b[1:]
    return matrix, list(labels)


def resample_groups(t_values, freq):
    """``DataArray.resample({t: freq})`` groups [U]: xarray builds a pandas Grouper with pandas' default
    ``closed`` / ``label`` for the frequency; every bin between the first and the last is kept, empty ones too.
    Returns (list of member index arrays, bin labels datetime64[ns])."""
    idx = pd.DatetimeIndex(np.asarray(t_values).astype("datetime64[ns]"))
    series = pd.Series(np.arange(len(idx)), index=idx)
    groups, labels = [], []
    for label, members in series.resample(freq):
        groups.append(np.asarray(members.values, dtype=np.int64))
        labels.append(label.to_datetime64())
    return groups, np.asarray(labels, dtype="datetime64[ns]")


def aggregate_temporal_period(array, dims, coords, reducer, period, dimension="t", **reducer_kwargs):
    """cubes/aggregate.py:206-253 on plain arrays: pandas-frequency periods through ``resample_groups`` (an empty
    bin gives NaN [U]), the synthetic ones through ``get_intervals`` + ``aggregate_temporal``."""
    if period not in _PANDAS_FREQ:
        intervals, labels = get_intervals(np.asarray(coords[dimension]), period)
        return aggregate_temporal(array, dims, coords, intervals, reducer, labels=labels, dimension=dimension,
                                  **reducer_kwargs)
    axis = dims.index(dimension)
    groups, labels = resample_groups(coords[dimension], _PANDAS_FREQ[period])
    moved = np.moveaxis(np.asarray(array), axis, 0)
    parts = [None if len(g) == 0 else np.asarray(reducer(moved[g], axis=0, **reducer_kwargs)) for g in groups]
    proto = next(p for p in parts if p is not None)
    if any(p is None for p in parts):
        dt = np.result_type(proto.dtype, np.float32)
        parts = [np.full(proto.shape, np.nan, dtype=dt) if p is None else p.astype(dt) for p in parts]
    out = np.moveaxis(np.stack(parts, axis=0), 0, axis)
    out_coords = dict(coords)
    out_coords[dimension] = labels
    return out, dims, out_coords


# ---- aggregate_spatial (cubes/aggregate.py:256-323) -----------------------------------------------------------
def point_in_polygon(xs, ys, rings):
    """Boolean [len(ys), len(xs)]: pixel centre inside the polygon by the even-odd crossing rule -- the membership
    rasterio.features.geometry_mask(all_touched=False) computes for xvec.zonal_stats (pixel CENTRE inside).
    PARITY UNPINNED: rasterio / GDAL are not installable here; centres exactly on an edge may be decided differently."""
    px, py = np.meshgrid(np.asarray(xs, dtype=np.float64), np.asarray(ys, dtype=np.float64))
    inside = np.zeros(px.shape, dtype=bool)
    for ring in rings:
        ring = np.asarray(ring, dtype=np.float64).reshape(-1, 2)
        for k in range(len(ring)):
            x1, y1 = ring[k]
            x2, y2 = ring[(k + 1) % len(ring)]
            cross = (y1 > py) != (y2 > py)
            with np.errstate(divide="ignore", invalid="ignore"):
                xi = (x2 - x1) * (py - y1) / (y2 - y1) + x1
            inside ^= cross & (px < xi)
    return inside


def geometry_rings(geometry):
    """GeoJSON Polygon / MultiPolygon -> list of rings ((n, 2) arrays); even-odd treats holes and parts alike."""
    if geometry["type"] == "Polygon":
        return [np.asarray(r, dtype=np.float64)[:, :2] for r in geometry["coordinates"]]
    if geometry["type"] == "MultiPolygon":
        return [np.asarray(r, dtype=np.float64)[:, :2] for poly in geometry["coordinates"] for r in poly]
    raise NotImplementedError(geometry["type"])


def aggregate_spatial(array, dims, coords, geometries, reducer, x_dim="x", y_dim="y", **reducer_kwargs):
    """cubes/aggregate.py:256-323 via xvec.zonal_stats(method="iterate") restated [U]: per geometry
    ``data.where(mask).reduce(reducer, dim=(y, x))``, results stacked along a leading ``geometry`` dimension.
    ``geometries``: list of GeoJSON geometry dicts.  Returns (array, dims)."""
    array = np.asarray(array)
    iy, ix = dims.index(y_dim), dims.index(x_dim)
    out = []
    for g in geometries:
        m = point_in_polygon(coords[x_dim], coords[y_dim], geometry_rings(g))
        shape = [1] * array.ndim
        shape[iy], shape[ix] = m.shape
        mm = m.reshape(shape) if iy < ix else m.T.reshape(shape)
        masked = np.where(mm, array.astype(np.result_type(array.dtype, np.float32)), np.nan)
        out.append(np.asarray(reducer(masked, axis=(min(iy, ix), max(iy, ix)), **reducer_kwargs)))
    return np.stack(out, axis=0), ("geometry",) + tuple(d for d in dims if d not in (x_dim, y_dim))
