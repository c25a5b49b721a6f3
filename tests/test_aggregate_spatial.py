"""aggregate_spatial (reference cubes/aggregate.py:256-323 -> xvec.zonal_stats(method="iterate")).

PARITY UNPINNED at the geometry boundary: rasterio / xvec / geopandas are not installable here, so the membership
rule (pixel CENTRE inside, even-odd) and the reduce-over-where()-masked-raster semantics are restated in
oracle/cubes.py; the reference's own test (tests/test_aggregate.py:186-240) checks a polygon covering the cube's
extent against ``mean(["x","y"])`` -- mirrored below.
"""

import functools
import warnings

import numpy as np
import pytest

from oracle import cubes as OC
from oracle import reducers as R

warnings.simplefilter("ignore", RuntimeWarning)

XS = 600000.0 + np.arange(96) * 10.0 + 5.0          # pixel centres, ascending x
YS = 5100000.0 - np.arange(70) * 10.0 - 5.0         # descending y, like a north-up raster
POLY = {"type": "Polygon", "coordinates": [
    [[600103.0, 5099850.0], [600903.0, 5099790.0], [600750.0, 5099400.0], [600200.0, 5099420.0], [600103.0, 5099850.0]],
    [[600400.0, 5099700.0], [600600.0, 5099700.0], [600600.0, 5099550.0], [600400.0, 5099550.0], [600400.0, 5099700.0]]]}
MULTI = {"type": "MultiPolygon", "coordinates": [
    [[[600010.0, 5099990.0], [600110.0, 5099990.0], [600110.0, 5099900.0], [600010.0, 5099900.0]]],
    [[[600800.0, 5099500.0], [600950.0, 5099450.0], [600870.0, 5099330.0]]]]}
OUTSIDE = {"type": "Polygon", "coordinates": [[[0.0, 0.0], [10.0, 0.0], [10.0, 10.0], [0.0, 10.0]]]}
WHOLE = {"type": "Polygon", "coordinates": [[[599990.0, 5100010.0], [600970.0, 5100010.0], [600970.0, 5099290.0],
                                             [599990.0, 5099290.0]]]}
FC = {"type": "FeatureCollection", "crs": {"type": "name", "properties": {"name": "EPSG:32633"}},
      "features": [{"type": "Feature", "properties": None, "geometry": g} for g in (POLY, MULTI, OUTSIDE, WHOLE)]}


def test_oracle_point_in_polygon_basics():
    m = OC.point_in_polygon(XS, YS, OC.geometry_rings(POLY))
    assert m.shape == (70, 96) and 1500 < m.sum() < 3500
    assert not m[35, 50] and m[35, 30]        # inside the hole / inside the ring around it
    hole = OC.point_in_polygon(XS, YS, [np.asarray(POLY["coordinates"][1])])
    assert hole.sum() == 20 * 15 and not (m & hole).any()
    assert OC.point_in_polygon(XS, YS, OC.geometry_rings(WHOLE)).all()
    assert not OC.point_in_polygon(XS, YS, OC.geometry_rings(OUTSIDE)).any()
    mm = OC.point_in_polygon(XS, YS, OC.geometry_rings(MULTI))
    assert mm[:10, :11].sum() == 90 and mm.sum() > 90


def _cube(P, wrap, dtype=np.float32):
    rng = np.random.default_rng(17)
    x = rng.normal(3000, 1000, size=(5, 3, 70, 96)).astype(dtype)
    if np.dtype(dtype).kind == "f":
        x[rng.random(x.shape) < 0.2] = np.nan
    cube = P.DataCube(wrap(x), ("t", "bands", "y", "x"), {"x": XS, "y": YS, "bands": ["B02", "B03", "B04"]},
                      attrs={"crs": "EPSG:32633"})
    return x, cube


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["numpy", "cuda", "device_array"])
def test_gpu_aggregate_spatial_vs_oracle(kind):
    import torch

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200.device_array import DeviceArray

    wrap = {"numpy": lambda a: a, "cuda": lambda a: torch.from_numpy(a).cuda(), "device_array": DeviceArray.from_numpy}[kind]
    x, cube = _cube(P, wrap)
    ref = P.ParameterReference(from_parameter="data")
    geoms = [POLY, MULTI, OUTSIDE, WHOLE]
    for name, ofn in (("mean", R.mean), ("sum", R._sum), ("min", R._min), ("max", R._max), ("sd", R.sd),
                      ("count", R.count), ("median", R.median)):
        fn = {"sum": P._sum, "min": P._min, "max": P._max}.get(name) or getattr(P, name)
        out = P.aggregate_spatial(data=cube, geometries=FC, reducer=functools.partial(P.process(fn), data=ref))
        want, wdims = OC.aggregate_spatial(x, cube.dims, {"x": XS, "y": YS}, geoms, ofn)
        assert out.dims == wdims == ("geometry", "t", "bands") and out.shape == (4, 5, 3)
        got = out.values
        if name in ("min", "max", "count", "median"):
            np.testing.assert_array_equal(got, want, err_msg=name)
        else:
            np.testing.assert_allclose(got, want, rtol=1e-6, equal_nan=True, err_msg=name)
        assert np.isnan(got[2]).all() or name in ("sum", "count")        # nothing under the outside polygon
    # the reference's own check (tests/test_aggregate.py:226-240): a polygon over the whole extent == mean over x, y
    out = P.aggregate_spatial(data=cube, geometries=WHOLE, reducer=functools.partial(P.process(P.mean), data=ref))
    np.testing.assert_allclose(out.values[0], np.nanmean(x.reshape(5, 3, -1), axis=2), rtol=1e-6)
    # ignore_nodata = False: the reference reduces the where()-masked RASTER, so anything outside the polygon poisons
    red = functools.partial(P.process(P.mean), data=ref, ignore_nodata=False)
    assert np.isnan(P.aggregate_spatial(data=cube, geometries=POLY, reducer=red).values).all()
    with pytest.raises(NotImplementedError):
        P.aggregate_spatial(data=cube, geometries={**FC, "crs": {"type": "name", "properties": {"name": "EPSG:4326"}}},
                            reducer=red)


@pytest.mark.gpu
def test_gpu_rasterize_equals_oracle_on_every_pixel():
    """Edge cases decided identically: vertices on pixel centres, horizontal / vertical edges through centres."""
    from openeo_processes_dask_b200 import engine

    xs, ys = np.arange(40) + 0.5, 30.0 - np.arange(30) - 0.5
    shapes = [
        [np.array([[3.5, 25.5], [30.5, 25.5], [30.5, 5.5], [3.5, 5.5]])],                      # edges through centres
        [np.array([[1.0, 29.0], [39.0, 20.0], [20.5, 0.5], [5.5, 10.5]]), np.array([[10.0, 20.0], [20.0, 20.0], [15.0, 12.0]])],
        [np.array([[0.5, 0.5], [39.5, 29.5], [39.5, 0.5]])],                                   # diagonal through centres
        [np.array([[12.0, 12.0], [12.0, 12.0], [12.0, 12.0]])],                                # degenerate
    ]
    for rings in shapes:
        got = engine.rasterize_polygon(xs, ys, rings).cpu().numpy().astype(bool)
        np.testing.assert_array_equal(got, OC.point_in_polygon(xs, ys, rings))
