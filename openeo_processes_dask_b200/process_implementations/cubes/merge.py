"""``merge_cubes`` with the reference's signature (reference ``cubes/merge.py:54-295``).

The reference stacks both cubes along a new ``__cubes__`` dimension (``xr.concat``), re-chunks, and reduces
that 2-long dimension with the overlap resolver -- a process of ``x`` and ``y`` (``add``, ``max`` of two,
``and`` / ``or`` on masks, ``eq`` ...), which the ``@process`` decorator feeds from ``named_parameters`` while
the stacked array it is handed positionally is dropped (``core.py:52-58``).  Here the stacked copy is never
built: the resolver is called with ``x`` / ``y`` views of the two cubes and runs as ONE ``eoc_binary`` launch;
a lower-dimensional cube (example 4) is broadcast by stride 0 inside the kernel instead of ``broadcast_like``.

Cases, as in the reference (spec examples 1-4, https://processes.openeo.org/#merge_cubes):

* same dimensions, same labels (``:87-121``): no resolver -> both cubes stacked along ``__cubes__``
  (labels ``cube1`` / ``cube2``); resolver -> ``resolver(x=cube1, y=cube2)`` per pixel.
* same dimensions, labels differ (``:122-239``): if no dimension has labels both in common AND on one side
  only, the cubes are joined by coordinates (``combine_by_coords``); this implementation covers the join
  along ONE dimension (disjoint bands or disjoint time steps -- the spec's example 1).  If exactly one
  dimension overlaps partially, the common labels are resolved with the resolver and joined with the rest of
  both cubes (example 2; ``OverlapResolverMissing`` without a resolver); more -> ``ValueError``.
* one or two dimensions exist in one cube only (``:241-291``): the lower-dimensional cube is broadcast over
  the other and every pixel goes through the resolver (``OverlapResolverMissing`` without one); ``x`` is
  always cube1.  More than two differing dimensions -> ``ValueError``.

Array bookkeeping (label lookups, gathers of sub-cubes, the final join) is host logic plus torch / numpy
indexing; the arithmetic is the resolver's kernel.
"""

from __future__ import annotations

from collections import namedtuple

import numpy as np

from ... import engine
from ...device_array import DeviceArray
from .cube import Coordinate, DataCube
from ..exceptions import OverlapResolverMissing

__all__ = ["merge_cubes"]

NEW_DIM_NAME = "__cubes__"
NEW_DIM_COORDS = ["cube1", "cube2"]
FLOAT_TOLERANCE = 1e-6  # reference merge.py:16

Overlap = namedtuple("Overlap", ["only_in_cube1", "only_in_cube2", "in_both"])


# ---- array plumbing (no arithmetic) ---------------------------------------------------------------------------
def _raw(data):
    return data.tensor if isinstance(data, DeviceArray) else data


def _like(ref, raw):
    return DeviceArray(raw) if isinstance(ref, DeviceArray) else raw


def _take(data, idx, axis):
    raw = _raw(data)
    if engine.is_torch(raw):
        import torch

        return _like(data, raw.index_select(axis, torch.as_tensor(np.asarray(idx), device=raw.device, dtype=torch.long)))
    return np.take(raw, np.asarray(idx), axis=axis)


def _concat(parts, axis, stack=False):
    raws = [_raw(p) for p in parts]
    if any(engine.is_torch(r) for r in raws):
        import torch

        dev = next(r for r in raws if engine.is_torch(r)).device
        raws = [r if engine.is_torch(r) else torch.as_tensor(r, device=dev) for r in raws]
        dt = raws[0].dtype
        for r in raws[1:]:
            dt = torch.promote_types(dt, r.dtype)
        raws = [r.to(dt) for r in raws]
        out = torch.stack(raws, dim=axis) if stack else torch.cat(raws, dim=axis)
    else:
        out = np.stack(raws, axis=axis) if stack else np.concatenate(raws, axis=axis)
    return _like(parts[0], out) if any(isinstance(p, DeviceArray) for p in parts) else out


def _labels(cube, dim):
    return np.asarray(cube[dim].values)


def _positions(labels, wanted):
    """index of every `wanted` label in `labels` (labels are unique along a dimension)."""
    order = {lab: i for i, lab in enumerate(labels.tolist())}
    return np.asarray([order[w] for w in np.asarray(wanted).tolist()], dtype=np.int64)


def _align_coordinates(cube1, cube2):
    """reference merge.py:24-51: float coordinates that differ by less than 1e-6 take cube1's values."""
    for dim in set(cube1.dims) & set(cube2.dims):
        c1, c2 = _labels(cube1, dim), _labels(cube2, dim)
        if not (np.issubdtype(c1.dtype, np.floating) and np.issubdtype(c2.dtype, np.floating)):
            continue
        if c1.shape != c2.shape:
            continue
        if c1.size and np.max(np.abs(c1 - c2)) < FLOAT_TOLERANCE:
            cube2 = cube2.assign_coords({dim: c1})
    return cube1, cube2


def _resolve(overlap_resolver, x, y, context, axis=0):
    """What ``stacked.reduce(overlap_resolver, dim="__cubes__", positional_parameters={}, named_parameters={x, y,
    context})`` amounts to (merge.py:106-117): the resolver's ParameterReferences x / y are filled from the named
    parameters; `axis` is dropped by @process unless the resolver's signature has it."""
    return overlap_resolver(positional_parameters={}, named_parameters={"x": x, "y": y, "context": context},
                            axis=axis, keepdims=False)


def _to_order(cube, dims):
    return cube if tuple(cube.dims) == tuple(dims) else cube.transpose(*dims)


def _with_data(template, data, dims=None, coords=None):
    return DataCube(data, template.dims if dims is None else dims, template.coords if coords is None else coords,
                    dict(template.attrs), template.name)


def _sub(cube, dim, labels):
    """cube.sel({dim: labels}) for a list of labels."""
    ax = cube.dims.index(dim)
    idx = _positions(_labels(cube, dim), labels)
    coords = {k: (Coordinate(k, c.dim, c.values[idx]) if c.dim == dim else c) for k, c in cube.coords.items()}
    if dim not in coords:
        coords[dim] = Coordinate(dim, dim, np.asarray(labels))
    return DataCube(_take(cube.data, idx, ax), cube.dims, coords, cube.attrs, cube.name)


def _join(parts, dim, order):
    """Concatenate cubes with disjoint labels along `dim` and arrange the labels as `order`."""
    first = parts[0]
    ax = first.dims.index(dim)
    parts = [_to_order(p, first.dims) for p in parts if p.shape[p.dims.index(dim)] > 0]
    data = _concat([p.data for p in parts], ax)
    labels = np.concatenate([_labels(p, dim) for p in parts])
    idx = _positions(labels, order)
    if not np.array_equal(idx, np.arange(len(idx))):
        data = _take(data, idx, ax)
    coords = {k: c for k, c in first.coords.items() if c.dim != dim}
    coords[dim] = Coordinate(dim, dim, np.asarray(order))
    return DataCube(data, first.dims, coords, dict(first.attrs), first.name)


# ---- the process ---------------------------------------------------------------------------------------------
def merge_cubes(cube1, cube2, overlap_resolver=None, context=None):
    if context is None:
        context = {}
    if not isinstance(cube1, type(cube2)):
        raise Exception(f"Provided cubes have incompatible types. cube1: {type(cube1)}, cube2: {type(cube2)}")
    cube1, cube2 = _align_coordinates(cube1, cube2)

    shared = [d for d in cube1.dims if d in cube2.dims]
    overlap = {}
    for dim in shared:
        l1, l2 = _labels(cube1, dim), _labels(cube2, dim)
        overlap[dim] = Overlap(np.setdiff1d(l1, l2), np.setdiff1d(l2, l1), np.intersect1d(l1, l2))
    differing_dims = set(cube1.dims) ^ set(cube2.dims)

    if len(differing_dims) == 0:
        cube2 = _to_order(cube2, cube1.dims)
        if all(len(o.only_in_cube1) == 0 and len(o.only_in_cube2) == 0 for o in overlap.values()):
            # example 3: same dimensions, same labels (xr.concat aligns cube2's label order to cube1's)
            for dim in cube1.dims:
                l1, l2 = _labels(cube1, dim), _labels(cube2, dim)
                if not np.array_equal(l1, l2):
                    cube2 = _sub(cube2, dim, l1)
            if overlap_resolver is None:
                data = _concat([cube1.data, cube2.data], 0, stack=True)
                coords = dict(cube1.coords)
                coords[NEW_DIM_NAME] = Coordinate(NEW_DIM_NAME, NEW_DIM_NAME, np.asarray(NEW_DIM_COORDS))
                return DataCube(data, (NEW_DIM_NAME,) + tuple(cube1.dims), coords, dict(cube1.attrs), cube1.name)
            return _with_data(cube1, _resolve(overlap_resolver, cube1.data, cube2.data, context))

        need_resolve = [d for d in shared if len(overlap[d].in_both) > 0
                        and (len(overlap[d].only_in_cube1) > 0 or len(overlap[d].only_in_cube2) > 0)]
        if len(need_resolve) == 0:
            # example 1: join by coordinates (combine_by_coords); one dimension with disjoint labels
            differ = [d for d in shared if len(overlap[d].only_in_cube1) > 0 or len(overlap[d].only_in_cube2) > 0]
            if len(differ) != 1:
                raise NotImplementedError(
                    "merge_cubes: joining cubes whose labels differ along more than one dimension "
                    f"({differ}) is not part of the B200 path")
            dim = differ[0]
            l1, l2 = _labels(cube1, dim), _labels(cube2, dim)
            if dim in cube1.openeo.band_dims:  # merge.py:137-160: cube1's bands, then cube2's new ones
                order = list(l1) + [b for b in l2 if b not in set(l1.tolist())]
            else:                              # combine_by_coords yields monotonically increasing labels
                order = np.sort(np.concatenate([l1, l2]))
            return _join([cube1, cube2], dim, order)
        if len(need_resolve) == 1:
            # example 2: resolve the common labels of that dimension, join with what each cube has alone
            if overlap_resolver is None or not callable(overlap_resolver):
                raise OverlapResolverMissing("Overlapping data cubes, but no overlap resolver has been specified.")
            dim = need_resolve[0]
            other = [d for d in shared if d != dim and
                     (len(overlap[d].only_in_cube1) > 0 or len(overlap[d].only_in_cube2) > 0)]
            if other:
                raise NotImplementedError(f"merge_cubes: labels also differ along {other}; not part of the B200 path")
            ov = overlap[dim]
            c1, c2 = _sub(cube1, dim, ov.in_both), _sub(cube2, dim, ov.in_both)
            merged = _with_data(c1, _resolve(overlap_resolver, c1.data, c2.data, context))
            parts = [merged, _sub(cube1, dim, ov.only_in_cube1), _sub(cube2, dim, ov.only_in_cube2)]
            order = np.sort(np.concatenate([ov.in_both, ov.only_in_cube1, ov.only_in_cube2]))
            return _join(parts, dim, order)
        raise ValueError("More than one overlapping dimension, merge not possible.")

    if len(differing_dims) <= 2:
        if overlap_resolver is None or not callable(overlap_resolver):
            raise OverlapResolverMissing("Overlapping data cubes, but no overlap resolver has been specified.")
        # example 4: broadcast the lower-dimensional cube over the other one
        cube1_is_lower = len(cube1.dims) < len(cube2.dims)
        lower, higher = (cube1, cube2) if cube1_is_lower else (cube2, cube1)
        if not set(lower.dims) <= set(higher.dims):
            raise NotImplementedError("merge_cubes: both cubes have dimensions the other lacks; not part of the B200 path")
        for dim in lower.dims:
            l_lo, l_hi = _labels(lower, dim), _labels(higher, dim)
            if not np.array_equal(l_lo, l_hi):
                if set(l_lo.tolist()) != set(l_hi.tolist()):
                    raise NotImplementedError(f"merge_cubes: labels of {dim!r} differ between the cubes")
                lower = _sub(lower, dim, l_hi)
        # a view of the lower cube in the higher cube's dimension order with 1-long axes where it has none:
        # eoc_binary broadcasts those with stride 0, nothing is expanded in memory
        lower = _to_order(lower, [d for d in higher.dims if d in lower.dims])
        raw = _raw(lower.data)
        shape = [higher.shape[i] if d in lower.dims else 1 for i, d in enumerate(higher.dims)]
        view = _like(lower.data, raw.reshape(shape))
        x, y = (view, higher.data) if cube1_is_lower else (higher.data, view)
        return _with_data(higher, _resolve(overlap_resolver, x, y, context))
    raise ValueError("Number of differing dimensions is >2, merge not possible.")
