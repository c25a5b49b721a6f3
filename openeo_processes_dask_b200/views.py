"""Host logic: turn "reduce axis k of an N-d array" into the strided 3-D view the kernels take.

A reduction over ``axis`` of an array with ``shape``/``strides`` (in elements) is expressed as
``[outer, R, inner]`` with element (o, r, i) at ``o*stride_outer + r*stride_R + i`` (include/eoc.h).
That needs (a) the dims after the reduced ones to be jointly C-contiguous (-> inner), (b) the reduced
dims to collapse into one stride (-> R) and (c) the dims before them to collapse into one stride
(-> outer).  Band selections such as ``cube[:, 3]`` of a (t,b,y,x) cube satisfy this without a
copy (stride_R = B*Y*X); anything else is made contiguous by the caller first.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence, Tuple


@dataclass(frozen=True)
class ReduceView:
    outer: int
    R: int
    inner: int
    stride_outer: int
    stride_R: int
    out_shape: Tuple[int, ...]      # shape of the reduced result (keepdims=False)
    keep_shape: Tuple[int, ...]     # shape with the reduced axes kept as length 1
    axis_pos: int                   # where a new axis (quantiles) is inserted in out_shape


def normalize_axes(axis, ndim: int) -> Tuple[int, ...]:
    if axis is None:
        return tuple(range(ndim))
    if isinstance(axis, (tuple, list)):
        axes = tuple(int(a) % ndim if -ndim <= int(a) < ndim else _bad_axis(a, ndim) for a in axis)
    else:
        a = int(axis)
        axes = (a % ndim if -ndim <= a < ndim else _bad_axis(a, ndim),)
    if len(set(axes)) != len(axes):
        raise ValueError("duplicate value in 'axis'")
    return tuple(sorted(axes))


def _bad_axis(a, ndim):
    raise ValueError(f"axis {a} is out of bounds for array of dimension {ndim}")


def _collapse(shape: Sequence[int], strides: Sequence[int]) -> Optional[Tuple[int, int]]:
    """Collapse dims into (count, stride) if they are laid out like one strided dim; None if not."""
    dims = [(n, s) for n, s in zip(shape, strides) if n != 1]
    if not dims:
        return 1, 0
    count, stride = dims[-1]
    for n, s in reversed(dims[:-1]):
        if s != stride * count:
            return None
        count *= n
    return count, stride


def plan_view(shape: Sequence[int], strides: Sequence[int], axis) -> Optional[ReduceView]:
    """Return the 3-D view for reducing ``axis`` (int, tuple or None), or None if a copy is needed.

    ``strides`` are in elements.  Reduced axes must be adjacent (after sorting).
    """
    ndim = len(shape)
    axes = normalize_axes(axis, ndim)
    if any(b - a != 1 for a, b in zip(axes, axes[1:])):
        return None
    first, last = axes[0], axes[-1]
    inner_c = _collapse(shape[last + 1:], strides[last + 1:])
    red_c = _collapse(shape[first:last + 1], strides[first:last + 1])
    outer_c = _collapse(shape[:first], strides[:first])
    if inner_c is None or red_c is None or outer_c is None:
        return None
    inner, inner_stride = inner_c
    if inner != 1 and inner_stride != 1:
        return None
    R, stride_R = red_c
    outer, stride_outer = outer_c
    if any(n == 0 for n in shape):
        inner = 1
        for n in shape[last + 1:]:
            inner *= n
        R = 1
        for n in shape[first:last + 1]:
            R *= n
        outer = 1
        for n in shape[:first]:
            outer *= n
    out_shape = tuple(shape[:first]) + tuple(shape[last + 1:])
    keep_shape = tuple(shape[:first]) + (1,) * len(axes) + tuple(shape[last + 1:])
    return ReduceView(int(outer), int(R), int(inner), int(stride_outer), int(stride_R), out_shape,
                      keep_shape, first)


def plan_view_pair(shape: Sequence[int], strides_a: Sequence[int], strides_b: Sequence[int], axis):
    """Views of TWO arrays of the same shape (the second typically a broadcast: some strides 0) over the
    same [outer, R, inner] index space, or None.

    Unlike ``plan_view`` the kept dims AFTER the reduced ones may be split: a leading part joins ``outer``
    (element offset o*stride_outer may well be smaller than r*stride_R -- the kernels do not care), the
    rest is ``inner``.  That is what lets a (t,y,x) mask address a (t,bands,y,x) cube reduced over t:
    outer = bands (mask stride 0), R = t, inner = y*x.  The split with the largest ``inner`` wins.
    Returns (view_a, view_b, n_outer_tail) where n_outer_tail kept dims after the reduced ones went to
    ``outer``; dense outputs [outer, inner] still reshape to the kept dims in their original order.
    """
    ndim = len(shape)
    axes = normalize_axes(axis, ndim)
    if any(b - a != 1 for a, b in zip(axes, axes[1:])) or any(n == 0 for n in shape):
        return None
    first, last = axes[0], axes[-1]
    out_shape = tuple(shape[:first]) + tuple(shape[last + 1:])
    keep_shape = tuple(shape[:first]) + (1,) * len(axes) + tuple(shape[last + 1:])
    for split in range(last + 1, ndim + 1):
        views = []
        for strides in (strides_a, strides_b):
            inner_c = _collapse(shape[split:], strides[split:])
            red_c = _collapse(shape[first:last + 1], strides[first:last + 1])
            outer_c = _collapse(tuple(shape[:first]) + tuple(shape[last + 1:split]),
                                tuple(strides[:first]) + tuple(strides[last + 1:split]))
            if inner_c is None or red_c is None or outer_c is None:
                break
            inner, inner_stride = inner_c
            if inner != 1 and inner_stride != 1:
                break
            views.append(ReduceView(int(outer_c[0]), int(red_c[0]), int(inner), int(outer_c[1]), int(red_c[1]),
                                    out_shape, keep_shape, first))
        if len(views) == 2:
            return views[0], views[1], split - (last + 1)
    return None


def contiguous_strides(shape: Sequence[int]) -> Tuple[int, ...]:
    strides = [1] * len(shape)
    for i in range(len(shape) - 2, -1, -1):
        strides[i] = strides[i + 1] * max(int(shape[i + 1]), 1)
    return tuple(strides)


def strip_bounds(n_rows: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous y-strip [y0, y1) of rank ``rank``: the first ``n_rows % world_size`` ranks get one
    extra row (10980 rows over 8 GPUs -> 4 x 1373 + 4 x 1372)."""
    if not 0 <= rank < world_size:
        raise ValueError("rank outside [0, world_size)")
    base, extra = divmod(int(n_rows), int(world_size))
    y0 = rank * base + min(rank, extra)
    return y0, y0 + base + (1 if rank < extra else 0)
