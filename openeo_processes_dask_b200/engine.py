"""Array-level engine: numpy arrays / torch CUDA tensors in, same kind out, libeoc_sm100 in between.

* ``numpy.ndarray``      -> host entry points (``eoc_reduce_host`` / ``eoc_select_host``): the strip
                            pipeline copies, reduces and copies back; result is a numpy array.
* ``torch.Tensor`` (cuda)-> device entry points on torch's current stream; result is a cuda tensor.

torch is used for device memory, streams and dtype casts only; every reduction runs in the CUDA
library.  There is no CPU implementation here: without the library or a GPU these functions raise.
"""

from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, Optional, Sequence, Tuple

import numpy as np

from . import _lib as L
from .views import ReduceView, contiguous_strides, plan_view

try:  # torch is the device-memory provider; numpy-only use still works without it
    import torch
except Exception:  # pragma: no cover
    torch = None

_TORCH_TO_NP = {}
_NP_TO_TORCH = {}
if torch is not None:
    _TORCH_TO_NP = {
        torch.float32: np.dtype(np.float32), torch.float64: np.dtype(np.float64),
        torch.uint8: np.dtype(np.uint8), torch.int16: np.dtype(np.int16), torch.int32: np.dtype(np.int32),
        torch.int64: np.dtype(np.int64), torch.bool: np.dtype(np.bool_), torch.int8: np.dtype(np.int8),
        torch.float16: np.dtype(np.float16),
    }
    for _name in ("uint16", "uint32", "uint64"):
        if hasattr(torch, _name):
            _TORCH_TO_NP[getattr(torch, _name)] = np.dtype(getattr(np, _name))
    _NP_TO_TORCH = {v: k for k, v in _TORCH_TO_NP.items()}

# dtypes the kernels read natively; everything else is cast first
_CAST_FIRST = {
    np.dtype(np.bool_): np.dtype(np.uint8), np.dtype(np.int8): np.dtype(np.int16),
    np.dtype(np.int64): np.dtype(np.float64), np.dtype(np.uint32): np.dtype(np.float64),
    np.dtype(np.uint64): np.dtype(np.float64), np.dtype(np.float16): np.dtype(np.float32),
}


def _strip_bytes() -> int:
    """Strip size of the host pipeline in bytes; 0 = the library's default (256 MB pinned / 64 MB pageable input).
    EOC_STRIP_BYTES overrides it (tests use it to force many strips on small cubes)."""
    import os

    try:
        return max(0, int(os.environ.get("EOC_STRIP_BYTES", "0")))
    except ValueError:
        return 0


def is_torch(x) -> bool:
    return torch is not None and isinstance(x, torch.Tensor)


def np_dtype_of(x) -> np.dtype:
    return _TORCH_TO_NP[x.dtype] if is_torch(x) else np.dtype(x.dtype)


def _astype(x, dt: np.dtype):
    if np_dtype_of(x) == dt:
        return x
    return x.to(_NP_TO_TORCH[dt]) if is_torch(x) else x.astype(dt)


def _native(x):
    """Cast to a dtype the kernels read (no-op for f32/f64/u8/i16/u16/i32)."""
    dt = np_dtype_of(x)
    if dt in L.DTYPE_CODE:
        return x
    if dt in _CAST_FIRST:
        return _astype(x, _CAST_FIRST[dt])
    raise TypeError(f"unsupported dtype {dt} for the B200 reducers")


def _strides_elems(x) -> Tuple[int, ...]:
    if is_torch(x):
        return tuple(int(s) for s in x.stride())
    return tuple(int(s) // x.itemsize for s in x.strides)


def _contiguous(x):
    return x.contiguous() if is_torch(x) else np.ascontiguousarray(x)


def _ptr(x) -> int:
    return int(x.data_ptr()) if is_torch(x) else int(x.ctypes.data)


def _empty_like_kind(ref, shape, dt: np.dtype):
    if is_torch(ref):
        return torch.empty(tuple(int(s) for s in shape), dtype=_NP_TO_TORCH[dt], device=ref.device)
    return np.empty(tuple(int(s) for s in shape), dtype=dt)


def _stream_of(ref) -> int:
    return int(torch.cuda.current_stream(ref.device).cuda_stream)


class _DeviceGuard:
    """Make the tensor's device current for the library call (it uses the CUDA current device)."""

    def __init__(self, ref):
        self.ctx = torch.cuda.device(ref.device) if is_torch(ref) else None

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


def acc_dtype(dt: np.dtype) -> np.dtype:
    return np.dtype(np.float32) if dt == np.float32 else np.dtype(np.float64)


def nd_acc_dtype(dt: np.dtype, integer_wrap: bool = False) -> np.dtype:
    """dtype ``eoc_nd_reduce`` computes and writes in: float32 / float64 for float cubes; integer cubes give
    float64 under numpy's integer semantics (``integer_wrap``), else the dtype of the cast-first workflow
    (float32 for 8/16-bit samples, float64 for int32)."""
    if integer_wrap and dt.kind in "iu":
        return np.dtype(np.float64)
    return np.dtype(np.float64) if dt in (np.dtype(np.float64), np.dtype(np.int32)) else np.dtype(np.float32)


def prepare(x, axis, allow_tail: bool = False) -> Tuple[object, ReduceView]:
    """Native dtype + a [outer, R, inner] view (copying to C order only if the layout demands it).

    ``allow_tail``: kept dims BEHIND the reduced ones may be counted into ``outer`` (views.plan_view_pair),
    which covers e.g. a row slice ``cube[:, :, y0:y1]`` of a (t,bands,y,x) cube reduced over t without a copy.
    Only for callers whose outputs are dense [outer, inner] (the kept dims keep their order)."""
    if not is_torch(x):
        x = np.asarray(x)
    elif not x.is_cuda:
        raise TypeError("torch tensors must live on a CUDA device (numpy arrays are the host interface)")
    x = _native(x)
    if x.ndim == 0:
        x = x.reshape(1)
    view = plan_view(tuple(x.shape), _strides_elems(x), axis)
    if view is None and allow_tail and all(n > 0 for n in x.shape):
        from .views import plan_view_pair

        pair = plan_view_pair(tuple(x.shape), _strides_elems(x), _strides_elems(x), axis)
        if pair is not None:
            view = pair[0]
    if view is None:
        from .views import normalize_axes

        axes = normalize_axes(axis, x.ndim)
        if any(b - a != 1 for a, b in zip(axes, axes[1:])):
            # non-adjacent reduced axes: move them together (in front of the rest of the kept tail)
            keep = [d for d in range(x.ndim) if d not in axes]
            lead = [d for d in keep if d < axes[0]]
            tail = [d for d in keep if d > axes[0]]
            perm = lead + list(axes) + tail
            x = x.permute(*perm) if is_torch(x) else np.transpose(x, perm)
            axis = tuple(range(len(lead), len(lead) + len(axes)))
        x = _contiguous(x)
        view = plan_view(tuple(x.shape), contiguous_strides(x.shape), axis)
    return x, view


def _c_view(x, v: ReduceView) -> L.View:
    return L.make_view(_ptr(x), L.DTYPE_CODE[np_dtype_of(x)], v.outer, v.R, v.inner, v.stride_outer, v.stride_R)


def reduce(x, ops: Iterable[int], axis, ignore_nodata: bool = True, ddof: int = 1) -> Dict[int, object]:
    """Streaming reducers ``ops`` over ``axis``; returns {op: array of shape view.out_shape}."""
    ops = list(ops)
    x, v = prepare(x, axis, allow_tail=is_torch(x))
    adt = acc_dtype(np_dtype_of(x))
    outs = {op: _empty_like_kind(x, (v.outer, v.inner), np.dtype(np.int64) if op in L.INT64_OPS else adt)
            for op in ops}
    cv = _c_view(x, v)
    mask = L.ops_mask(ops)
    optr = L.outs_array({op: _ptr(t) for op, t in outs.items()})
    lib = L.lib()
    if v.outer * v.inner > 0:
        if is_torch(x):
            with _DeviceGuard(x):
                ws_bytes = lib.eoc_reduce_workspace_bytes(C.byref(cv), mask)
                ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=x.device)
                L.check(lib.eoc_reduce(C.byref(cv), mask, int(ignore_nodata), int(ddof), optr,
                                       C.c_void_p(ws.data_ptr()), ws_bytes, C.c_void_p(_stream_of(x))))
        else:
            try:
                L.check(lib.eoc_reduce_host(C.byref(cv), mask, int(ignore_nodata), int(ddof), optr, _strip_bytes()))
            except L.EocError as e:
                if e.code != -2:
                    raise
                # long-row (spatial) reduction: stage the whole array on the device instead
                dev = reduce(torch.from_numpy(x).cuda(), ops, _axes_of(v, x), ignore_nodata, ddof)
                return {op: t.cpu().numpy() for op, t in dev.items()}
    return {op: t.reshape(v.out_shape) for op, t in outs.items()}


def _axes_of(v: ReduceView, x) -> Tuple[int, ...]:
    n_red = len(v.keep_shape) - len(v.out_shape)
    return tuple(range(v.axis_pos, v.axis_pos + n_red))


def select(x, axis, ignore_nodata: bool = True, want_median: bool = True,
           probs: Optional[Sequence[float]] = None, quant_dtype=np.float64,
           ops: Optional[Iterable[int]] = None, ddof: int = 1):
    """Order statistics over ``axis``: (median or None, quantiles [.., nq at axis_pos, ..] or None).

    With ``ops`` (streaming reducer ids) the same pass also produces those reducers
    (``eoc_select_fused``) and the return value is (median, quantiles, {op: array}).
    """
    fused = list(ops) if ops is not None else None
    x, v = prepare(x, axis)
    if v.R > L.SELECT_MAX_R:
        # 513 .. 1024 float32 samples of a device cube: the eight-run kernel takes them when it can address the view
        # (the library answers EOC_ERR_UNSUPPORTED otherwise); everything else goes the slower general path
        nprobs = 0 if probs is None else len(probs)
        fast_dt = np_dtype_of(x) in (np.dtype(np.float32), np.dtype(np.uint16), np.dtype(np.int16), np.dtype(np.uint8))
        if not (is_torch(x) and v.R <= 1024 and fast_dt and not fused and nprobs <= 4):
            return _select_general(x, v, ignore_nodata, want_median, probs, quant_dtype, fused, ddof)
        try:
            return _select_device(x, v, ignore_nodata, want_median, probs, quant_dtype, fused, ddof)
        except L.EocError as e:
            if e.code != -2:
                raise
        # the view is not one the eight-run kernel can address (the reduced axis is the contiguous one, or `inner` is
        # short): a transposed copy with the reduced axis in front is -- two passes of HBM, ~100 x cheaper than the
        # general path (median over x of a (100, 4, 512, 1024) cube: 8 GB/s before)
        rest = int(v.outer * v.inner)
        if v.axis_pos is not None and rest >= 65536 and rest % 4 == 0 and len(v.out_shape) == x.dim() - 1:
            axis_i = v.axis_pos
            xt = x.movedim(axis_i, 0).contiguous()
            xt2, vt = prepare(xt, 0)
            try:
                res = _select_device(xt2, vt, ignore_nodata, want_median, probs, quant_dtype, fused, ddof)
                med, qout = res[0], res[1]
                if qout is not None:      # [nq, rest...] -> the quantile axis where the reduced axis was
                    qout = qout.movedim(0, axis_i)
                return (med, qout, {}) if fused is not None else (med, qout)
            except L.EocError as e:
                if e.code != -2:
                    raise
        return _select_general(x, v, ignore_nodata, want_median, probs, quant_dtype, fused, ddof)
    return _select_device(x, v, ignore_nodata, want_median, probs, quant_dtype, fused, ddof)


def _select_device(x, v, ignore_nodata, want_median, probs, quant_dtype, fused, ddof):
    """``select`` through eoc_select_fused / eoc_select_fused_host on a prepared view."""
    adt = acc_dtype(np_dtype_of(x))
    probs = np.ascontiguousarray(np.asarray(probs if probs is not None else [], dtype=np.float64))
    nq = int(probs.size)
    if nq > L.SELECT_MAX_Q:
        raise NotImplementedError(f"more than {L.SELECT_MAX_Q} probabilities per call")
    qdt = np.dtype(quant_dtype)
    med = _empty_like_kind(x, (v.outer, v.inner), adt) if want_median else None
    qout = _empty_like_kind(x, (v.outer, nq, v.inner), qdt) if nq else None
    cv = _c_view(x, v)
    pp = probs.ctypes.data_as(C.POINTER(C.c_double)) if nq else None
    qcode = L.F64 if qdt == np.float64 else L.F32
    lib = L.lib()
    fouts = {}
    if fused:
        fouts = {op: _empty_like_kind(x, (v.outer, v.inner), np.dtype(np.int64) if op in L.INT64_OPS else adt)
                 for op in fused}
    if fused and (v.R == 0 or (med is None and qout is None)):
        # nothing to select (or an empty axis): the streaming entry point owns those cases
        red = reduce(x, fused, _axes_of(v, x), ignore_nodata, ddof)
        fouts = {op: t.reshape(v.outer, v.inner) for op, t in red.items()}
    if v.outer * v.inner > 0 and v.R > 0 and (med is not None or qout is not None):
        mptr = C.c_void_p(_ptr(med)) if med is not None else None
        qptr = C.c_void_p(_ptr(qout)) if qout is not None else None
        mask = L.ops_mask(fused or [])
        optr = L.outs_array({op: _ptr(t) for op, t in fouts.items()})
        if is_torch(x):
            with _DeviceGuard(x):
                L.check(lib.eoc_select_fused(C.byref(cv), int(ignore_nodata), mptr, nq, pp, qptr, qcode, mask,
                                             int(ddof), optr, C.c_void_p(_stream_of(x))))
        else:
            L.check(lib.eoc_select_fused_host(C.byref(cv), int(ignore_nodata), mptr, nq, pp, qptr, qcode, mask,
                                              int(ddof), optr, _strip_bytes()))
    elif v.R == 0:
        for t in (med, qout):
            if t is not None:
                t[...] = float("nan")
    if med is not None:
        med = med.reshape(v.out_shape)
    if qout is not None:
        lead = v.out_shape[:v.axis_pos]
        qout = qout.reshape(tuple(lead) + (nq,) + tuple(v.out_shape[v.axis_pos:]))
    if fused is not None:
        return med, qout, {op: t.reshape(v.out_shape) for op, t in fouts.items()}
    return med, qout


def _general(x, v: ReduceView, ignore_nodata, med, qout, probs, qcode, sort_out=None, desc=False, nan_first=False):
    """One call of ``eoc_select_general`` on a CUDA tensor (any R): the workspace is a torch allocation of the
    recommended size."""
    cv = _c_view(x, v)
    lib = L.lib()
    nq = int(probs.size)
    pp = probs.ctypes.data_as(C.POINTER(C.c_double)) if nq else None
    with _DeviceGuard(x):
        ws_bytes = int(lib.eoc_select_general_workspace_bytes(C.byref(cv)))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=x.device)
        L.check(lib.eoc_select_general(
            C.byref(cv), int(ignore_nodata), C.c_void_p(_ptr(med)) if med is not None else None, nq, pp,
            C.c_void_p(_ptr(qout)) if qout is not None else None, qcode,
            C.c_void_p(_ptr(sort_out)) if sort_out is not None else None, int(desc), int(nan_first),
            C.c_void_p(ws.data_ptr()), ws_bytes, C.c_void_p(_stream_of(x))))


def _select_general(x, v: ReduceView, ignore_nodata, want_median, probs, quant_dtype, fused, ddof):
    """``select`` for an axis longer than SELECT_MAX_R: the slower general path (``eoc_select_general``); a numpy
    array is staged on the device as a whole; fused streaming reducers run as their own launch."""
    was_numpy = not is_torch(x)
    if was_numpy:
        x = torch.from_numpy(np.ascontiguousarray(x)).cuda()
        v = plan_view(tuple(x.shape), contiguous_strides(x.shape), _axes_of(v, x))
    adt = acc_dtype(np_dtype_of(x))
    probs = np.ascontiguousarray(np.asarray(probs if probs is not None else [], dtype=np.float64))
    nq = int(probs.size)
    if nq > L.SELECT_MAX_Q:
        raise NotImplementedError(f"more than {L.SELECT_MAX_Q} probabilities per call")
    qdt = np.dtype(quant_dtype)
    med = _empty_like_kind(x, (v.outer, v.inner), adt) if want_median else None
    qout = _empty_like_kind(x, (v.outer, nq, v.inner), qdt) if nq else None
    if v.outer * v.inner > 0 and (med is not None or qout is not None):
        _general(x, v, ignore_nodata, med, qout, probs, L.F64 if qdt == np.float64 else L.F32)
    fouts = {}
    if fused:
        fouts = reduce(x, fused, _axes_of(v, x), ignore_nodata, ddof)
    if med is not None:
        med = med.reshape(v.out_shape)
    if qout is not None:
        qout = qout.reshape(tuple(v.out_shape[:v.axis_pos]) + (nq,) + tuple(v.out_shape[v.axis_pos:]))
    if was_numpy:
        med = med.cpu().numpy() if med is not None else None
        qout = qout.cpu().numpy() if qout is not None else None
        fouts = {op: t.cpu().numpy() for op, t in fouts.items()}
    if fused is not None:
        return med, qout, fouts
    return med, qout


def segmented_reduce(x, axis: int, groups: Sequence[Sequence[int]], ops: Iterable[int],
                     ignore_nodata: bool = True, ddof: int = 1) -> Dict[int, object]:
    """Reduce index groups of ``axis`` independently -> arrays with len(groups) along ``axis``."""
    ops = list(ops)
    was_numpy = not is_torch(x)
    x, v = prepare(x, int(axis))
    adt = acc_dtype(np_dtype_of(x))
    nseg = len(groups)
    offsets = np.zeros(nseg + 1, dtype=np.int32)
    offsets[1:] = np.cumsum([len(g) for g in groups])
    members = np.ascontiguousarray(np.asarray([i for g in groups for i in g], dtype=np.int32))
    if members.size and (members.min() < 0 or members.max() >= v.R):
        raise IndexError("group member outside the reduced axis")
    if was_numpy:  # host strip pipeline: H2D of strip k+1 overlaps the segmented kernel and the D2H of strip k
        outs = {op: np.empty((v.outer, nseg, v.inner), dtype=np.dtype(np.int64) if op in L.INT64_OPS else adt) for op in ops}
        if v.outer * v.inner > 0 and nseg > 0:
            cv = _c_view(x, v)
            mem = members if members.size else np.zeros(1, np.int32)
            L.check(L.lib().eoc_segmented_reduce_host(
                C.byref(cv), nseg, C.c_void_p(offsets.ctypes.data), C.c_void_p(mem.ctypes.data),
                int(max((len(g) for g in groups), default=0)), L.ops_mask(ops), int(ignore_nodata), int(ddof),
                L.outs_array({op: _ptr(t) for op, t in outs.items()}), _strip_bytes()))
        shape = tuple(v.out_shape[:v.axis_pos]) + (nseg,) + tuple(v.out_shape[v.axis_pos:])
        return {op: t.reshape(shape) for op, t in outs.items()}
    d_off = torch.from_numpy(offsets).to(x.device)
    d_mem = torch.from_numpy(members if members.size else np.zeros(1, np.int32)).to(x.device)
    outs = {op: _empty_like_kind(x, (v.outer, nseg, v.inner), np.dtype(np.int64) if op in L.INT64_OPS else adt)
            for op in ops}
    cv = _c_view(x, v)
    optr = L.outs_array({op: _ptr(t) for op, t in outs.items()})
    if v.outer * v.inner > 0 and nseg > 0:
        with _DeviceGuard(x):
            L.check(L.lib().eoc_segmented_reduce(
                C.byref(cv), nseg, C.c_void_p(d_off.data_ptr()), C.c_void_p(d_mem.data_ptr()),
                int(max((len(g) for g in groups), default=0)), L.ops_mask(ops), int(ignore_nodata), int(ddof),
                optr, C.c_void_p(_stream_of(x))))
    shape = tuple(v.out_shape[:v.axis_pos]) + (nseg,) + tuple(v.out_shape[v.axis_pos:])
    return {op: t.reshape(shape) for op, t in outs.items()}


def nd_reduce(a, b, axis, ops: Iterable[int], ignore_nodata: bool = True, ddof: int = 1, want_nd: bool = False,
              integer_wrap: bool = False):
    """(a-b)/(a+b) fused with reducers over ``axis`` (``axis=None`` and no ops: elementwise only).
    ``integer_wrap``: integer operands follow numpy (difference and sum wrap in their dtype, float64 quotient)
    instead of being converted to float first.

    Returns (outs dict, nd array or None).
    """
    ops = list(ops)
    was_numpy = not is_torch(a)
    if was_numpy:
        a, b = _native(np.asarray(a)), _native(np.asarray(b))
    if tuple(a.shape) != tuple(b.shape) or a.dtype != b.dtype:
        raise ValueError("normalized difference operands must have the same shape and dtype")
    if a.ndim == 0:
        a, b = a.reshape(1), b.reshape(1)
    elementwise_only = not ops
    red_axis = 0 if elementwise_only else axis
    a, va = prepare(a, red_axis)
    b, vb = prepare(b, red_axis)
    adt = nd_acc_dtype(np_dtype_of(a), integer_wrap)
    outs = {op: _empty_like_kind(a, (va.outer, va.inner), np.dtype(np.int64) if op in L.INT64_OPS else adt)
            for op in ops}
    nd = _empty_like_kind(a, (va.outer, va.R, va.inner), adt) if (want_nd or elementwise_only) else None
    ca, cb = _c_view(a, va), _c_view(b, vb)
    optr = L.outs_array({op: _ptr(t) for op, t in outs.items()})
    if was_numpy:
        if va.outer * va.inner * va.R > 0:
            L.check(L.lib().eoc_nd_reduce_host(C.byref(ca), C.byref(cb), L.ops_mask(ops), int(ignore_nodata), int(ddof),
                                               optr, C.c_void_p(_ptr(nd)) if nd is not None else None,
                                               int(bool(integer_wrap)), _strip_bytes()))
    elif va.outer * va.inner * va.R > 0:
        with _DeviceGuard(a):
            L.check(L.lib().eoc_nd_reduce(C.byref(ca), C.byref(cb), L.ops_mask(ops), int(ignore_nodata), int(ddof),
                                          optr, C.c_void_p(_ptr(nd)) if nd is not None else None,
                                          int(bool(integer_wrap)), C.c_void_p(_stream_of(a))))
    res = {op: t.reshape(va.out_shape) for op, t in outs.items()}
    if nd is not None:
        nd = nd.reshape(tuple(a.shape))
    return res, nd


# ---- axis-kept array processes ----------------------------------------------------------
def scan(x, op: int, axis, ignore_nodata: bool = True):
    """Cumulative sum / product / min / max along ``axis`` (``eoc_scan``); result has the shape of ``x`` and the
    compute dtype.  numpy input is staged on the device and the result copied back."""
    was_numpy = not is_torch(x)
    x, v = prepare(x, axis)
    out = _empty_like_kind(x, (v.outer, v.R, v.inner), acc_dtype(np_dtype_of(x)))
    if v.outer * v.R * v.inner > 0:
        cv = _c_view(x, v)
        if was_numpy:
            L.check(L.lib().eoc_scan_host(C.byref(cv), int(op), int(ignore_nodata), C.c_void_p(_ptr(out)), _strip_bytes()))
        else:
            with _DeviceGuard(x):
                L.check(L.lib().eoc_scan(C.byref(cv), int(op), int(ignore_nodata), C.c_void_p(_ptr(out)),
                                         C.c_void_p(_stream_of(x))))
    return out.reshape(tuple(x.shape))


def sort(x, axis, descending: bool = False, nan_first: bool = False):
    """Every column sorted along ``axis`` (``eoc_sort``): NaNs last (numpy's order) or first; result has the
    shape of ``x`` and the compute dtype."""
    was_numpy = not is_torch(x)
    if was_numpy:
        xn, vn = prepare(x, axis)
        if vn.R <= L.SELECT_MAX_R:  # host strip pipeline
            out = np.empty((vn.outer, vn.R, vn.inner), dtype=acc_dtype(np_dtype_of(xn)))
            if out.size > 0:
                cv = _c_view(xn, vn)
                L.check(L.lib().eoc_sort_host(C.byref(cv), int(descending), int(nan_first), C.c_void_p(_ptr(out)),
                                              _strip_bytes()))
            return out.reshape(tuple(xn.shape))
        x = torch.from_numpy(np.ascontiguousarray(_native(np.asarray(x)))).cuda()  # any length: staged as a whole
    x, v = prepare(x, axis)
    out = _empty_like_kind(x, (v.outer, v.R, v.inner), acc_dtype(np_dtype_of(x)))
    if v.R > L.SELECT_MAX_R and v.outer * v.inner > 0:  # any length: the slower general path
        _general(x, v, True, None, None, np.zeros(0), L.F64, sort_out=out, desc=descending, nan_first=nan_first)
    elif v.outer * v.R * v.inner > 0:
        cv = _c_view(x, v)
        with _DeviceGuard(x):
            L.check(L.lib().eoc_sort(C.byref(cv), int(descending), int(nan_first), C.c_void_p(_ptr(out)),
                                     C.c_void_p(_stream_of(x))))
    out = out.reshape(tuple(x.shape))
    return out.cpu().numpy() if was_numpy else out


def order(x, axis):
    """Stable ascending argsort along ``axis`` with the NaNs last (``eoc_order``): int64, shape of ``x``."""
    was_numpy = not is_torch(x)
    if was_numpy:
        x = torch.from_numpy(np.ascontiguousarray(_native(np.asarray(x)))).cuda()
    x, v = prepare(x, axis)
    if v.R > L.SELECT_MAX_R:
        raise NotImplementedError(f"order along an axis of length {v.R} > {L.SELECT_MAX_R} is not implemented on the device")
    out = _empty_like_kind(x, (v.outer, v.R, v.inner), np.dtype(np.int64))
    if v.outer * v.R * v.inner > 0:
        cv = _c_view(x, v)
        with _DeviceGuard(x):
            L.check(L.lib().eoc_order(C.byref(cv), C.c_void_p(_ptr(out)), C.c_void_p(_stream_of(x))))
    out = out.reshape(tuple(x.shape))
    return out.cpu().numpy() if was_numpy else out


def take(x, index, axis):
    """``np.take(x, index, axis)`` for a 1-D integer ``index`` (``eoc_take``): a gather of planes, input dtype kept."""
    was_numpy = not is_torch(x)
    if was_numpy:
        x = torch.from_numpy(np.ascontiguousarray(np.asarray(x))).cuda()
    if np_dtype_of(x) not in L.DTYPE_CODE and x.element_size() in (1, 2, 4, 8):
        code = {1: L.U8, 2: L.U16, 4: L.F32, 8: L.F64}[x.element_size()]   # bytes are bytes for a gather
    else:
        code = L.DTYPE_CODE[np_dtype_of(x)]
    if any(st < 0 for st in x.stride()):
        x = x.contiguous()
    view = plan_view(tuple(x.shape), _strides_elems(x), axis)
    if view is None:
        x = x.contiguous()
        view = plan_view(tuple(x.shape), contiguous_strides(x.shape), axis)
    idx = np.ascontiguousarray(np.asarray(index, dtype=np.int64).reshape(-1))
    n = int(x.shape[axis])
    if idx.size and (idx.min() < -n or idx.max() >= n):
        raise IndexError(f"index {int(idx.max() if idx.max() >= n else idx.min())} is out of bounds for axis {axis} with size {n}")
    d_idx = torch.from_numpy(idx if idx.size else np.zeros(1, np.int64)).to(x.device)
    out = torch.empty((view.outer, idx.size, view.inner), dtype=x.dtype, device=x.device)
    if out.numel() > 0:
        cv = L.make_view(_ptr(x), code, view.outer, view.R, view.inner, view.stride_outer, view.stride_R)
        with _DeviceGuard(x):
            L.check(L.lib().eoc_take(C.byref(cv), C.c_void_p(d_idx.data_ptr()), idx.size, C.c_void_p(out.data_ptr()),
                                     C.c_void_p(_stream_of(x))))
    ax = axis % x.ndim
    out = out.reshape(tuple(x.shape[:ax]) + (idx.size,) + tuple(x.shape[ax + 1:]))
    return out.cpu().numpy() if was_numpy else out


def interpolate_linear(x, axis, positions=None):
    """Linear interpolation of the non-finite gaps along ``axis`` (``eoc_interpolate_linear``); ``positions``:
    R float64 sample positions (None: 0..R-1).  Result has the shape and dtype of ``x`` (float32 / float64)."""
    was_numpy = not is_torch(x)
    if was_numpy:  # host strip pipeline: numpy in -> numpy out, copies overlapped with the kernel
        xn, vn = prepare(x, axis)
        out = np.empty((vn.outer, vn.R, vn.inner), dtype=np_dtype_of(xn))
        if out.size > 0:
            pos = None
            if positions is not None:
                pos = np.ascontiguousarray(np.asarray(positions, dtype=np.float64))
                if pos.size != vn.R:
                    raise ValueError("positions must have one entry per sample along the axis")
            cv = _c_view(xn, vn)
            L.check(L.lib().eoc_interpolate_linear_host(C.byref(cv), C.c_void_p(pos.ctypes.data) if pos is not None else None,
                                                        C.c_void_p(_ptr(out)), _strip_bytes()))
        return out.reshape(tuple(xn.shape))
    x, v = prepare(x, axis)
    out = _empty_like_kind(x, (v.outer, v.R, v.inner), np_dtype_of(x))
    if v.outer * v.R * v.inner > 0:
        cv = _c_view(x, v)
        xs = None
        if positions is not None:
            xs = torch.from_numpy(np.ascontiguousarray(np.asarray(positions, dtype=np.float64))).to(x.device)
            if xs.numel() != v.R:
                raise ValueError("positions must have one entry per sample along the axis")
        with _DeviceGuard(x):
            L.check(L.lib().eoc_interpolate_linear(C.byref(cv), C.c_void_p(xs.data_ptr()) if xs is not None else None,
                                                   C.c_void_p(_ptr(out)), C.c_void_p(_stream_of(x))))
    out = out.reshape(tuple(x.shape))
    return out.cpu().numpy() if was_numpy else out


def elementwise(x, op: int, params):
    """``eoc_elementwise`` over the whole array; result has the shape of ``x`` and the compute dtype."""
    was_numpy = not is_torch(x)
    pp = (C.c_double * 4)(*([float(p) for p in params] + [0.0] * (4 - len(params))))
    if was_numpy:  # host strip pipeline
        xn = np.ascontiguousarray(_native(np.asarray(x)))
        out = np.empty(xn.shape, dtype=acc_dtype(xn.dtype))
        if xn.size > 0:
            L.check(L.lib().eoc_elementwise_host(C.c_void_p(xn.ctypes.data), L.DTYPE_CODE[np.dtype(xn.dtype)], xn.size, int(op),
                                                 pp, C.c_void_p(out.ctypes.data), _strip_bytes()))
        return out
    x = _contiguous(_native(x))
    out = _empty_like_kind(x, tuple(x.shape), acc_dtype(np_dtype_of(x)))
    if x.numel() > 0:
        with _DeviceGuard(x):
            L.check(L.lib().eoc_elementwise(C.c_void_p(x.data_ptr()), L.DTYPE_CODE[np_dtype_of(x)], x.numel(), int(op), pp,
                                            C.c_void_p(out.data_ptr()), C.c_void_p(_stream_of(x))))
    return out.cpu().numpy() if was_numpy else out


def apply_tape(x, y, steps, compute: np.dtype):
    """``eoc_apply_tape``: the accumulator program ``steps`` = [(code, operand, imm)] over the samples of ``x`` (and
    ``y``, same shape and dtype, or None); result in ``compute`` (float32 / float64), same shape.  numpy in ->
    numpy out."""
    was_numpy = not is_torch(x)
    compute = np.dtype(compute)
    if was_numpy and (y is None or not is_torch(y)):  # host strip pipeline
        xn = np.ascontiguousarray(_native(np.asarray(x)))
        yn = None if y is None else np.ascontiguousarray(_native(np.asarray(y)))
        if yn is not None and (yn.shape != xn.shape or yn.dtype != xn.dtype):
            raise ValueError("apply_tape: x and y differ in shape, dtype or device")
        if len(steps) > L.TAPE_MAX_OPS:
            raise ValueError(f"apply_tape: {len(steps)} steps, at most {L.TAPE_MAX_OPS}")
        out = np.empty(xn.shape, dtype=compute)
        hops = (L.TapeOp * max(len(steps), 1))()
        for k, (code, operand, imm) in enumerate(steps):
            hops[k].code, hops[k].operand, hops[k].imm = int(code), int(operand), float(imm)
        if xn.size > 0:
            L.check(L.lib().eoc_apply_tape_host(C.c_void_p(xn.ctypes.data), C.c_void_p(yn.ctypes.data) if yn is not None else None,
                                                L.DTYPE_CODE[np.dtype(xn.dtype)], xn.size, hops, len(steps),
                                                C.c_void_p(out.ctypes.data), L.DTYPE_CODE[compute], _strip_bytes()))
        return out
    if was_numpy:
        x = torch.from_numpy(np.ascontiguousarray(_native(np.asarray(x)))).cuda()
        if y is not None:
            y = torch.from_numpy(np.ascontiguousarray(_native(np.asarray(y)))).to(x.device)
    if y is not None and (tuple(y.shape) != tuple(x.shape) or y.dtype != x.dtype or y.device != x.device):
        raise ValueError("apply_tape: x and y differ in shape, dtype or device")
    if len(steps) > L.TAPE_MAX_OPS:
        raise ValueError(f"apply_tape: {len(steps)} steps, at most {L.TAPE_MAX_OPS}")
    # one dense traversal order for both inputs: x's layout when it is dense in some order, else C order
    if y is not None and x.stride() != y.stride():
        x, y = _contiguous(x), _contiguous(y)
    if not (x.is_contiguous() or _dense_any_order(x)):
        x = _contiguous(x)
        if y is not None:
            y = _contiguous(y)
    out = torch.empty_strided(tuple(x.shape), x.stride(), dtype=_NP_TO_TORCH[compute], device=x.device)
    ops = (L.TapeOp * max(len(steps), 1))()
    for k, (code, operand, imm) in enumerate(steps):
        ops[k].code, ops[k].operand, ops[k].imm = int(code), int(operand), float(imm)
    if x.numel() > 0:
        with _DeviceGuard(x):
            L.check(L.lib().eoc_apply_tape(C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()) if y is not None else None,
                                           L.DTYPE_CODE[np_dtype_of(x)], x.numel(), ops, len(steps),
                                           C.c_void_p(out.data_ptr()), L.DTYPE_CODE[compute], C.c_void_p(_stream_of(x))))
    return out.cpu().numpy() if was_numpy else out


def _dense_any_order(t) -> bool:
    """the tensor covers numel() consecutive elements in SOME axis order (a permuted contiguous block)"""
    dims = sorted(((st, sz) for st, sz in zip(t.stride(), t.shape) if sz > 1), key=lambda p: p[0])
    expect = 1
    for st, sz in dims:
        if st != expect:
            return False
        expect *= sz
    return True


_BIN_ARITH = (L.BIN_ADD, L.BIN_SUBTRACT, L.BIN_MULTIPLY, L.BIN_DIVIDE)


def _is_host_scalar(v) -> bool:
    return not is_torch(v) and np.ndim(v) == 0


def binary_result_dtype(op: int, common: np.dtype) -> np.dtype:
    """dtype ``eoc_binary`` writes for operands of dtype ``common`` (numpy's: math.py:98-115, np.where(.., nan))."""
    if op in _BIN_ARITH:
        return np.dtype(np.float64) if (op == L.BIN_DIVIDE and common.kind in "iu") else common
    return np.dtype(np.float64)


def binary(x, y, op: int):
    """``eoc_binary``: x OP y with numpy broadcasting and numpy's (NEP 50) type promotion; python scalars are
    weak.  numpy / scalar operands -> numpy result, a CUDA tensor on either side -> CUDA tensor."""
    any_torch = is_torch(x) or is_torch(y)
    sx, sy = _is_host_scalar(x), _is_host_scalar(y)
    if sx and sy:  # two host scalars: still one (tiny) launch, no host arithmetic in the product path
        x, sx = np.asarray(x).reshape(1), False
        if x.dtype.kind in "iu" and x.dtype.itemsize == 8:  # python int: int32 when it fits (result value is the same)
            x = x.astype(np.int32) if abs(int(x[0])) < 2**31 else x.astype(np.float64)
        shape0 = ()
    else:
        shape0 = None
    ax = x if (sx or is_torch(x)) else np.asarray(x)
    ay = y if (sy or is_torch(y)) else np.asarray(y)
    dts = [v if s else np_dtype_of(v) for v, s in ((ax, sx), (ay, sy))]
    dts = [np.dtype(np.uint8) if (isinstance(d, np.dtype) and d == np.bool_) else d for d in dts]
    dts = [int(d) if isinstance(d, (bool, np.bool_)) else d for d in dts]
    common = np.result_type(*dts)
    if common == np.float16:
        common = np.dtype(np.float32)
    if common not in L.DTYPE_CODE:
        raise TypeError(f"binary process on operands of common dtype {common} is not supported by the B200 kernels")
    dev = (x if is_torch(x) else y).device if any_torch else None

    def as_dev(v):
        if not is_torch(v):
            v = np.asarray(v)
            if v.dtype == np.bool_:
                v = v.view(np.uint8)
            v = torch.from_numpy(np.ascontiguousarray(v)).to(dev if dev is not None else "cuda")
        elif v.dtype == torch.bool:
            v = v.view(torch.uint8) if v.is_contiguous() else v.to(torch.uint8)
        return _astype(v, common)

    scalar, side = 0.0, 0
    if sx:
        scalar, side = float(np.asarray(x, dtype=common)), 1
        tb = as_dev(ay)
        ta = tb
    elif sy:
        scalar, side = float(np.asarray(y, dtype=common)), 2
        ta = as_dev(ax)
        tb = ta
    else:
        ta, tb = as_dev(ax), as_dev(ay)
    if side:
        ta = tb = _contiguous(ta)
        shape = tuple(ta.shape)
        sa = sb = contiguous_strides(shape)
    else:
        shape = tuple(np.broadcast_shapes(tuple(ta.shape), tuple(tb.shape)))
        if len(shape) > L.BINARY_MAX_DIMS:
            raise ValueError(f"binary process: more than {L.BINARY_MAX_DIMS} dimensions")

        def bstrides(t):
            if any(st < 0 for st in t.stride()):
                t = t.contiguous()
            pad = len(shape) - t.dim()
            st = [0] * pad + [int(v) for v in t.stride()]
            sh = [1] * pad + [int(v) for v in t.shape]
            return t, tuple(0 if sh[d] == 1 and shape[d] != 1 else st[d] for d in range(len(shape)))

        ta, sa = bstrides(ta)
        tb, sb = bstrides(tb)
    out = torch.empty(shape, dtype=_NP_TO_TORCH[binary_result_dtype(op, common)], device=ta.device)
    if out.numel() > 0:
        nd = len(shape)
        arr = C.c_int64 * max(nd, 1)
        with _DeviceGuard(ta):
            L.check(L.lib().eoc_binary(C.c_void_p(ta.data_ptr()), C.c_void_p(tb.data_ptr()), L.DTYPE_CODE[common], nd,
                                       arr(*shape), arr(*sa), arr(*sb), int(op), float(scalar), side,
                                       C.c_void_p(out.data_ptr()), C.c_void_p(_stream_of(ta))))
    if shape0 is not None:
        out = out.reshape(shape0)
    return out if any_torch else out.cpu().numpy()


# ---- mask fused into the reduce -----------------------------------------------------
def mask_canonical(m):
    """uint8 CUDA tensor, 1 where the mask is truthy and not NaN (the samples ``data.where(_not(mask), r)``
    replaces, cubes/mask.py:112 + logic.py:51-54), same shape, dense.  bool tensors are reinterpreted."""
    if not is_torch(m):
        m = torch.from_numpy(np.ascontiguousarray(np.asarray(m))).cuda()
    if m.dtype == torch.bool:
        return m.contiguous().view(torch.uint8)
    m = _contiguous(_native(m))
    out = torch.empty(tuple(m.shape), dtype=torch.uint8, device=m.device)
    with _DeviceGuard(m):
        L.check(L.lib().eoc_mask_canonical(C.c_void_p(m.data_ptr()), L.DTYPE_CODE[np_dtype_of(m)], m.numel(),
                                           C.c_void_p(out.data_ptr()), C.c_void_p(_stream_of(m))))
    return out


def _masked_views(x, m, axis):
    """Native data tensor, mask tensor expanded to its shape, and the paired [outer, R, inner] views."""
    from .views import plan_view_pair

    x = _native(x)
    if x.ndim == 0:
        x, m = x.reshape(1), m.reshape(1)
    if tuple(m.shape) != tuple(x.shape):
        m = m.expand(tuple(x.shape))
    pair = plan_view_pair(tuple(x.shape), _strides_elems(x), _strides_elems(m), axis)
    if pair is None:  # layouts that do not pair up: dense copies of both
        x, m = _contiguous(x), m.contiguous()
        pair = plan_view_pair(tuple(x.shape), _strides_elems(x), _strides_elems(m), axis)
    if pair is None:
        raise NotImplementedError("masked reduction over non-adjacent axes")
    vx, vm, tail = pair
    return x, m, vx, vm, tail


def _elementwise_axis(x, m) -> int:
    """For the materialised ``mask`` process any axis may play R: take one whose paired view needs no
    re-ordering of the dense output (no kept dim behind R in ``outer``) and has the longest ``inner``."""
    from .views import plan_view_pair

    if x.ndim == 0:
        return 0
    mm = m.expand(tuple(x.shape)) if tuple(m.shape) != tuple(x.shape) else m
    best, best_inner = 0, -1
    for ax in range(x.ndim):
        pair = plan_view_pair(tuple(x.shape), _strides_elems(x), _strides_elems(mm), ax)
        if pair is not None and pair[2] == 0 and pair[0].inner > best_inner:
            best, best_inner = ax, pair[0].inner
    return best


def _mask_c_view(m, v: ReduceView) -> L.View:
    return L.make_view(_ptr(m), L.U8, v.outer, v.R, v.inner, v.stride_outer, v.stride_R)


def masked_reduce(x, m, replacement, ops: Iterable[int], axis, ignore_nodata: bool = True, ddof: int = 1,
                  want_masked: bool = False):
    """``where(mask, replacement, x)`` fused with streaming reducers over ``axis`` (``eoc_masked_reduce``).

    ``x``: CUDA tensor; ``m``: canonical uint8 CUDA tensor broadcastable to ``x`` (``expand`` views with stride
    0 are used as they are).  Returns ({op: array}, masked cube or None); with no ops only the masked cube.
    """
    ops = list(ops)
    elementwise_only = not ops
    if elementwise_only:
        axis = _elementwise_axis(x, m)
    x, m, vx, vm, tail = _masked_views(x, m, axis)
    adt = acc_dtype(np_dtype_of(x))
    outs = {op: _empty_like_kind(x, (vx.outer, vx.inner), np.dtype(np.int64) if op in L.INT64_OPS else adt)
            for op in ops}
    dense = None
    if want_masked or elementwise_only:
        dense = _empty_like_kind(x, (vx.outer, vx.R, vx.inner), adt)
    optr = L.outs_array({op: _ptr(t) for op, t in outs.items()})
    cx, cm = _c_view(x, vx), _mask_c_view(m, vm)
    if vx.outer * vx.R * vx.inner > 0:
        with _DeviceGuard(x):
            L.check(L.lib().eoc_masked_reduce(C.byref(cx), C.byref(cm), float(replacement), L.ops_mask(ops),
                                              int(ignore_nodata), int(ddof), optr,
                                              C.c_void_p(_ptr(dense)) if dense is not None else None,
                                              C.c_void_p(_stream_of(x))))
    res = {op: t.reshape(vx.out_shape) for op, t in outs.items()}
    if dense is not None:
        # dense is [outer, R, inner]; with outer dims that sit behind R in the cube, put R back in place
        dense = _unview(dense, x, axis, tail)
    return res, dense


def _unview(dense, x, axis, tail: int):
    """dense [outer, R, inner] of a paired view -> the cube's own dim order (``tail`` kept dims that follow
    the reduced ones in the cube were counted into ``outer``)."""
    from .views import normalize_axes

    axes = normalize_axes(axis, x.ndim)
    first, last = axes[0], axes[-1]
    shape = tuple(x.shape)
    lead, red = shape[:first], shape[first:last + 1]
    tail_outer, inner = shape[last + 1:last + 1 + tail], shape[last + 1 + tail:]
    t = dense.reshape(*lead, *tail_outer, *red, *inner)
    if tail:
        nl, nt, nr = len(lead), len(tail_outer), len(red)
        perm = (list(range(nl)) + list(range(nl + nt, nl + nt + nr)) + list(range(nl, nl + nt))
                + list(range(nl + nt + nr, t.ndim)))
        t = t.permute(*perm).contiguous()
    return t


def masked_select(x, m, axis, ignore_nodata: bool = True, want_median: bool = True,
                  probs: Optional[Sequence[float]] = None, quant_dtype=np.float64, replacement: float = float("nan")):
    """Median / quantiles of ``where(mask, replacement, x)`` over ``axis`` in ONE pass (``eoc_masked_select``).
    Raises ``_lib.EocError`` with code -2 when the fused kernel does not cover the case -- a replacement other
    than NaN among them -- (the caller then materialises the masked cube and calls ``select``)."""
    x, m, vx, vm, tail = _masked_views(x, m, axis)
    adt = acc_dtype(np_dtype_of(x))
    probs = np.ascontiguousarray(np.asarray(probs if probs is not None else [], dtype=np.float64))
    nq = int(probs.size)
    qdt = np.dtype(quant_dtype)
    med = _empty_like_kind(x, (vx.outer, vx.inner), adt) if want_median else None
    qout = _empty_like_kind(x, (vx.outer, nq, vx.inner), qdt) if nq else None
    cx, cm = _c_view(x, vx), _mask_c_view(m, vm)
    pp = probs.ctypes.data_as(C.POINTER(C.c_double)) if nq else None
    with _DeviceGuard(x):
        L.check(L.lib().eoc_masked_select(C.byref(cx), C.byref(cm), float(replacement), int(ignore_nodata),
                                          C.c_void_p(_ptr(med)) if med is not None else None, nq, pp,
                                          C.c_void_p(_ptr(qout)) if qout is not None else None,
                                          L.F64 if qdt == np.float64 else L.F32, C.c_void_p(_stream_of(x))))
    if med is not None:
        med = med.reshape(vx.out_shape)
    if qout is not None:
        # [outer, nq, inner]: the quantile axis sits behind the outer dims; move it to the reduced axis' place
        n_outer_dims = vx.axis_pos + tail
        shape = list(vx.out_shape[:n_outer_dims]) + [nq] + list(vx.out_shape[n_outer_dims:])
        qout = qout.reshape(shape)
        if tail:
            qout = qout.movedim(n_outer_dims, vx.axis_pos)
    return med, qout


# ---- zonal statistics (aggregate_spatial) ---------------------------------------------------
def rasterize_polygon(xs, ys, rings, device="cuda"):
    """uint8 CUDA tensor [len(ys), len(xs)]: 1 where the pixel centre (xs[x], ys[y]) is inside the polygon given
    by ``rings`` (a list of (n, 2) float arrays: exterior rings and holes alike) by the even-odd rule
    (``eoc_rasterize_polygon``)."""
    xs = np.ascontiguousarray(np.asarray(xs, dtype=np.float64))
    ys = np.ascontiguousarray(np.asarray(ys, dtype=np.float64))
    rings = [np.asarray(r, dtype=np.float64).reshape(-1, 2) for r in rings]
    offs = np.zeros(len(rings) + 1, dtype=np.int32)
    offs[1:] = np.cumsum([len(r) for r in rings])
    verts = np.ascontiguousarray(np.concatenate(rings, axis=0) if rings else np.zeros((0, 2)))
    dev = torch.device(device)
    d_xs, d_ys = torch.from_numpy(xs).to(dev), torch.from_numpy(ys).to(dev)
    d_v = torch.from_numpy(verts if verts.size else np.zeros((1, 2))).to(dev)
    d_o = torch.from_numpy(offs).to(dev)
    mask = torch.empty((len(ys), len(xs)), dtype=torch.uint8, device=dev)
    if mask.numel() > 0:
        with torch.cuda.device(dev):
            L.check(L.lib().eoc_rasterize_polygon(
                C.c_void_p(d_xs.data_ptr()), C.c_void_p(d_ys.data_ptr()), len(xs), len(ys), C.c_void_p(d_v.data_ptr()),
                C.c_void_p(d_o.data_ptr()), len(rings), C.c_void_p(mask.data_ptr()),
                C.c_void_p(int(torch.cuda.current_stream(dev).cuda_stream))))
    return mask


def zonal_reduce(x, mask, window, ops: Iterable[int], ignore_nodata: bool = True, ddof: int = 1, full_len=None):
    """Streaming reducers of the samples of ``x[..., y0:y1, x0:x1]`` under ``mask`` ([y1-y0, x1-x0] uint8), one
    value per leading index (``eoc_zonal_reduce``).  ``x``: CUDA tensor whose last two dims are (y, x) with unit x
    stride and uniformly strided leading dims (made contiguous otherwise).  ``full_len``: the number of samples the
    reference reduces over per plane (default: the whole (y, x) raster of ``x``)."""
    ops = list(ops)
    x = _native(x)
    lead = tuple(x.shape[:-2])
    Y, X = int(x.shape[-2]), int(x.shape[-1])
    if not x.is_contiguous():
        x = x.contiguous()
    y0, y1, x0, x1 = (int(v) for v in window)
    H, W = y1 - y0, x1 - x0
    planes = int(np.prod(lead)) if lead else 1
    adt = acc_dtype(np_dtype_of(x))
    outs = {op: torch.empty((planes,), dtype=_NP_TO_TORCH[np.dtype(np.int64) if op in L.INT64_OPS else adt], device=x.device)
            for op in ops}
    if planes > 0:
        if H * W == 0:  # nothing under the mask: what the reducers give for an all-NaN slice
            sub = torch.full(lead + (1, 1), float("nan"), dtype=_NP_TO_TORCH[adt], device=x.device)
            return reduce(sub, ops, (-2, -1), ignore_nodata, ddof)
        mask = mask.contiguous()
        lib = L.lib()
        es = x.element_size()
        base = x.data_ptr() + (y0 * X + x0) * es
        with _DeviceGuard(x):
            ws_bytes = int(lib.eoc_zonal_reduce_workspace_bytes(planes, H))
            ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=x.device)
            L.check(lib.eoc_zonal_reduce(
                C.c_void_p(base), L.DTYPE_CODE[np_dtype_of(x)], planes, Y * X, X, H, W, C.c_void_p(mask.data_ptr()), W,
                int(full_len if full_len is not None else Y * X), L.ops_mask(ops), int(ignore_nodata), int(ddof),
                L.outs_array({op: _ptr(t) for op, t in outs.items()}), C.c_void_p(ws.data_ptr()), ws_bytes,
                C.c_void_p(_stream_of(x))))
    return {op: t.reshape(lead) for op, t in outs.items()}


def synth_fill(out, planes: int, plane_stride: int, first_index: int, strip_len: int, seed: int,
               nan_prob: float = 0.0) -> None:
    """Fill a contiguous cuda tensor ``out`` (planes x strip_len) with the deterministic test cube."""
    thr = int(nan_prob * (1 << 24))
    with _DeviceGuard(out):
        L.check(L.lib().eoc_synth_fill(C.c_void_p(out.data_ptr()), L.DTYPE_CODE[np_dtype_of(out)], planes,
                                       plane_stride, first_index, strip_len, seed, thr,
                                       C.c_void_p(_stream_of(out))))


def launch_count(reset: bool = False) -> int:
    return int(L.lib().eoc_launch_count(int(reset)))
