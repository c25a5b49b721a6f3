"""DeviceArray: a CUDA buffer that xarray can hold in ``DataArray.data``.

xarray only keeps duck arrays that implement ``__array_function__`` / ``__array_ufunc__`` (a bare
torch tensor would be coerced through ``np.asarray``), and the reference's reducers are literally
``np.nan*(data, axis=...)`` calls on ``DataArray.data`` (``math.py:118-209``), which is how dask plugs
in.  DeviceArray is that second seam: it routes the numpy functions of the reducer
path to the CUDA library, so even an unmodified ``np.nanmean(device_array, axis=0)`` runs on the GPU.

It is deliberately small: shape/dtype/ndim, basic indexing and transposition (views, via torch),
``np.asarray`` (explicit device->host copy) and the reducer-path functions.  Anything else raises
``TypeError`` instead of silently falling back to the CPU.
"""

from __future__ import annotations

import numpy as np

from . import engine
from . import _lib as L

try:
    import torch
except Exception:  # pragma: no cover
    torch = None

_HANDLED = {}


def _implements(*np_funcs):
    def deco(fn):
        for f in np_funcs:
            _HANDLED[f] = fn
        return fn

    return deco


class DeviceArray:
    __slots__ = ("tensor",)
    __array_priority__ = 100

    def __init__(self, tensor):
        if torch is None or not isinstance(tensor, torch.Tensor) or not tensor.is_cuda:
            raise TypeError("DeviceArray wraps a torch CUDA tensor")
        self.tensor = tensor

    # -- construction ---------------------------------------------------------------------------
    @classmethod
    def from_numpy(cls, array, device="cuda"):
        return cls(torch.from_numpy(np.ascontiguousarray(array)).to(device))

    # -- duck-array protocol --------------------------------------------------------------------
    @property
    def shape(self):
        return tuple(self.tensor.shape)

    @property
    def ndim(self):
        return self.tensor.ndim

    @property
    def dtype(self):
        return engine.np_dtype_of(self.tensor)

    @property
    def size(self):
        return int(self.tensor.numel())

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, key):
        return DeviceArray(self.tensor[key])

    def transpose(self, *axes):
        if len(axes) == 1 and isinstance(axes[0], (tuple, list)):
            axes = tuple(axes[0])
        if not axes:
            axes = tuple(reversed(range(self.ndim)))
        return DeviceArray(self.tensor.permute(*axes))

    @property
    def T(self):
        return self.transpose()

    def astype(self, dtype, copy=True):
        if np.dtype(dtype) == np.float32 and self.dtype in LazyCast.CASTABLE:
            return LazyCast(self.tensor)  # converted in the consumer's load stage (or on first other use)
        return DeviceArray(engine._astype(self.tensor, np.dtype(dtype)))

    def permute(self, *axes):
        return self.transpose(*axes)

    def __array__(self, dtype=None, copy=None):
        host = self.tensor.cpu().numpy()
        return host.astype(dtype) if dtype is not None else host

    def __repr__(self):
        return f"DeviceArray(shape={self.shape}, dtype={self.dtype}, device={self.tensor.device})"

    def __array_function__(self, func, types, args, kwargs):
        if func not in _HANDLED:
            raise TypeError(f"numpy.{func.__name__} is not implemented for DeviceArray (no CPU fallback)")
        return _HANDLED[func](*args, **kwargs)

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        if method != "__call__" or kwargs.get("out") is not None:
            return NotImplemented
        tensors = [x.tensor if isinstance(x, DeviceArray) else x for x in inputs]
        arith = {np.subtract: L.BIN_SUBTRACT, np.add: L.BIN_ADD, np.true_divide: L.BIN_DIVIDE, np.multiply: L.BIN_MULTIPLY}
        if ufunc in arith:  # x OP y in one launch of eoc_binary, numpy's promotion and broadcasting
            return DeviceArray(engine.binary(tensors[0], tensors[1], arith[ufunc]))
        table = {np.isnan: torch.isnan, np.isinf: torch.isinf}
        if ufunc not in table:
            return NotImplemented
        if ufunc in (np.isnan, np.isinf) and not tensors[0].is_floating_point():
            return DeviceArray(torch.zeros_like(tensors[0], dtype=torch.bool))
        return DeviceArray(table[ufunc](*tensors))

    # arithmetic used by normalized_difference(x, y) == (x - y) / (x + y): elementwise plumbing
    def __sub__(self, other):
        return np.subtract(self, other)

    def __add__(self, other):
        return np.add(self, other)

    def __truediv__(self, other):
        return np.true_divide(self, other)

    def __mul__(self, other):
        return np.multiply(self, other)


class LazyCast(DeviceArray):
    """An integer device cube that reads as float32: ``cube.astype("float32")`` of uint8 / int16 / uint16 digital
    numbers without the float32 copy (twice the bytes of a uint16 cube).

    The cast is what a workflow applies before band math on raw Sentinel-2 data (numpy's integer arithmetic would
    wrap: ``uint16(1000) - uint16(2000) == 64536``).  Consumers that can convert in their load stage take
    ``source``: ``ndvi`` -> ``LazyNormalizedDifference`` -> ``eoc_nd_reduce`` reads the two uint16 band planes and
    computes in float32.  Everything else sees ``.tensor``, the float32 copy, made on first use and kept.
    Indexing and transposition stay lazy (they are views of ``source``).
    """

    __slots__ = ("source", "_dense")
    CASTABLE = (np.dtype(np.uint8), np.dtype(np.int16), np.dtype(np.uint16))

    def __init__(self, source):
        if isinstance(source, DeviceArray):
            source = source.tensor
        if torch is None or not isinstance(source, torch.Tensor) or not source.is_cuda:
            raise TypeError("LazyCast wraps a torch CUDA tensor")
        if engine.np_dtype_of(source) not in self.CASTABLE:
            raise TypeError("LazyCast: uint8 / int16 / uint16 sources only (exact in float32)")
        self.source = source
        self._dense = None

    @property
    def tensor(self):
        if self._dense is None:
            self._dense = self.source.to(torch.float32)
        return self._dense

    @property
    def shape(self):
        return tuple(self.source.shape)

    @property
    def ndim(self):
        return self.source.ndim

    @property
    def dtype(self):
        return np.dtype(np.float32)

    @property
    def size(self):
        return int(self.source.numel())

    def __getitem__(self, key):
        return LazyCast(self.source[key])

    def transpose(self, *axes):
        if len(axes) == 1 and isinstance(axes[0], (tuple, list)):
            axes = tuple(axes[0])
        if not axes:
            axes = tuple(reversed(range(self.ndim)))
        return LazyCast(self.source.permute(*axes))

    def permute(self, *axes):
        return self.transpose(*axes)

    def astype(self, dtype, copy=True):
        if np.dtype(dtype) == np.float32:
            return self
        return DeviceArray(engine._astype(self.tensor, np.dtype(dtype)))

    def __repr__(self):
        return f"LazyCast(shape={self.shape}, {engine.np_dtype_of(self.source)} -> float32, device={self.source.device})"


def _wrap(x):
    return DeviceArray(x) if engine.is_torch(x) else x


def _reducer(op, nan_aware):
    def impl(a, axis=None, keepdims=False, ddof=0, **unsupported):
        _reject(unsupported)
        res = engine.reduce(a.tensor, [op], axis, ignore_nodata=nan_aware, ddof=ddof)[op]
        return _wrap(_keep(res, a, axis, keepdims))

    return impl


def _reject(kwargs):
    bad = {k: v for k, v in kwargs.items() if v is not None and v is not np._NoValue}
    if bad:
        raise TypeError(f"unsupported arguments for a DeviceArray reduction: {sorted(bad)}")


def _keep(res, a, axis, keepdims):
    if keepdims is True:
        from .views import normalize_axes

        shape = list(a.shape)
        for ax in normalize_axes(axis, a.ndim):
            shape[ax] = 1
        return res.reshape(shape)
    return res


for _np_nan, _np_plain, _op in (
    (np.nansum, np.sum, L.OP_SUM), (np.nanmean, np.mean, L.OP_MEAN), (np.nanmin, np.min, L.OP_MIN),
    (np.nanmax, np.max, L.OP_MAX), (np.nanvar, np.var, L.OP_VARIANCE), (np.nanstd, np.std, L.OP_SD),
    (np.nanprod, np.prod, L.OP_PRODUCT),
):
    _implements(_np_nan)(_reducer(_op, True))
    _implements(_np_plain)(_reducer(_op, False))


@_implements(np.nanmedian)
def _nanmedian(a, axis=None, keepdims=False, **unsupported):
    _reject(unsupported)
    return _wrap(_keep(engine.select(a.tensor, axis, True, True)[0], a, axis, keepdims))


@_implements(np.median)
def _median(a, axis=None, keepdims=False, **unsupported):
    _reject(unsupported)
    return _wrap(_keep(engine.select(a.tensor, axis, False, True)[0], a, axis, keepdims))


def _quantile(a, q, axis, nan_aware, method, keepdims):
    if method != "linear":
        raise TypeError("only method='linear' is implemented for DeviceArray quantiles")
    q = np.asarray(q, dtype=np.float64)
    res = engine.select(a.tensor, axis, nan_aware, False, q.ravel())[1]
    # numpy puts the q axis first (and drops it for scalar q)
    from .views import normalize_axes

    pos = normalize_axes(axis, a.ndim)[0]
    res = res.movedim(pos, 0)
    if q.ndim == 0:
        res = res[0]
    if keepdims is True:
        raise TypeError("keepdims is not implemented for DeviceArray quantiles")
    return _wrap(res)


@_implements(np.nanquantile)
def _nanquantile(a, q, axis=None, method="linear", keepdims=False, **unsupported):
    _reject(unsupported)
    return _quantile(a, q, axis, True, method, keepdims)


@_implements(np.quantile)
def _plainquantile(a, q, axis=None, method="linear", keepdims=False, **unsupported):
    _reject(unsupported)
    return _quantile(a, q, axis, False, method, keepdims)


@_implements(np.moveaxis)
def _moveaxis(a, source, destination):
    return DeviceArray(a.tensor.movedim(source, destination))


@_implements(np.transpose)
def _transpose(a, axes=None):
    return a.transpose(*(axes or ()))


@_implements(np.take)
def _take(a, indices, axis=None, **unsupported):
    _reject(unsupported)
    if axis is None:
        raise TypeError("np.take without axis is not implemented for DeviceArray")
    if np.ndim(indices) == 0:
        return DeviceArray(a.tensor.select(axis, int(indices)))
    idx = torch.as_tensor(np.asarray(indices), device=a.tensor.device)
    return DeviceArray(a.tensor.index_select(axis, idx))


@_implements(np.shape)
def _shape(a):
    return a.shape


@_implements(np.ndim)
def _ndim(a):
    return a.ndim


class LazyNormalizedDifference(DeviceArray):
    """(a - b) / (a + b) of two same-shape device arrays, not yet evaluated.  A ``DeviceArray`` (so the result of
    ``ndvi`` on a DeviceArray-backed cube is again an instance of the input's array type, as the reference's
    ``general_output_checks`` asserts) whose ``.tensor`` is computed on first use.

    Produced by ``ndvi``; the streaming reducers of ``process_implementations.math`` recognise it and run the
    fused kernel ``eoc_nd_reduce`` (two band planes in, one image out, the NDVI cube never exists in memory).
    ``materialize()`` / ``np.asarray`` evaluate it elementwise with the same kernel.

    The arithmetic is the reference's (``math.py:412-414`` on what ``DataArray.data`` holds):
    * float operands: float arithmetic in their dtype;
    * raw integer operands: numpy's -- ``a - b`` and ``a + b`` WRAP in the operands' dtype, the true division
      gives float64 (``integer_wrap``); uint16(1000), uint16(2000) -> 21.512;
    * ``LazyCast`` operands (``cube.astype("float32")`` of uint8 / int16 / uint16): float32 arithmetic on the
      exactly converted samples -- the integer planes are read as they are and converted in the load stage.
    Fused and materialised evaluation give the same bits in each of the three cases.
    """

    __slots__ = ("a", "b", "wrapped", "integer_wrap", "_dense")

    def __init__(self, a, b):
        self._dense = None
        self.wrapped = isinstance(a, DeviceArray)
        cast = isinstance(a, LazyCast) and isinstance(b, LazyCast)
        if cast:
            self.a, self.b = a.source, b.source
        else:
            self.a = a.tensor if isinstance(a, DeviceArray) else a
            self.b = b.tensor if isinstance(b, DeviceArray) else b
        self.integer_wrap = (not cast) and engine.np_dtype_of(self.a).kind in "iu"
        if tuple(self.a.shape) != tuple(self.b.shape):
            raise ValueError("normalized difference operands differ in shape")

    @staticmethod
    def supports(a, b) -> bool:
        if isinstance(a, LazyCast) != isinstance(b, LazyCast):
            return False
        ra = a.source if isinstance(a, LazyCast) else (a.tensor if isinstance(a, DeviceArray) else a)
        rb = b.source if isinstance(b, LazyCast) else (b.tensor if isinstance(b, DeviceArray) else b)
        return (engine.is_torch(ra) and engine.is_torch(rb) and ra.is_cuda and rb.is_cuda
                and ra.dtype == rb.dtype and tuple(ra.shape) == tuple(rb.shape)
                and engine.np_dtype_of(ra) in L.DTYPE_CODE)

    @property
    def shape(self):
        return tuple(self.a.shape)

    @property
    def ndim(self):
        return self.a.ndim

    def __len__(self):
        return self.shape[0] if self.ndim else 0

    @property
    def dtype(self):
        return engine.nd_acc_dtype(engine.np_dtype_of(self.a), self.integer_wrap)

    def _like(self, a, b):
        out = object.__new__(LazyNormalizedDifference)
        out.a, out.b, out.wrapped, out.integer_wrap, out._dense = a, b, self.wrapped, self.integer_wrap, None
        return out

    @property
    def tensor(self):
        if self._dense is None:
            _, self._dense = engine.nd_reduce(self.a, self.b, None, [], want_nd=True, integer_wrap=self.integer_wrap)
        return self._dense

    @property
    def size(self):
        return int(self.a.numel())

    def astype(self, dtype, copy=True):
        return DeviceArray(engine._astype(self.tensor, np.dtype(dtype)))

    def __getitem__(self, key):
        return self._like(self.a[key], self.b[key])

    def permute(self, *axes):
        return self._like(self.a.permute(*axes), self.b.permute(*axes))

    transpose = permute

    def materialize(self):
        return DeviceArray(self.tensor) if self.wrapped else self.tensor

    def reduce(self, op, axis, ignore_nodata=True, ddof=1):
        res, _ = engine.nd_reduce(self.a, self.b, axis, [op], ignore_nodata=ignore_nodata, ddof=ddof,
                                  integer_wrap=self.integer_wrap)
        return DeviceArray(res[op]) if self.wrapped else res[op]

    def __array__(self, dtype=None, copy=None):
        dense = self.materialize()
        dense = dense.cpu().numpy() if engine.is_torch(dense) else np.asarray(dense)
        return dense if dtype is None else dense.astype(dtype)

    def __array_function__(self, func, types, args, kwargs):
        args = tuple(x.materialize() if isinstance(x, LazyNormalizedDifference) else x for x in args)
        return func(*args, **kwargs)

    def __repr__(self):
        return f"LazyNormalizedDifference(shape={self.shape}, dtype={self.dtype})"


class LazyMasked(DeviceArray):
    """``where(mask, replacement, data)`` of a device cube and a canonical uint8 mask, not yet evaluated (a
    ``DeviceArray`` whose ``.tensor`` is computed on first use, like ``LazyNormalizedDifference``).

    Produced by ``mask(..., fuse=True)`` (reference ``cubes/mask.py:112``: ``data.where(_not(mask),
    replacement)``).  The reducers of ``process_implementations.math`` recognise it: streaming reducers run
    ``eoc_masked_reduce`` (mask byte consumed in the load stage), ``median`` / ``quantiles`` run
    ``eoc_masked_select`` (mask staged next to the TMA tile) -- the masked cube never exists in memory.
    ``materialize()`` / ``np.asarray`` evaluate it elementwise with the same kernel.

    ``mask`` is kept as an ``expand`` view (stride 0 over the dims it lacks): a (t,y,x) cloud mask is read as
    is for every band of a (t,bands,y,x) cube.
    """

    __slots__ = ("x", "m", "wrapped", "replacement", "_dense")

    def __init__(self, data, mask_u8, replacement=float("nan")):
        self._dense = None
        self.wrapped = isinstance(data, DeviceArray)
        self.x = data.tensor if self.wrapped else data
        self.m = mask_u8
        self.replacement = float(replacement)

    @property
    def shape(self):
        return tuple(self.x.shape)

    @property
    def ndim(self):
        return self.x.ndim

    def __len__(self):
        return self.shape[0] if self.ndim else 0

    @property
    def dtype(self):
        return self.result_dtype(engine.np_dtype_of(self.x), self.replacement)

    @staticmethod
    def result_dtype(in_dt, replacement):
        """dtype of xarray's ``where(cond, data, other)``: floats keep theirs; integers stay integer only
        for an integral replacement, NaN (the default) promotes them to float64."""
        if in_dt.kind == "f":
            return in_dt
        integral = replacement == replacement and float(replacement).is_integer()
        return in_dt if integral else np.dtype(np.float64)

    def _wrap(self, t):
        return DeviceArray(t) if self.wrapped else t

    def _like(self, x, m):
        out = LazyMasked(x, m, self.replacement)
        out.wrapped = self.wrapped
        return out

    def __getitem__(self, key):
        m = self.m.expand(tuple(self.x.shape)) if tuple(self.m.shape) != tuple(self.x.shape) else self.m
        return self._like(self.x[key], m[key])

    def permute(self, *axes):
        m = self.m.expand(tuple(self.x.shape)) if tuple(self.m.shape) != tuple(self.x.shape) else self.m
        return self._like(self.x.permute(*axes), m.permute(*axes))

    transpose = permute

    @property
    def tensor(self):
        if self._dense is None:
            _, dense = engine.masked_reduce(self.x, self.m, self.replacement, [], None, want_masked=True)
            self._dense = engine._astype(dense, self.dtype)
        return self._dense

    @property
    def size(self):
        return int(self.x.numel())

    def astype(self, dtype, copy=True):
        return DeviceArray(engine._astype(self.tensor, np.dtype(dtype)))

    def materialize(self):
        return self._wrap(self.tensor)

    def reduce(self, op, axis, ignore_nodata=True, ddof=1):
        res, _ = engine.masked_reduce(self.x, self.m, self.replacement, [op], axis, ignore_nodata, ddof)
        return self._wrap(res[op])

    def select(self, axis, ignore_nodata=True, want_median=True, probs=None):
        """(median or None, quantiles or None): one fused pass when the kernel covers the case, else the masked
        cube is materialised first (two more passes over the data, same result)."""
        try:
            med, q = engine.masked_select(self.x, self.m, axis, ignore_nodata, want_median, probs,
                                           replacement=self.replacement)
        except L.EocError as e:
            if e.code != -2:
                raise
            dense = self.materialize()
            dense = dense.tensor if isinstance(dense, DeviceArray) else dense
            med, q = engine.select(dense, axis, ignore_nodata, want_median, probs)
        return (self._wrap(med) if med is not None else None), (self._wrap(q) if q is not None else None)

    def __array__(self, dtype=None, copy=None):
        dense = self.materialize()
        dense = dense.cpu().numpy() if engine.is_torch(dense) else np.asarray(dense)
        return dense if dtype is None else dense.astype(dtype)

    def __array_function__(self, func, types, args, kwargs):
        args = tuple(x.materialize() if isinstance(x, LazyMasked) else x for x in args)
        return func(*args, **kwargs)

    def __repr__(self):
        return f"LazyMasked(shape={self.shape}, dtype={self.dtype}, replacement={self.replacement})"
