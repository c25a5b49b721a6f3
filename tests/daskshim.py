"""A ``dask.array``-shaped stand-in for the tests (dask is not installable in the build image).

It implements the part of ``dask.array`` the host layer touches -- and nothing else -- with the same names and
argument meaning, so every ``_is_dask`` branch of ``process_implementations`` can be driven on CPU and GPU:

    Array.shape / ndim / dtype / chunks / numblocks, len()
    Array.rechunk({axis: -1} | shape)       merge the chunks of an axis (or of all axes)
    Array.map_blocks(fn, drop_axis=, new_axis=, chunks=, dtype=)
    Array.flatten(), Array.compute()
    dask.array.from_array(x, chunks=), map_blocks(fn, *arrays, dtype=), expand_dims, stack, take

An ``Array`` is LAZY like the real one: a chunk grid plus a function that produces one block on demand; nothing
runs before ``compute()``.  ``Array.calls`` counts block evaluations so tests can assert laziness, and
``BLOCK_LOG`` records the shape of every block handed to a mapped function (one entry per kernel-side call).

``install()`` registers the modules ``dask``, ``dask.array`` and ``dask.array.core`` in ``sys.modules`` unless a
real dask is importable; ``Array.__module__`` is ``dask.array.core`` like the real class.
"""

from __future__ import annotations

import itertools
import sys
import types

import numpy as np

BLOCK_LOG = []


def _normalize_chunks(chunks, shape):
    if isinstance(chunks, int):
        chunks = (chunks,) * len(shape)
    out = []
    for c, n in zip(chunks, shape):
        if isinstance(c, int):
            c = n if c == -1 else c
            full, rest = divmod(n, c) if c else (0, 0)
            c = (c,) * full + ((rest,) if rest else ())
            if n == 0:
                c = (0,)
        out.append(tuple(int(v) for v in c))
    return tuple(out)


class Array:
    """Lazy chunked array: ``block(index_tuple) -> ndarray`` over the grid described by ``chunks``."""

    calls = 0  # block evaluations, all arrays

    def __init__(self, block, chunks, dtype):
        self._block = block
        self.chunks = tuple(tuple(c) for c in chunks)
        self.dtype = np.dtype(dtype)

    @property
    def shape(self):
        return tuple(sum(c) for c in self.chunks)

    @property
    def ndim(self):
        return len(self.chunks)

    @property
    def numblocks(self):
        return tuple(len(c) for c in self.chunks)

    @property
    def size(self):
        return int(np.prod(self.shape))

    def __len__(self):
        return self.shape[0]

    def __repr__(self):
        return f"dask.array<shim, shape={self.shape}, dtype={self.dtype}, chunks={self.chunks}>"

    def __array__(self, dtype=None, copy=None):  # like dask: np.asarray computes
        out = self.compute()
        return out if dtype is None else out.astype(dtype)

    # -- evaluation --------------------------------------------------------------------------------------
    def _get(self, idx):
        Array.calls += 1
        blk = self._block(idx)
        want = tuple(self.chunks[d][i] for d, i in enumerate(idx))
        got = tuple(np.shape(blk))
        if got != want:
            raise ValueError(f"block {idx} has shape {got}, the chunk grid says {want}")
        return np.asarray(blk)

    def compute(self):
        if self.ndim == 0:
            return np.asarray(self._get(()))
        out = np.empty(self.shape, dtype=self.dtype)
        starts = [np.concatenate([[0], np.cumsum(c)]) for c in self.chunks]
        for idx in itertools.product(*[range(n) for n in self.numblocks]):
            sl = tuple(slice(int(starts[d][i]), int(starts[d][i + 1])) for d, i in enumerate(idx))
            out[sl] = self._get(idx)
        return out

    # -- structure ---------------------------------------------------------------------------------------
    def rechunk(self, chunks):
        if isinstance(chunks, dict):
            new = list(self.chunks)
            for ax, c in chunks.items():
                ax = ax % self.ndim
                new[ax] = _normalize_chunks((c,), (self.shape[ax],))[0]
        else:
            new = _normalize_chunks(tuple(chunks), self.shape)
        new = tuple(new)
        if new == self.chunks:
            return self
        old = self
        old_starts = [np.concatenate([[0], np.cumsum(c)]) for c in old.chunks]
        new_starts = [np.concatenate([[0], np.cumsum(c)]) for c in new]

        def block(idx):
            lo = [int(new_starts[d][i]) for d, i in enumerate(idx)]
            hi = [int(new_starts[d][i + 1]) for d, i in enumerate(idx)]
            out = np.empty([h - l for l, h in zip(lo, hi)], dtype=old.dtype)
            ranges = []
            for d in range(old.ndim):
                ranges.append([j for j in range(len(old.chunks[d]))
                               if old_starts[d][j] < hi[d] and old_starts[d][j + 1] > lo[d]] or [0])
            for oidx in itertools.product(*ranges):
                src = old._get(oidx)
                ssl, dsl = [], []
                for d, j in enumerate(oidx):
                    a, b = int(old_starts[d][j]), int(old_starts[d][j + 1])
                    s, e = max(a, lo[d]), min(b, hi[d])
                    ssl.append(slice(s - a, e - a))
                    dsl.append(slice(s - lo[d], e - lo[d]))
                out[tuple(dsl)] = src[tuple(ssl)]
            return out

        return Array(block, new, self.dtype)

    def map_blocks(self, fn, *args, drop_axis=(), new_axis=(), chunks=None, dtype=None, **kwargs):
        return map_blocks(fn, self, *args, drop_axis=drop_axis, new_axis=new_axis, chunks=chunks, dtype=dtype,
                          **kwargs)

    def flatten(self):
        src = self

        def block(idx):
            return src.compute().reshape(-1)

        return Array(block, ((self.size,),), self.dtype)

    def __getitem__(self, key):
        src = self
        probe = np.empty(self.shape, dtype=np.bool_)[key]

        def block(idx):
            return src.compute()[key]

        return Array(block, tuple((n,) for n in probe.shape), self.dtype)


def from_array(x, chunks):
    x = np.asarray(x)
    chunks = _normalize_chunks(chunks, x.shape)
    starts = [np.concatenate([[0], np.cumsum(c)]) for c in chunks]

    def block(idx):
        return x[tuple(slice(int(starts[d][i]), int(starts[d][i + 1])) for d, i in enumerate(idx))]

    return Array(block, chunks, x.dtype)


def map_blocks(fn, *args, drop_axis=(), new_axis=(), chunks=None, dtype=None, **kwargs):
    arrays = [a for a in args if isinstance(a, Array)]
    lead = arrays[0]
    for a in arrays[1:]:
        if a.chunks != lead.chunks:
            raise NotImplementedError("shim: mapped arrays must share one chunk grid")
    drop = sorted({d % lead.ndim for d in ([drop_axis] if isinstance(drop_axis, int) else list(drop_axis or ()))})
    new = sorted([new_axis] if isinstance(new_axis, int) else list(new_axis or ()))
    for d in drop:
        if len(lead.chunks[d]) != 1:
            raise ValueError("map_blocks(drop_axis=...): the dropped axis must be in one chunk")
    kept = [c for d, c in enumerate(lead.chunks) if d not in drop]
    if chunks is not None:
        out_chunks = tuple(tuple(c) for c in chunks)
    else:
        out_chunks = list(kept)
        for ax in new:
            out_chunks.insert(ax, (1,))
        out_chunks = tuple(out_chunks)

    def block(idx):
        idx = list(idx)
        for ax in reversed(new):
            idx.pop(ax)
        for d in drop:
            idx.insert(d, 0)
        ins = [a._get(tuple(idx)) if isinstance(a, Array) else a for a in args]
        BLOCK_LOG.append(tuple(np.shape(ins[0])))
        return np.asarray(fn(*ins, **kwargs))

    return Array(block, out_chunks, dtype if dtype is not None else lead.dtype)


def expand_dims(a, axis):
    axes = sorted([axis] if isinstance(axis, int) else list(axis))
    chunks = list(a.chunks)
    for ax in axes:
        chunks.insert(ax, (1,))

    def block(idx):
        inner = tuple(i for d, i in enumerate(idx) if d not in axes)
        return np.expand_dims(a._get(inner), tuple(axes))

    return Array(block, tuple(chunks), a.dtype)


def stack(arrays, axis=0):
    arrays = list(arrays)
    lead = arrays[0]
    chunks = list(lead.chunks)
    chunks.insert(axis, (1,) * len(arrays))

    def block(idx):
        idx = list(idx)
        k = idx.pop(axis)
        return np.expand_dims(arrays[k]._get(tuple(idx)), axis)

    return Array(block, tuple(chunks), np.result_type(*[a.dtype for a in arrays]))


def take(a, index, axis=0):
    ax = axis % a.ndim
    src = a.rechunk({ax: -1})
    return map_blocks(lambda b: np.take(b, index, axis=ax), src, drop_axis=[ax], dtype=a.dtype)


def install():
    """Register the shim as ``dask`` unless the real package is importable.  Returns the ``dask.array`` module."""
    try:
        import dask.array as real  # noqa: F401

        if not getattr(real, "__eoc_shim__", False):
            return real
        return real
    except Exception:
        pass
    dask = types.ModuleType("dask")
    da = types.ModuleType("dask.array")
    core = types.ModuleType("dask.array.core")
    Array.__module__ = "dask.array.core"
    for mod in (da, core):
        mod.Array = Array
    da.from_array, da.map_blocks, da.expand_dims, da.stack, da.take = from_array, map_blocks, expand_dims, stack, take
    da.core = core
    da.__eoc_shim__ = True
    dask.array = da
    sys.modules.update({"dask": dask, "dask.array": da, "dask.array.core": core})
    return da
