"""Deterministic synthetic cubes (TEST INFRASTRUCTURE; numpy twin of csrc/synth.cu).

Element value is a pure function of (seed, linear index in the logical C-order
of the FULL cube), so any y-strip on any GPU regenerates the same numbers
(SURVEY.md section 8d):

    h   = splitmix64(seed * 0x9E3779B97F4A7C15 + idx)
    NaN iff (h & 0xFFFFFF) < nan_prob * 2**24          (float dtypes only)
    val = (h >> 40) % 10000                            (Sentinel-2-like DN)
"""

from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
GOLDEN = 0x9E3779B97F4A7C15


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + np.uint64(GOLDEN)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def nan_threshold(nan_prob: float) -> int:
    """Integer threshold used by both twins: NaN iff (h & 0xFFFFFF) < threshold."""
    return int(nan_prob * (1 << 24))


def synth_values(seed: int, first_index: int, n: int, dtype=np.float32, nan_prob: float = 0.0):
    """n consecutive elements starting at logical linear index ``first_index``."""
    idx = np.arange(first_index, first_index + n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        base = np.uint64((seed * GOLDEN) & 0xFFFFFFFFFFFFFFFF)
        h = splitmix64(base + idx)
    val = ((h >> np.uint64(40)) % np.uint64(10000)).astype(np.int64)
    dtype = np.dtype(dtype)
    out = val.astype(dtype)
    if dtype.kind == "f" and nan_prob > 0.0:
        thr = np.uint64(nan_threshold(nan_prob))
        out[(h & np.uint64(0xFFFFFF)) < thr] = np.nan
    return out


def synth_cube(seed: int, shape, dtype=np.float32, nan_prob: float = 0.0):
    """Whole cube of ``shape`` in C order."""
    n = int(np.prod(shape))
    return synth_values(seed, 0, n, dtype, nan_prob).reshape(shape)


def synth_window(seed, full_shape, window, dtype=np.float32, nan_prob=0.0):
    """A hyper-rectangular window (tuple of slices) of the full cube without
    materialising the cube: index arithmetic only."""
    full_shape = tuple(int(s) for s in full_shape)
    ranges = []
    for sl, dim in zip(window, full_shape):
        start, stop, step = sl.indices(dim)
        assert step == 1
        ranges.append(np.arange(start, stop, dtype=np.uint64))
    strides = np.ones(len(full_shape), dtype=np.uint64)
    for i in range(len(full_shape) - 2, -1, -1):
        strides[i] = strides[i + 1] * np.uint64(full_shape[i + 1])
    idx = np.zeros([len(r) for r in ranges], dtype=np.uint64)
    for ax, r in enumerate(ranges):
        shp = [1] * len(ranges)
        shp[ax] = len(r)
        idx = idx + (r * strides[ax]).reshape(shp)
    flat = idx.ravel()
    with np.errstate(over="ignore"):
        base = np.uint64((seed * GOLDEN) & 0xFFFFFFFFFFFFFFFF)
        h = splitmix64(base + flat)
    val = ((h >> np.uint64(40)) % np.uint64(10000)).astype(np.int64)
    dtype = np.dtype(dtype)
    out = val.astype(dtype)
    if dtype.kind == "f" and nan_prob > 0.0:
        out[(h & np.uint64(0xFFFFFF)) < np.uint64(nan_threshold(nan_prob))] = np.nan
    return out.reshape(idx.shape)
