"""Builds and loads the host model of the kernels' column arithmetic (test infrastructure)."""

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(ROOT, "build", "hostmodel", "libeoc_hostmodel.so")
SRC = os.path.join(HERE, "hostmodel.cpp")
HDR = os.path.join(ROOT, "openeo_processes_dask_b200", "csrc", "column_math.cuh")
INC = os.path.join(ROOT, "openeo_processes_dask_b200", "csrc", "generated", "sort_networks.inc")


def load():
    newest = max(os.path.getmtime(p) for p in (SRC, HDR, INC))
    if not os.path.exists(SO) or os.path.getmtime(SO) < newest:
        os.makedirs(os.path.dirname(SO), exist_ok=True)
        subprocess.run(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
                        "-x", "c++", SRC, "-o", SO], check=True)
    return C.CDLL(SO)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def reduce_columns(lib, x, ignore=True, ddof=1):
    """x: (R, inner) float32/float64 -> dict of per-column results."""
    x = np.ascontiguousarray(x)
    R, inner = x.shape
    fn = lib.hm_reduce_f32 if x.dtype == np.float32 else lib.hm_reduce_f64
    out = {k: np.empty(inner, dtype=x.dtype) for k in ("sum", "mean", "min", "max", "product", "variance", "sd")}
    cnt = np.empty(inner, dtype=np.int64)
    fn(_p(x), R, inner, int(ignore), ddof, *[_p(out[k]) for k in ("sum", "mean", "min", "max", "product", "variance", "sd")], _p(cnt))
    out["count"] = cnt
    return out


def order_columns(lib, x, ignore=True):
    """x: (R, inner) -> first / last / any / all per column (any / all as 0 / 1 / NaN in x's dtype)."""
    x = np.ascontiguousarray(x)
    R, inner = x.shape
    fn = lib.hm_order_f32 if x.dtype == np.float32 else lib.hm_order_f64
    out = {k: np.empty(inner, dtype=x.dtype) for k in ("first", "last", "any", "all")}
    fn(_p(x), R, inner, int(ignore), *[_p(out[k]) for k in ("first", "last", "any", "all")])
    return out


def median_columns(lib, x, N, ignore=True):
    x = np.ascontiguousarray(x)
    R, inner = x.shape
    fn = lib.hm_median_f32 if x.dtype == np.float32 else lib.hm_median_f64
    out = np.empty(inner, dtype=x.dtype)
    assert fn(_p(x), R, inner, N, int(ignore), _p(out)) == 0
    return out


def quantile_columns(lib, x, N, probs, ignore=True):
    x = np.ascontiguousarray(x)
    R, inner = x.shape
    probs = np.ascontiguousarray(probs, dtype=np.float64)
    fn = lib.hm_quantiles_f32 if x.dtype == np.float32 else lib.hm_quantiles_f64
    q = np.empty((len(probs), inner), dtype=np.float64)
    med = np.empty(inner, dtype=x.dtype)
    assert fn(_p(x), R, inner, N, int(ignore), len(probs), _p(probs), _p(q), _p(med)) == 0
    return q, med


def split_columns(lib, x, NH, probs, ignore=True):
    """The split selection (two runs of NH + rank search): (quantiles [nq, inner] float64, median [inner])."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    R, inner = x.shape
    probs = np.ascontiguousarray(probs, dtype=np.float64)
    q = np.empty((len(probs), inner), dtype=np.float64)
    med = np.empty(inner, dtype=np.float32)
    assert lib.hm_split_f32(_p(x), R, inner, NH, int(ignore), len(probs), _p(probs), _p(q), _p(med)) == 0
    return q, med


def quad_columns(lib, x, NH, ignore=True):
    """k_select_quad's column math (four raw-sorted runs of NH with NaN sentinels + rank elimination): median [inner]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    R, inner = x.shape
    med = np.empty(inner, dtype=np.float32)
    assert lib.hm_quad_f32(_p(x), R, inner, NH, int(ignore), _p(med)) == 0
    return med


def multi_columns(lib, x, NH, ignore=True):
    """k_select_multi's column math (eight raw-sorted runs of NH with NaN sentinels + rank elimination): median [inner]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    R, inner = x.shape
    med = np.empty(inner, dtype=np.float32)
    assert lib.hm_multi_f32(_p(x), R, inner, NH, int(ignore), _p(med)) == 0
    return med


def kruns_columns(lib, x, NH, K, probs, ignore=True):
    """K sorted runs of NH + rank elimination (long columns): (quantiles [nq, inner] float64, median [inner])."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    R, inner = x.shape
    probs = np.ascontiguousarray(probs, dtype=np.float64)
    q = np.empty((len(probs), inner), dtype=np.float64)
    med = np.empty(inner, dtype=np.float32)
    assert lib.hm_kruns_f32(_p(x), R, inner, NH, K, int(ignore), len(probs), _p(probs), _p(q), _p(med)) == 0
    return q, med
