"""eoc_reduce_rows_static: the long-row (spatial) reduction with its scratch in a static __device__ array.

The entry point has no counterpart in the reference (added for callers that
cannot call cudaMalloc).  Called here through the C ABI (ctypes) and compared with the existing entry point
``eoc_reduce`` given a caller-allocated workspace -- same kernels, so the results must be bit-identical (NaN masks
included; the tolerance of the spatial tests, rtol 1e-6, is checked against the oracle besides) -- on two calls in a
row that share the array: the second call is a different, smaller problem, so stale partials of the first call would
show.  Collected last on purpose (``zz``).
"""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call_both(L, torch, x, ops, ignore_nodata=1, ddof=1):
    """x: CUDA tensor (rows, n) contiguous -> ({op: static result}, {op: eoc_reduce result})"""
    lib = L.lib()
    rows, n = x.shape
    dt = L.DTYPE_CODE[np.dtype(str(x.dtype).replace("torch.", ""))]
    view = L.make_view(int(x.data_ptr()), dt, rows, n, 1, n, 1)
    acc = torch.float32 if x.dtype == torch.float32 else torch.float64
    mask = L.ops_mask(ops)

    def outs():
        return {op: torch.full((rows,), -12345, dtype=torch.int64 if op in L.INT64_OPS else acc, device=x.device)
                for op in ops}

    stream = C.c_void_p(int(torch.cuda.current_stream(x.device).cuda_stream))
    a = outs()
    L.check(lib.eoc_reduce_rows_static(C.byref(view), mask, ignore_nodata, ddof,
                                       L.outs_array({op: int(t.data_ptr()) for op, t in a.items()}), stream))
    ws_bytes = lib.eoc_reduce_workspace_bytes(C.byref(view), mask)
    assert 0 < ws_bytes <= L.ROWS_STATIC_BYTES
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    b = outs()
    L.check(lib.eoc_reduce(C.byref(view), mask, ignore_nodata, ddof,
                           L.outs_array({op: int(t.data_ptr()) for op, t in b.items()}),
                           C.c_void_p(ws.data_ptr()), ws_bytes, stream))
    torch.cuda.synchronize()
    return {op: t.cpu().numpy() for op, t in a.items()}, {op: t.cpu().numpy() for op, t in b.items()}


def test_rows_static_matches_eoc_reduce_on_two_calls_in_a_row():
    import torch

    from openeo_processes_dask_b200 import _lib as L

    ops = [L.OP_SUM, L.OP_MEAN, L.OP_MIN, L.OP_MAX, L.OP_COUNT, L.OP_VARIANCE, L.OP_SD, L.OP_NOT_NAN]
    rng = np.random.default_rng(2026)
    cases = [((24, 300_000), np.float32, 1), ((5, 70_001), np.float32, 0), ((7, 40_960), np.float64, 1)]
    for shape, dtype, ignore in cases:  # calls in a row on the same device share the scratch array
        x = rng.normal(3000, 1000, size=shape).astype(dtype)
        x[rng.random(shape) < 0.2] = np.nan
        x[0, :] = np.nan                      # an all-NaN row
        if ignore == 0:
            x[1, :] = np.abs(x[1, :])
            x[1, np.isnan(x[1, :])] = 1.0     # a row without NaN survives ignore_nodata = 0
        d = torch.from_numpy(x).cuda()
        got, want = _call_both(L, torch, d, ops, ignore_nodata=ignore)
        for op in ops:
            np.testing.assert_array_equal(got[op], want[op], err_msg=f"op {op} shape {shape}")
        with np.errstate(all="ignore"):
            ref_mean = np.nanmean(x.astype(np.float64), axis=1) if ignore else np.mean(x.astype(np.float64), axis=1)
            ref_sd = (np.nanstd(x.astype(np.float64), axis=1, ddof=1) if ignore
                      else np.std(x.astype(np.float64), axis=1, ddof=1))
        np.testing.assert_allclose(got[L.OP_MEAN], ref_mean, rtol=1e-6, equal_nan=True)
        np.testing.assert_allclose(got[L.OP_SD], ref_sd, rtol=1e-5, equal_nan=True)
        assert (got[L.OP_NOT_NAN] == np.sum(~np.isnan(x), axis=1)).all()


def test_rows_static_rejects_what_does_not_fit():
    import torch

    from openeo_processes_dask_b200 import _lib as L

    lib = L.lib()
    rows, n = 9000, 4096                      # 9000 rows x 1 chunk x 64 B = 576000 B > 512 KiB
    x = torch.zeros((rows, n), dtype=torch.float32, device="cuda")
    out = torch.empty(rows, dtype=torch.float32, device="cuda")
    view = L.make_view(int(x.data_ptr()), L.F32, rows, n, 1, n, 1)
    rc = lib.eoc_reduce_rows_static(C.byref(view), L.ops_mask([L.OP_SUM]), 1, 1,
                                    L.outs_array({L.OP_SUM: int(out.data_ptr())}), None)
    assert rc == -2 and b"EOC_ROWS_STATIC_BYTES" in lib.eoc_last_error()
    view = L.make_view(int(x.data_ptr()), L.F32, 1, rows, n, rows * n, n)   # a column view: no scratch needed
    rc = lib.eoc_reduce_rows_static(C.byref(view), L.ops_mask([L.OP_SUM]), 1, 1,
                                    L.outs_array({L.OP_SUM: int(out.data_ptr())}), None)
    assert rc == -2 and b"eoc_reduce" in lib.eoc_last_error()
    torch.cuda.synchronize()
