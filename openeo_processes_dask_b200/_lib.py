"""ctypes binding of libeoc_sm100.so (declarations mirror include/eoc.h one to one).

There is no CPU fallback: if the library is missing, ``lib()`` raises; if it is present but no
CUDA device is, the first entry point called returns EOC_ERR_CUDA and ``check()`` raises.
"""

from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# EOC_LIBRARY_PATH: load another build of the same library (kernel tuning variants, tools/build_variants.py)
LIB_PATH = os.environ.get("EOC_LIBRARY_PATH") or os.path.join(_HERE, "_native", "libeoc_sm100.so")

# eoc_dtype
F32, F64, U8, I16, U16, I32 = range(6)
DTYPE_CODE = {
    np.dtype(np.float32): F32,
    np.dtype(np.float64): F64,
    np.dtype(np.uint8): U8,
    np.dtype(np.int16): I16,
    np.dtype(np.uint16): U16,
    np.dtype(np.int32): I32,
}
# eoc_op
(OP_SUM, OP_MEAN, OP_MIN, OP_MAX, OP_COUNT, OP_VARIANCE, OP_SD, OP_PRODUCT, OP_COUNT_ALL, OP_NOT_NAN,
 OP_FIRST, OP_LAST, OP_ANY, OP_ALL) = range(14)
NUM_OPS = 14
INT64_OPS = (OP_COUNT, OP_NOT_NAN)
SCAN_SUM, SCAN_PRODUCT, SCAN_MIN, SCAN_MAX = range(4)
ELEM_CLIP, ELEM_LINEAR_SCALE_RANGE, ELEM_ABSOLUTE, ELEM_SGN, ELEM_SQRT, ELEM_FLOOR, ELEM_CEIL, ELEM_TRUNC = range(8)
(BIN_ADD, BIN_SUBTRACT, BIN_MULTIPLY, BIN_DIVIDE, BIN_EQ, BIN_NEQ, BIN_GT, BIN_GTE, BIN_LT, BIN_LTE, BIN_AND, BIN_OR,
 BIN_XOR) = range(13)
BINARY_MAX_DIMS = 8
ABI_VERSION = 3  # EOC_ABI_VERSION of include/eoc.h
SELECT_MAX_R = 512
SELECT_MAX_Q = 16

EXPORTS = [
    "eoc_abi_version", "eoc_last_error", "eoc_device_info", "eoc_reduce_workspace_bytes", "eoc_reduce",
    "eoc_select", "eoc_segmented_reduce", "eoc_nd_reduce", "eoc_synth_fill", "eoc_reduce_host",
    "eoc_select_host", "eoc_launch_count", "eoc_select_fused", "eoc_select_fused_host", "eoc_selftest_nd_u16",
    "eoc_masked_reduce", "eoc_masked_select", "eoc_mask_canonical", "eoc_scan", "eoc_sort", "eoc_interpolate_linear", "eoc_elementwise",
    "eoc_binary", "eoc_reduce_rows_static", "eoc_order", "eoc_take", "eoc_select_general_workspace_bytes", "eoc_select_general", "eoc_rasterize_polygon",
    "eoc_zonal_reduce_workspace_bytes", "eoc_zonal_reduce", "eoc_segmented_reduce_host", "eoc_nd_reduce_host",
    "eoc_scan_host", "eoc_sort_host", "eoc_host_release", "eoc_apply_tape", "eoc_interpolate_linear_host",
    "eoc_elementwise_host", "eoc_apply_tape_host",
]
ROWS_STATIC_BYTES = 512 * 1024  # EOC_ROWS_STATIC_BYTES


class TapeOp(C.Structure):
    """struct eoc_tape_op"""

    _fields_ = [("code", C.c_int32), ("operand", C.c_int32), ("imm", C.c_double)]


TAPE_MAX_OPS = 24
(TAPE_LOAD, TAPE_STORE_T0, TAPE_STORE_T1, TAPE_ADD, TAPE_SUB, TAPE_RSUB, TAPE_MUL, TAPE_DIV, TAPE_RDIV, TAPE_MAXNUM,
 TAPE_MINNUM, TAPE_POW, TAPE_RPOW, TAPE_MOD, TAPE_RMOD, TAPE_ATAN2, TAPE_RATAN2, TAPE_ABS, TAPE_SGN, TAPE_SQRT,
 TAPE_FLOOR, TAPE_CEIL, TAPE_TRUNC, TAPE_ROUND, TAPE_ROUND_NEG, TAPE_EXP, TAPE_LN, TAPE_SIN, TAPE_COS, TAPE_TAN,
 TAPE_ASIN, TAPE_ACOS, TAPE_ATAN, TAPE_SINH, TAPE_COSH, TAPE_TANH, TAPE_ASINH, TAPE_ACOSH, TAPE_ATANH) = range(39)


class View(C.Structure):
    """struct eoc_view"""

    _fields_ = [
        ("data", C.c_void_p),
        ("dtype", C.c_int32),
        ("outer", C.c_int64),
        ("R", C.c_int64),
        ("inner", C.c_int64),
        ("stride_outer", C.c_int64),
        ("stride_R", C.c_int64),
    ]


class EocError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libeoc_sm100 error {code}: {message}")
        self.code = code


_lib = None
_lock = threading.Lock()


def _declare(lib):
    vp, i32, i64, u32, u64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_size_t
    outs_t = C.POINTER(C.c_void_p)
    lib.eoc_abi_version.restype = C.c_int
    lib.eoc_last_error.restype = C.c_char_p
    lib.eoc_device_info.argtypes = [C.POINTER(C.c_int32), C.POINTER(i64), C.POINTER(i64)]
    lib.eoc_reduce_workspace_bytes.argtypes = [C.POINTER(View), u32]
    lib.eoc_reduce_workspace_bytes.restype = sz
    lib.eoc_reduce.argtypes = [C.POINTER(View), u32, i32, i32, outs_t, vp, sz, vp]
    lib.eoc_select.argtypes = [C.POINTER(View), i32, vp, i32, C.POINTER(C.c_double), vp, i32, vp]
    lib.eoc_segmented_reduce.argtypes = [C.POINTER(View), i32, vp, vp, i32, u32, i32, i32, outs_t, vp]
    lib.eoc_nd_reduce.argtypes = [C.POINTER(View), C.POINTER(View), u32, i32, i32, outs_t, vp, i32, vp]
    lib.eoc_synth_fill.argtypes = [vp, i32, i64, i64, i64, i64, u64, u32, vp]
    lib.eoc_reduce_host.argtypes = [C.POINTER(View), u32, i32, i32, outs_t, sz]
    lib.eoc_select_host.argtypes = [C.POINTER(View), i32, vp, i32, C.POINTER(C.c_double), vp, i32, sz]
    lib.eoc_select_fused.argtypes = [C.POINTER(View), i32, vp, i32, C.POINTER(C.c_double), vp, i32, u32, i32,
                                     outs_t, vp]
    lib.eoc_select_fused_host.argtypes = [C.POINTER(View), i32, vp, i32, C.POINTER(C.c_double), vp, i32, u32, i32,
                                          outs_t, sz]
    lib.eoc_selftest_nd_u16.argtypes = [vp, vp]
    lib.eoc_selftest_nd_u16.restype = C.c_int
    lib.eoc_masked_reduce.argtypes = [C.POINTER(View), C.POINTER(View), C.c_double, u32, i32, i32, outs_t, vp, vp]
    lib.eoc_masked_select.argtypes = [C.POINTER(View), C.POINTER(View), C.c_double, i32, vp, i32,
                                      C.POINTER(C.c_double), vp, i32, vp]
    lib.eoc_mask_canonical.argtypes = [vp, i32, i64, vp, vp]
    lib.eoc_scan.argtypes = [C.POINTER(View), i32, i32, vp, vp]
    lib.eoc_scan.restype = C.c_int
    lib.eoc_sort.argtypes = [C.POINTER(View), i32, i32, vp, vp]
    lib.eoc_sort.restype = C.c_int
    lib.eoc_interpolate_linear.argtypes = [C.POINTER(View), vp, vp, vp]
    lib.eoc_interpolate_linear.restype = C.c_int
    lib.eoc_elementwise.argtypes = [vp, i32, i64, i32, C.POINTER(C.c_double), vp, vp]
    lib.eoc_elementwise.restype = C.c_int
    lib.eoc_binary.argtypes = [vp, vp, i32, i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), i32, C.c_double, i32,
                               vp, vp]
    lib.eoc_binary.restype = C.c_int
    lib.eoc_reduce_rows_static.argtypes = [C.POINTER(View), u32, i32, i32, outs_t, vp]
    lib.eoc_reduce_rows_static.restype = C.c_int
    lib.eoc_order.argtypes = [C.POINTER(View), vp, vp]
    lib.eoc_order.restype = C.c_int
    lib.eoc_take.argtypes = [C.POINTER(View), vp, i64, vp, vp]
    lib.eoc_take.restype = C.c_int
    lib.eoc_select_general_workspace_bytes.argtypes = [C.POINTER(View)]
    lib.eoc_select_general_workspace_bytes.restype = sz
    lib.eoc_select_general.argtypes = [C.POINTER(View), i32, vp, i32, C.POINTER(C.c_double), vp, i32, vp, i32, i32, vp,
                                       sz, vp]
    lib.eoc_select_general.restype = C.c_int
    lib.eoc_rasterize_polygon.argtypes = [vp, vp, i64, i64, vp, vp, i32, vp, vp]
    lib.eoc_rasterize_polygon.restype = C.c_int
    lib.eoc_zonal_reduce_workspace_bytes.argtypes = [i64, i64]
    lib.eoc_zonal_reduce_workspace_bytes.restype = sz
    lib.eoc_zonal_reduce.argtypes = [vp, i32, i64, i64, i64, i64, i64, vp, i64, i64, u32, i32, i32, outs_t, vp, sz, vp]
    lib.eoc_zonal_reduce.restype = C.c_int
    lib.eoc_segmented_reduce_host.argtypes = [C.POINTER(View), i32, vp, vp, i32, u32, i32, i32, outs_t, sz]
    lib.eoc_nd_reduce_host.argtypes = [C.POINTER(View), C.POINTER(View), u32, i32, i32, outs_t, vp, i32, sz]
    lib.eoc_scan_host.argtypes = [C.POINTER(View), i32, i32, vp, sz]
    lib.eoc_sort_host.argtypes = [C.POINTER(View), i32, i32, vp, sz]
    for name in ("eoc_segmented_reduce_host", "eoc_nd_reduce_host", "eoc_scan_host", "eoc_sort_host", "eoc_host_release"):
        getattr(lib, name).restype = C.c_int
    lib.eoc_apply_tape.argtypes = [vp, vp, i32, i64, C.POINTER(TapeOp), i32, vp, i32, vp]
    lib.eoc_apply_tape.restype = C.c_int
    lib.eoc_interpolate_linear_host.argtypes = [C.POINTER(View), vp, vp, sz]
    lib.eoc_elementwise_host.argtypes = [vp, i32, i64, i32, C.POINTER(C.c_double), vp, sz]
    lib.eoc_apply_tape_host.argtypes = [vp, vp, i32, i64, C.POINTER(TapeOp), i32, vp, i32, sz]
    for name in ("eoc_interpolate_linear_host", "eoc_elementwise_host", "eoc_apply_tape_host"):
        getattr(lib, name).restype = C.c_int
    lib.eoc_launch_count.argtypes = [i32]
    lib.eoc_launch_count.restype = i64
    for name in ("eoc_device_info", "eoc_reduce", "eoc_select", "eoc_segmented_reduce", "eoc_nd_reduce",
                 "eoc_synth_fill", "eoc_reduce_host", "eoc_select_host", "eoc_select_fused",
                 "eoc_select_fused_host", "eoc_masked_reduce", "eoc_masked_select", "eoc_mask_canonical"):
        getattr(lib, name).restype = C.c_int


def lib():
    """The loaded library.  Raises if it has not been built -- there is no fallback path."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise ImportError(
                        f"{LIB_PATH} is missing: build it with `python -m openeo_processes_dask_b200.build` "
                        "(there is no CPU fallback for the reducer kernels)")
                handle = C.CDLL(LIB_PATH)
                _declare(handle)
                if handle.eoc_abi_version() != ABI_VERSION:
                    raise ImportError("libeoc_sm100.so ABI version mismatch")
                _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise EocError(rc, lib().eoc_last_error().decode("utf-8", "replace"))


def make_view(ptr: int, dtype_code: int, outer: int, R: int, inner: int, stride_outer: int, stride_R: int) -> View:
    return View(C.c_void_p(ptr), dtype_code, outer, R, inner, stride_outer, stride_R)


def outs_array(mapping: dict) -> "C.Array":
    """mapping: op id -> raw pointer (int).  Returns void*[NUM_OPS]."""
    arr = (C.c_void_p * NUM_OPS)()
    for op, ptr in mapping.items():
        arr[op] = C.c_void_p(ptr)
    return arr


def ops_mask(ops) -> int:
    m = 0
    for op in ops:
        m |= 1 << op
    return m
