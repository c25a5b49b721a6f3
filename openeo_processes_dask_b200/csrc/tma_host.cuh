// Host side of TMA: the CUtensorMap of an [outer, R, inner] view (shared by tile.cu and select_tile.cu).
// cuTensorMapEncodeTiled is resolved at run time through cudaGetDriverEntryPoint, so the library links
// against the (static) CUDA runtime only.
#pragma once

#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "eoc_common.cuh"

namespace eoc {

inline PFN_cuTensorMapEncodeTiled tma_encode_fn() {
  static std::once_flag once;
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    else
      cudaGetLastError();
  });
  return fn;
}

inline CUtensorMapDataType tma_dtype(int dt) {
  switch (dt) {
    case EOC_F32: return CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    case EOC_F64: return CU_TENSOR_MAP_DATA_TYPE_FLOAT64;
    case EOC_U8: return CU_TENSOR_MAP_DATA_TYPE_UINT8;
    case EOC_I16: return CU_TENSOR_MAP_DATA_TYPE_UINT16;
    case EOC_U16: return CU_TENSOR_MAP_DATA_TYPE_UINT16;
    default: return CU_TENSOR_MAP_DATA_TYPE_INT32;
  }
}

// TMA constraints on the view: 16-byte aligned base and strides, extents and coordinates within int32
inline bool tma_view_ok(const eoc_view& in) {
  const size_t es = dtype_size(in.dtype);
  if (!tma_encode_fn()) return false;
  if ((uintptr_t)in.data % 16 != 0 || (in.stride_R * es) % 16 != 0) return false;
  if (in.outer > 1 && (in.stride_outer * es) % 16 != 0) return false;
  if (in.inner >= (1LL << 31) || in.outer >= (1LL << 31) || in.R >= (1LL << 31)) return false;
  if ((uint64_t)in.stride_R * es >= (1ull << 40) || (uint64_t)in.stride_outer * es >= (1ull << 40)) return false;
  return true;
}

// The TMA-staged kernels walk tiles of W adjacent columns of ONE outer slice, so a view with a short `inner` -- the
// reference's test-fixture layout (x, y, t, bands): inner = bands = 4 -- would fill 4 of 128 lanes per tile (measured:
// median of a (1024, 1024, 100, 4) cube over t at 164 GB/s against 3323 for the per-thread-load kernel).  They apply
// only when the tiles are at least three quarters full.
inline bool tma_tiles_fill(const eoc_view& in, int W) {
  const int64_t tiles = (in.inner + W - 1) / W;
  return in.inner * 4 >= tiles * (int64_t)W * 3;
}

// 3-D tensor {inner, R, outer}, box {box_w, box_r, 1}, no swizzle, out-of-bounds elements read as zero
inline int tma_make_map(const eoc_view& in, int box_w, int box_r, CUtensorMap* map) {
  const size_t es = dtype_size(in.dtype);
  const cuuint64_t gdim[3] = {(cuuint64_t)in.inner, (cuuint64_t)in.R, (cuuint64_t)in.outer};
  const cuuint64_t so = in.outer > 1 ? (cuuint64_t)in.stride_outer : (cuuint64_t)in.stride_R * (cuuint64_t)in.R;
  const cuuint64_t gstr[2] = {(cuuint64_t)in.stride_R * es, so * es};
  const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_r, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult rc = tma_encode_fn()(map, tma_dtype(in.dtype), 3, const_cast<void*>(in.data), gdim, gstr, box,
                                      estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) EOC_FAIL(EOC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)rc);
  return EOC_OK;
}

}  // namespace eoc
