// Shared host/device helpers for libeoc_sm100.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/eoc.h"

namespace eoc {

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;

#define EOC_FAIL(code, ...)        \
  do {                             \
    ::eoc::set_error(__VA_ARGS__); \
    return (code);                 \
  } while (0)

#define EOC_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::eoc::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                       __LINE__);                                                        \
      return EOC_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

#define EOC_LAUNCH_CHECK()                     \
  do {                                         \
    ::eoc::g_launches.fetch_add(1);            \
    EOC_CUDA(cudaPeekAtLastError());           \
  } while (0)

inline size_t dtype_size(int dt) {
  switch (dt) {
    case EOC_F32: return 4;
    case EOC_F64: return 8;
    case EOC_U8: return 1;
    case EOC_I16: return 2;
    case EOC_U16: return 2;
    case EOC_I32: return 4;
    default: return 0;
  }
}
// compute dtype: f32 stays f32, everything else is reduced in f64 (see eoc.h)
inline int acc_dtype(int dt) { return dt == EOC_F32 ? EOC_F32 : EOC_F64; }
// compute dtype of the normalized difference: the "uint16 -> fp32" cast of the workflow
inline int nd_acc_dtype(int dt) {
  return (dt == EOC_F64 || dt == EOC_I32) ? EOC_F64 : EOC_F32;
}

int check_view(const eoc_view* v, const char* what);

// number of SMs of the current device (cached per device)
int sm_count();

// ---- launchers implemented in the .cu files --------------------------------------------------
struct OutPtrs {
  void* p[EOC_NUM_OPS];
};

int launch_reduce(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                  void* workspace, size_t workspace_bytes, cudaStream_t s);
// the order-dependent reducers (first, last, any, all): column kernel only
constexpr uint32_t kOrderOps =
    (1u << EOC_OP_FIRST) | (1u << EOC_OP_LAST) | (1u << EOC_OP_ANY) | (1u << EOC_OP_ALL);
size_t reduce_workspace_bytes(const eoc_view& in, uint32_t ops);
int launch_segmented(const eoc_view& in, int nseg, const int32_t* seg_offsets, const int32_t* members,
                     int max_seg_len, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                     cudaStream_t s);
int launch_nd_reduce(const eoc_view& a, const eoc_view& b, uint32_t ops, int ignore_nodata, int ddof,
                     const OutPtrs& outs, void* nd_out, bool integer_wrap, cudaStream_t s);
// eoc_sort rides on the selection kernels: the whole sorted column as an extra output
struct SortSpec {
  void* out = nullptr;  // dense [outer, R, inner] in the compute dtype
  int descending = 0, nan_first = 0;
};
int launch_select(const eoc_view& in, int ignore_nodata, void* median_out, int nq, const double* probs,
                  void* quant_out, int quant_dtype, uint32_t ops, int ddof, const OutPtrs& outs, cudaStream_t s,
                  const SortSpec& sort = SortSpec());
int launch_masked_reduce(const eoc_view& in, const eoc_view& mk, double replacement, uint32_t ops,
                         int ignore_nodata, int ddof, const OutPtrs& outs, void* masked_out, cudaStream_t s);
int launch_mask_canonical(const void* src, int dtype, int64_t n, uint8_t* dst, cudaStream_t s);
int launch_masked_select(const eoc_view& in, const eoc_view& mk, double replacement, int ignore_nodata,
                         void* median_out, int nq, const double* probs, void* quant_out, int quant_dtype,
                         cudaStream_t s);
int launch_scan(const eoc_view& in, int op, int ignore_nodata, void* out, cudaStream_t s);
int launch_interpolate_linear(const eoc_view& in, const double* xs, void* out, cudaStream_t s);
int launch_elementwise(const void* in, int dtype, int64_t n, int op, const double* params, void* out, cudaStream_t s);
int launch_binary(const void* a, const void* b, int dtype, int ndim, const int64_t* shape, const int64_t* sa,
                  const int64_t* sb, int op, double scalar, int scalar_side, void* out, cudaStream_t s);
int launch_synth(void* out, int dtype, int64_t planes, int64_t plane_stride, int64_t first_index,
                 int64_t strip_len, uint64_t seed, uint32_t nan_threshold, cudaStream_t s);

}  // namespace eoc
