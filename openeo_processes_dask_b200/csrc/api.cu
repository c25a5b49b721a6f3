// extern "C" surface of libeoc_sm100.so (see include/eoc.h for the contract of every entry point).
#include <cstring>
#include <mutex>

#include "eoc_common.cuh"

namespace eoc {

static thread_local char tl_error[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static std::mutex mu;
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return 148;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      cudaGetLastError();
      n = 148;
    }
    cached[dev] = n;
  }
  return cached[dev];
}

int check_view(const eoc_view* v, const char* what) {
  if (!v) EOC_FAIL(EOC_ERR_INVALID, "%s: view is NULL", what);
  if (dtype_size(v->dtype) == 0) EOC_FAIL(EOC_ERR_INVALID, "%s: unknown dtype %d", what, v->dtype);
  if (v->outer < 0 || v->R < 0 || v->inner < 0)
    EOC_FAIL(EOC_ERR_INVALID, "%s: negative extent (%lld, %lld, %lld)", what, (long long)v->outer,
             (long long)v->R, (long long)v->inner);
  if (v->outer * v->R * v->inner > 0 && !v->data) EOC_FAIL(EOC_ERR_INVALID, "%s: data is NULL", what);
  if (v->stride_R < 0 || v->stride_outer < 0) EOC_FAIL(EOC_ERR_INVALID, "%s: negative stride", what);
  return EOC_OK;
}

static OutPtrs gather_outs(uint32_t ops, void* const* outs) {
  OutPtrs o;
  for (int id = 0; id < EOC_NUM_OPS; ++id) o.p[id] = (outs && (ops & (1u << id))) ? outs[id] : nullptr;
  return o;
}

}  // namespace eoc

using namespace eoc;

extern "C" {

int eoc_abi_version(void) { return EOC_ABI_VERSION; }

const char* eoc_last_error(void) { return tl_error; }

int64_t eoc_launch_count(int reset) {
  return reset ? g_launches.exchange(0) : g_launches.load();
}

int eoc_device_info(int32_t* sms, int64_t* total_bytes, int64_t* free_bytes) {
  size_t f = 0, t = 0;
  EOC_CUDA(cudaMemGetInfo(&f, &t));
  if (sms) *sms = sm_count();
  if (total_bytes) *total_bytes = (int64_t)t;
  if (free_bytes) *free_bytes = (int64_t)f;
  return EOC_OK;
}

size_t eoc_reduce_workspace_bytes(const eoc_view* in, uint32_t ops) {
  if (!in) return 0;
  return reduce_workspace_bytes(*in, ops);
}

int eoc_reduce(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
               void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_view(in, "eoc_reduce")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_reduce: outs is NULL");
  return launch_reduce(*in, ops, ignore_nodata, ddof, gather_outs(ops, outs), workspace, workspace_bytes,
                       (cudaStream_t)stream);
}

int eoc_select(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
               void* quant_out, int quant_dtype, void* stream) {
  if (int rc = check_view(in, "eoc_select")) return rc;
  return launch_select(*in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, 0u, 1,
                       gather_outs(0u, nullptr), (cudaStream_t)stream);
}

int eoc_select_fused(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                     void* quant_out, int quant_dtype, uint32_t ops, int ddof, void* const* outs,
                     void* stream) {
  if (int rc = check_view(in, "eoc_select_fused")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_select_fused: outs is NULL");
  return launch_select(*in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, ops, ddof,
                       gather_outs(ops, outs), (cudaStream_t)stream);
}

int eoc_segmented_reduce(const eoc_view* in, int nseg, const int32_t* seg_offsets, const int32_t* members,
                         int max_seg_len, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                         void* stream) {
  if (int rc = check_view(in, "eoc_segmented_reduce")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_segmented_reduce: outs is NULL");
  return launch_segmented(*in, nseg, seg_offsets, members, max_seg_len, ops, ignore_nodata, ddof,
                          gather_outs(ops, outs), (cudaStream_t)stream);
}

int eoc_nd_reduce(const eoc_view* a, const eoc_view* b, uint32_t ops, int ignore_nodata, int ddof,
                  void* const* outs, void* nd_out, int integer_wrap, void* stream) {
  if (int rc = check_view(a, "eoc_nd_reduce(a)")) return rc;
  if (int rc = check_view(b, "eoc_nd_reduce(b)")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_nd_reduce: outs is NULL");
  return launch_nd_reduce(*a, *b, ops, ignore_nodata, ddof, gather_outs(ops, outs), nd_out, integer_wrap != 0,
                          (cudaStream_t)stream);
}

int eoc_masked_reduce(const eoc_view* in, const eoc_view* mask, double replacement, uint32_t ops, int ignore_nodata,
                      int ddof, void* const* outs, void* masked_out, void* stream) {
  if (int rc = check_view(in, "eoc_masked_reduce(in)")) return rc;
  if (int rc = check_view(mask, "eoc_masked_reduce(mask)")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_masked_reduce: outs is NULL");
  return launch_masked_reduce(*in, *mask, replacement, ops, ignore_nodata, ddof, gather_outs(ops, outs), masked_out,
                              (cudaStream_t)stream);
}

int eoc_masked_select(const eoc_view* in, const eoc_view* mask, double replacement, int ignore_nodata,
                      void* median_out, int nq, const double* probs, void* quant_out, int quant_dtype,
                      void* stream) {
  if (int rc = check_view(in, "eoc_masked_select(in)")) return rc;
  if (int rc = check_view(mask, "eoc_masked_select(mask)")) return rc;
  return launch_masked_select(*in, *mask, replacement, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype,
                              (cudaStream_t)stream);
}

int eoc_mask_canonical(const void* src, int dtype, int64_t n, uint8_t* dst, void* stream) {
  if (n < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_mask_canonical: negative length");
  if (n > 0 && (!src || !dst)) EOC_FAIL(EOC_ERR_INVALID, "eoc_mask_canonical: NULL pointer");
  return launch_mask_canonical(src, dtype, n, dst, (cudaStream_t)stream);
}

int eoc_sort(const eoc_view* in, int descending, int nan_first, void* out, void* stream) {
  if (int rc = check_view(in, "eoc_sort")) return rc;
  if (!out && in->outer * in->R * in->inner > 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_sort: out is NULL");
  if (in->outer * in->R * in->inner == 0) return EOC_OK;
  SortSpec spec;
  spec.out = out;
  spec.descending = descending;
  spec.nan_first = nan_first;
  return launch_select(*in, 1, nullptr, 0, nullptr, nullptr, EOC_F64, 0u, 1, gather_outs(0u, nullptr),
                       (cudaStream_t)stream, spec);
}

int eoc_interpolate_linear(const eoc_view* in, const double* x, void* out, void* stream) {
  if (int rc = check_view(in, "eoc_interpolate_linear")) return rc;
  return launch_interpolate_linear(*in, x, out, (cudaStream_t)stream);
}

int eoc_elementwise(const void* in, int dtype, int64_t n, int op, const double* params, void* out, void* stream) {
  if (n < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_elementwise: negative length");
  return launch_elementwise(in, dtype, n, op, params, out, (cudaStream_t)stream);
}

int eoc_binary(const void* a, const void* b, int dtype, int ndim, const int64_t* shape, const int64_t* a_strides,
               const int64_t* b_strides, int op, double scalar, int scalar_side, void* out, void* stream) {
  return launch_binary(a, b, dtype, ndim, shape, a_strides, b_strides, op, scalar, scalar_side, out,
                       (cudaStream_t)stream);
}

int eoc_scan(const eoc_view* in, int op, int ignore_nodata, void* out, void* stream) {
  if (int rc = check_view(in, "eoc_scan")) return rc;
  return launch_scan(*in, op, ignore_nodata, out, (cudaStream_t)stream);
}

int eoc_synth_fill(void* out, int dtype, int64_t planes, int64_t plane_stride, int64_t first_index,
                   int64_t strip_len, uint64_t seed, uint32_t nan_threshold, void* stream) {
  return launch_synth(out, dtype, planes, plane_stride, first_index, strip_len, seed, nan_threshold,
                      (cudaStream_t)stream);
}

}  // extern "C"
