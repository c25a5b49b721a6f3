// mask -> reduce fused: eoc_masked_reduce, eoc_mask_canonical.
// The reference materialises data.where(_not(mask), replacement) (cubes/mask.py:112) -- one read and one
// write of the whole cube -- before the reduce reads it again; here the mask byte is consumed in the load
// stage of the column kernel.
#include <type_traits>

#include "column_launch.cuh"

namespace eoc {

namespace {

template <typename In>
int launch_masked_t(const eoc_view& in, const eoc_view& mk, double replacement, const ColumnParams& P,
                    void* masked_out, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  MaskedSource<In, A> src{(const In*)in.data, (const uint8_t*)mk.data, in.stride_outer, in.stride_R,
                          mk.stride_outer, mk.stride_R, (A)replacement, (A*)masked_out};
  constexpr int VM = VecMax<In>::value;
  auto ok = [&](int vec) {
    return aligned_for(in.data, in.stride_outer, in.stride_R, in.inner, sizeof(In), vec) &&
           aligned_for(mk.data, mk.stride_outer, mk.stride_R, mk.inner, 1, vec) &&
           ((uintptr_t)masked_out % (sizeof(A) * vec) == 0);
  };
  if (ok(VM)) return launch_columns_g<MaskedSource<In, A>, A, VM>(src, P, s);
  if constexpr (VM > 4) {
    if (ok(4)) return launch_columns_g<MaskedSource<In, A>, A, 4>(src, P, s);
  }
  return launch_columns_g<MaskedSource<In, A>, A, 1>(src, P, s);
}

// dst[i] = 1 where src[i] is truthy and not NaN (the samples data.where(_not(mask), ...) replaces), else 0
template <typename T>
__global__ void __launch_bounds__(256) k_mask_canonical(const T* __restrict__ src, uint8_t* __restrict__ dst, int64_t n) {
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
    const T v = src[i];
    dst[i] = (v != (T)0 && v == v) ? 1 : 0;
  }
}

template <typename T>
int launch_canonical_t(const void* src, uint8_t* dst, int64_t n, cudaStream_t s) {
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  k_mask_canonical<T><<<(unsigned)blocks, 256, 0, s>>>((const T*)src, dst, n);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace

int launch_masked_reduce(const eoc_view& in, const eoc_view& mk, double replacement, uint32_t ops,
                         int ignore_nodata, int ddof, const OutPtrs& outs, void* masked_out, cudaStream_t s) {
  if (int rc = check_outs(ops, outs)) return rc;
  if (mk.dtype != EOC_U8) EOC_FAIL(EOC_ERR_INVALID, "the mask must be EOC_U8 (see eoc_mask_canonical)");
  if (mk.outer != in.outer || mk.R != in.R || mk.inner != in.inner)
    EOC_FAIL(EOC_ERR_INVALID, "data and mask views differ in shape (broadcast with stride 0, not with extent 1)");
  if (ops == 0 && !masked_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute: ops == 0 and masked_out == NULL");
  if (in.R > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld too large", (long long)in.R);
  const ColumnParams P = make_params(in, ops, ignore_nodata, ddof, outs);
  switch (in.dtype) {
    case EOC_F32: return launch_masked_t<float>(in, mk, replacement, P, masked_out, s);
    case EOC_F64: return launch_masked_t<double>(in, mk, replacement, P, masked_out, s);
    case EOC_U8: return launch_masked_t<uint8_t>(in, mk, replacement, P, masked_out, s);
    case EOC_I16: return launch_masked_t<int16_t>(in, mk, replacement, P, masked_out, s);
    case EOC_U16: return launch_masked_t<uint16_t>(in, mk, replacement, P, masked_out, s);
    case EOC_I32: return launch_masked_t<int32_t>(in, mk, replacement, P, masked_out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

int launch_mask_canonical(const void* src, int dtype, int64_t n, uint8_t* dst, cudaStream_t s) {
  if (n == 0) return EOC_OK;
  switch (dtype) {
    case EOC_F32: return launch_canonical_t<float>(src, dst, n, s);
    case EOC_F64: return launch_canonical_t<double>(src, dst, n, s);
    case EOC_U8: return launch_canonical_t<uint8_t>(src, dst, n, s);
    case EOC_I16: return launch_canonical_t<int16_t>(src, dst, n, s);
    case EOC_U16: return launch_canonical_t<uint16_t>(src, dst, n, s);
    case EOC_I32: return launch_canonical_t<int32_t>(src, dst, n, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", dtype);
}

}  // namespace eoc
