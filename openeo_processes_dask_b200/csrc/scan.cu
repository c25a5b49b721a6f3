// "Axis kept" array processes of apply_dimension: cumulative sum / product / min / max
// along R of an [outer, R, inner] view, output dense [outer, R, inner].
// Reference: math.py:125-174 -- np.nancumsum / np.nancumprod / np.minimum.accumulate on a NaN-filled copy,
// then result[nan_mask] = nan.  One thread owns VEC adjacent columns and walks R with a running value,
// U rows requested ahead; reads and writes are both coalesced across the warp: 2 * R * sizeof bytes per
// column, HBM-bound.  The running value is updated exactly as numpy does (first element taken as is, a NaN
// slot contributes the neutral element +0 / 1 / +-inf, sequential order), so results are bit-identical.
#include <type_traits>

#include "column_launch.cuh"

namespace eoc {

namespace {

template <typename A, int OP> struct ScanOp;
template <typename A> struct ScanOp<A, EOC_SCAN_SUM> {
  static __device__ __forceinline__ A neutral() { return (A)0; }
  static __device__ __forceinline__ A step(A acc, A x) { return add_rn(acc, x); }
};
template <typename A> struct ScanOp<A, EOC_SCAN_PRODUCT> {
  static __device__ __forceinline__ A neutral() { return (A)1; }
  static __device__ __forceinline__ A step(A acc, A x) { return mul_rn(acc, x); }
};
// np.minimum / np.maximum propagate NaN (ignore_nodata = false); with the NaN slots filled they never see one
template <typename A> struct ScanOp<A, EOC_SCAN_MIN> {
  static __device__ __forceinline__ A neutral() { return Lim<A>::inf(); }
  static __device__ __forceinline__ A step(A acc, A x) { return (is_nan(acc) || is_nan(x)) ? Lim<A>::nan() : min_(acc, x); }
};
template <typename A> struct ScanOp<A, EOC_SCAN_MAX> {
  static __device__ __forceinline__ A neutral() { return -Lim<A>::inf(); }
  static __device__ __forceinline__ A step(A acc, A x) { return (is_nan(acc) || is_nan(x)) ? Lim<A>::nan() : max_(acc, x); }
};

template <typename In, typename A, int VEC, int OP, int U>
__global__ void __launch_bounds__(256) k_scan(const PlainSource<In, A> src, int64_t outer, int R, int64_t inner,
                                              int ignore_nodata, A* __restrict__ out) {
  const int64_t ivec = inner / VEC;
  const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= outer * ivec) return;
  const int64_t o = item / ivec;
  const int64_t i = (item - o * ivec) * VEC;
  A acc[VEC];
  A* dst = out + o * R * inner + i;
  auto consume = [&](const A (&x)[VEC], int r) {
    A y[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const bool nan = is_nan(x[k]);
      const A v = (ignore_nodata && nan) ? ScanOp<A, OP>::neutral() : x[k];
      acc[k] = (r == 0) ? v : ScanOp<A, OP>::step(acc[k], v);
      y[k] = nan ? Lim<A>::nan() : acc[k];
    }
    st_vec<A, VEC>(dst + (int64_t)r * inner, y);
  };
  int r = 0;
  for (; r + U <= R; r += U) {
    A x[U][VEC];
#pragma unroll
    for (int u = 0; u < U; ++u) src.template load<VEC>(o, r + u, i, x[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) consume(x[u], r + u);
  }
  for (; r < R; ++r) {
    A x[VEC];
    src.template load<VEC>(o, r, i, x);
    consume(x, r);
  }
}

template <typename In, typename A, int VEC>
int launch_scan_v(const eoc_view& in, int op, int ignore_nodata, A* out, cudaStream_t s) {
  const int64_t items = in.outer * (in.inner / VEC);
  if (items == 0 || in.R == 0) return EOC_OK;
  const int64_t blocks = (items + 255) / 256;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  PlainSource<In, A> src{(const In*)in.data, in.stride_outer, in.stride_R};
  constexpr int U = 4;
#define EOC_SCAN_GO(OPV) \
  k_scan<In, A, VEC, OPV, U><<<(unsigned)blocks, 256, 0, s>>>(src, in.outer, (int)in.R, in.inner, ignore_nodata, out)
  switch (op) {
    case EOC_SCAN_SUM: EOC_SCAN_GO(EOC_SCAN_SUM); break;
    case EOC_SCAN_PRODUCT: EOC_SCAN_GO(EOC_SCAN_PRODUCT); break;
    case EOC_SCAN_MIN: EOC_SCAN_GO(EOC_SCAN_MIN); break;
    case EOC_SCAN_MAX: EOC_SCAN_GO(EOC_SCAN_MAX); break;
    default: EOC_FAIL(EOC_ERR_INVALID, "unknown scan op %d", op);
  }
#undef EOC_SCAN_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In>
int launch_scan_t(const eoc_view& in, int op, int ignore_nodata, void* out, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  constexpr int VM = (sizeof(A) == 4) ? 4 : 2;  // 16-byte stores of the compute dtype
  const bool al = aligned_for(in.data, in.stride_outer, in.stride_R, in.inner, sizeof(In), VM) &&
                  ((uintptr_t)out % (sizeof(A) * VM) == 0);
  if (al) return launch_scan_v<In, A, VM>(in, op, ignore_nodata, (A*)out, s);
  return launch_scan_v<In, A, 1>(in, op, ignore_nodata, (A*)out, s);
}

}  // namespace

int launch_scan(const eoc_view& in, int op, int ignore_nodata, void* out, cudaStream_t s) {
  if (!out && in.outer * in.R * in.inner > 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_scan: out is NULL");
  if (in.R > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld too large", (long long)in.R);
  switch (in.dtype) {
    case EOC_F32: return launch_scan_t<float>(in, op, ignore_nodata, out, s);
    case EOC_F64: return launch_scan_t<double>(in, op, ignore_nodata, out, s);
    case EOC_U8: return launch_scan_t<uint8_t>(in, op, ignore_nodata, out, s);
    case EOC_I16: return launch_scan_t<int16_t>(in, op, ignore_nodata, out, s);
    case EOC_U16: return launch_scan_t<uint16_t>(in, op, ignore_nodata, out, s);
    case EOC_I32: return launch_scan_t<int32_t>(in, op, ignore_nodata, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

}  // namespace eoc
