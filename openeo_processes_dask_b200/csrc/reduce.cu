// Launchers of the streaming reducers (eoc_reduce, eoc_segmented_reduce, eoc_nd_reduce).
#include <cstdlib>
#include <type_traits>

#include "column_launch.cuh"
#include "rows_kernel.cuh"

namespace eoc {

namespace {

template <typename In>
int launch_plain_t(const eoc_view& in, const ColumnParams& P, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  PlainSource<In, A> src{(const In*)in.data, in.stride_outer, in.stride_R};
  constexpr int VM = VecMax<In>::value;
  // vector path also needs the outputs aligned: out index (o*nseg+s)*inner + i, i % VEC == 0
  if (aligned_for(in.data, in.stride_outer, in.stride_R, in.inner, sizeof(In), VM))
    return launch_columns_g<PlainSource<In, A>, A, VM>(src, P, s);
  if constexpr (VM > 4) {
    if (aligned_for(in.data, in.stride_outer, in.stride_R, in.inner, sizeof(In), 4))
      return launch_columns_g<PlainSource<In, A>, A, 4>(src, P, s);
  }
  return launch_columns_g<PlainSource<In, A>, A, 1>(src, P, s);
}

template <typename In, typename A, bool WRAP = false>
int launch_nd_t(const eoc_view& a, const eoc_view& b, const ColumnParams& P, void* nd_out,
                cudaStream_t s) {
  using Src = NdSource<In, A, WRAP>;
  Src src{(const In*)a.data, (const In*)b.data, a.stride_outer, a.stride_R,
          b.stride_outer, b.stride_R, (A*)nd_out};
  constexpr int VM = VecMax<In>::value;
  const bool al = aligned_for(a.data, a.stride_outer, a.stride_R, a.inner, sizeof(In), VM) &&
                  aligned_for(b.data, b.stride_outer, b.stride_R, b.inner, sizeof(In), VM) &&
                  ((uintptr_t)nd_out % (sizeof(A) * VM) == 0);
  if (al) return launch_columns_g<Src, A, VM>(src, P, s);
  if constexpr (VM > 4) {
    const bool al4 = aligned_for(a.data, a.stride_outer, a.stride_R, a.inner, sizeof(In), 4) &&
                     aligned_for(b.data, b.stride_outer, b.stride_R, b.inner, sizeof(In), 4) &&
                     ((uintptr_t)nd_out % (sizeof(A) * 4) == 0);
    if (al4) return launch_columns_g<Src, A, 4>(src, P, s);
  }
  return launch_columns_g<Src, A, 1>(src, P, s);
}

int dispatch_plain(const eoc_view& in, const ColumnParams& P, cudaStream_t s) {
  // variance / sd read the column twice: stage it in shared memory by TMA when it fits
  if ((P.ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) && !(P.ops & kOrderOps)) {
    bool handled = false;
    if (int rc = launch_tile(in, P, s, &handled)) return rc;
    if (handled) return EOC_OK;
  }
  switch (in.dtype) {
    case EOC_F32: return launch_plain_t<float>(in, P, s);
    case EOC_F64: return launch_plain_t<double>(in, P, s);
    case EOC_U8: return launch_plain_t<uint8_t>(in, P, s);
    case EOC_I16: return launch_plain_t<int16_t>(in, P, s);
    case EOC_U16: return launch_plain_t<uint16_t>(in, P, s);
    case EOC_I32: return launch_plain_t<int32_t>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

}  // namespace

// Path choice: the column kernel needs enough columns to fill the GPU; a reduction with few, very
// long columns (spatial reduce: outer*inner small, R = Y*X) goes to the row-cooperative kernel.
static bool use_rows_kernel(const eoc_view& in) {
  return in.inner == 1 && in.stride_R == 1 && in.R >= 4096 && in.outer * in.inner < 65536;
}

size_t reduce_workspace_bytes(const eoc_view& in, uint32_t ops) {
  if (!use_rows_kernel(in) || (ops & kOrderOps)) return 0;
  return rows_workspace_bytes(in);
}

int launch_reduce(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                  void* workspace, size_t workspace_bytes, cudaStream_t s) {
  if (int rc = check_outs(ops, outs)) return rc;
  if (in.R > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld too large", (long long)in.R);
  if (use_rows_kernel(in) && !(ops & kOrderOps))  // the order-dependent reducers always take the column kernel
    return launch_rows(in, ops, ignore_nodata, ddof, outs, workspace, workspace_bytes, s);
  // many short contiguous rows (the reduced axis is the last one): a warp per row (EOC_ROWS_WARP=0: the column kernel).
  // Measured on B200: mean / sd over x of a (100, 4, 512, 1024) cube 946 / 484 -> 4384 / 3876 GB/s; rows of 100 samples
  // ((y, x, t) over t) are better off with one thread per row (2052 / 1109 vs 1651 / 988), hence the 256-sample floor.
  if (in.inner == 1 && in.stride_R == 1 && in.R >= 256 && !(ops & kOrderOps)) {
    const char* e = getenv("EOC_ROWS_WARP");
    if (!(e && e[0] == '0')) return launch_rows_warp(in, ops, ignore_nodata, ddof, outs, s);
  }
  return dispatch_plain(in, make_params(in, ops, ignore_nodata, ddof, outs), s);
}

int launch_segmented(const eoc_view& in, int nseg, const int32_t* seg_offsets, const int32_t* members,
                     int max_seg_len, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                     cudaStream_t s) {
  (void)max_seg_len;
  if (int rc = check_outs(ops, outs)) return rc;
  if (nseg < 1 || !seg_offsets) EOC_FAIL(EOC_ERR_INVALID, "segmented reduce needs nseg >= 1 and offsets");
  ColumnParams P = make_params(in, ops, ignore_nodata, ddof, outs);
  P.nseg = nseg;
  P.seg_offsets = seg_offsets;
  P.members = members;
  return dispatch_plain(in, P, s);
}

int launch_nd_reduce(const eoc_view& a, const eoc_view& b, uint32_t ops, int ignore_nodata, int ddof,
                     const OutPtrs& outs, void* nd_out, bool integer_wrap, cudaStream_t s) {
  if (int rc = check_outs(ops, outs)) return rc;
  if (a.dtype != b.dtype || a.outer != b.outer || a.R != b.R || a.inner != b.inner)
    EOC_FAIL(EOC_ERR_INVALID, "normalized difference operands differ in dtype or shape");
  if (ops == 0 && !nd_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute: ops == 0 and nd_out == NULL");
  const ColumnParams P = make_params(a, ops, ignore_nodata, ddof, outs);
  if (integer_wrap) {
    switch (a.dtype) {
      case EOC_U8: return launch_nd_t<uint8_t, double, true>(a, b, P, nd_out, s);
      case EOC_I16: return launch_nd_t<int16_t, double, true>(a, b, P, nd_out, s);
      case EOC_U16: return launch_nd_t<uint16_t, double, true>(a, b, P, nd_out, s);
      case EOC_I32: return launch_nd_t<int32_t, double, true>(a, b, P, nd_out, s);
      default: break;  // float cubes: the flag does not apply
    }
  }
  switch (a.dtype) {
    case EOC_F32: return launch_nd_t<float, float>(a, b, P, nd_out, s);
    case EOC_F64: return launch_nd_t<double, double>(a, b, P, nd_out, s);
    case EOC_U8: return launch_nd_t<uint8_t, float>(a, b, P, nd_out, s);
    case EOC_I16: return launch_nd_t<int16_t, float>(a, b, P, nd_out, s);
    case EOC_U16: return launch_nd_t<uint16_t, float>(a, b, P, nd_out, s);
    case EOC_I32: return launch_nd_t<int32_t, double>(a, b, P, nd_out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", a.dtype);
}

}  // namespace eoc

// ---- self test of the unsigned normalized-difference fast path --------------------------------------------
namespace eoc {
namespace {
__global__ void __launch_bounds__(256) k_selftest_nd_u16(unsigned long long* mismatches) {
  unsigned long long bad = 0;
  const uint64_t total = 1ull << 32;
  for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t a = (uint32_t)(e >> 16), b = (uint32_t)(e & 0xffffu);
    const float fa = (float)a, fb = (float)b;
    const float want = __fdiv_rn(__fsub_rn(fa, fb), __fadd_rn(fa, fb));
    const float got = nd_unsigned(u2f_exact(a), u2f_exact(b));
    const bool same = (want != want) ? (got != got) : (__float_as_uint(want) == __float_as_uint(got));
    bad += same ? 0 : 1;
  }
  if (bad) atomicAdd(mismatches, bad);
}
}  // namespace
}  // namespace eoc

extern "C" int eoc_selftest_nd_u16(unsigned long long* mismatches, void* stream) {
  if (!mismatches) EOC_FAIL(EOC_ERR_INVALID, "eoc_selftest_nd_u16: mismatches is NULL");
  eoc::k_selftest_nd_u16<<<eoc::sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(mismatches);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}
