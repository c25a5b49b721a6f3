// Kernel B: few, very long, contiguous columns (inner == 1, stride_R == 1): the spatial reduce
// reduce_spatial / reduce_dimension(..., "x"|"y") of cubes/reduce.py:49-62 on a (t,b,y,x) cube, where
// outer = T*B and R = Y*X.  A CTA reduces one contiguous chunk of one row cooperatively (coalesced
// 128-bit loads, warp-shuffle combine), a second tiny kernel folds the per-chunk partials.
//
// numpy sums a contiguous axis pairwise in the input precision; here partial sums are kept in
// float64, which is at least as accurate, so results agree within the 1e-6 contract but are not
// bit-identical.  Variance uses shifted sums S1 = sum(x-K), S2 = sum((x-K)^2) in float64 with
// K = the first finite sample of the row (shared by all chunks, so partials simply add).
#pragma once

#include "column_kernel.cuh"

namespace eoc {

struct RowPartial {
  double sum, s1, s2, prod;
  long long n, nfin;
  double mn, mx;
};

__device__ __forceinline__ void rp_init(RowPartial& p) {
  p.sum = 0.0;
  p.s1 = 0.0;
  p.s2 = 0.0;
  p.prod = 1.0;
  p.n = 0;
  p.nfin = 0;
  p.mn = Lim<double>::nan();
  p.mx = Lim<double>::nan();
}
// Accumulator groups of the long-row kernels (compile-time, like the column kernel's): a mean needs one float64 add
// per sample, not the eight statistics -- the float64 pipe bounded aggregate_spatial at 0.16 of the HBM roof with all
// of them (bench r2).  The not-NaN count is always kept (ignore_nodata = 0 poisons on it).
constexpr uint32_t RG_SUM = 1u;   // sum
constexpr uint32_t RG_VAR = 2u;   // shifted sums s1, s2
constexpr uint32_t RG_MM = 4u;    // min, max
constexpr uint32_t RG_RARE = 8u;  // product, finite count
constexpr uint32_t RG_ALL = 15u;
inline uint32_t row_groups_for(uint32_t ops) {
  uint32_t g = 0;
  if (ops & ((1u << EOC_OP_SUM) | (1u << EOC_OP_MEAN))) g |= RG_SUM;
  if (ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) g |= RG_VAR;
  if (ops & ((1u << EOC_OP_MIN) | (1u << EOC_OP_MAX))) g |= RG_MM;
  if (ops & ((1u << EOC_OP_PRODUCT) | (1u << EOC_OP_COUNT))) g |= RG_RARE;
  // the instantiated combinations: sum | sum + variance | extrema | everything
  if ((g & ~RG_SUM) == 0) return RG_SUM;
  if ((g & ~(RG_SUM | RG_VAR)) == 0) return RG_SUM | RG_VAR;
  if (g == RG_MM) return RG_MM;
  return RG_ALL;
}
template <uint32_t RG = RG_ALL> __device__ __forceinline__ void rp_add(RowPartial& p, double x, double K) {
  const bool nan = x != x;
  p.n += nan ? 0 : 1;
  if constexpr ((RG & RG_SUM) != 0) p.sum += nan ? 0.0 : x;
  if constexpr ((RG & RG_VAR) != 0) {
    const double d = nan ? 0.0 : x - K;
    p.s1 += d;
    p.s2 += d * d;
  }
  if constexpr ((RG & RG_RARE) != 0) {
    p.prod *= nan ? 1.0 : x;
    p.nfin += (fabs(x) <= Lim<double>::maxv()) ? 1 : 0;
  }
  if constexpr ((RG & RG_MM) != 0) {
    p.mn = fmin(p.mn, x);
    p.mx = fmax(p.mx, x);
  }
}
__device__ __forceinline__ void rp_merge(RowPartial& a, const RowPartial& b) {
  a.sum += b.sum;
  a.s1 += b.s1;
  a.s2 += b.s2;
  a.prod *= b.prod;
  a.n += b.n;
  a.nfin += b.nfin;
  a.mn = fmin(a.mn, b.mn);
  a.mx = fmax(a.mx, b.mx);
}
__device__ __forceinline__ RowPartial rp_shfl_down(const RowPartial& p, int d) {
  RowPartial q;
  q.sum = __shfl_down_sync(0xffffffffu, p.sum, d);
  q.s1 = __shfl_down_sync(0xffffffffu, p.s1, d);
  q.s2 = __shfl_down_sync(0xffffffffu, p.s2, d);
  q.prod = __shfl_down_sync(0xffffffffu, p.prod, d);
  q.n = __shfl_down_sync(0xffffffffu, p.n, d);
  q.nfin = __shfl_down_sync(0xffffffffu, p.nfin, d);
  q.mn = __shfl_down_sync(0xffffffffu, p.mn, d);
  q.mx = __shfl_down_sync(0xffffffffu, p.mx, d);
  return q;
}

// The shift K of the shifted sums (variance without cancellation): the FIRST FINITE sample among the first 4096 of
// the row, found by the whole CTA (thread t looks at t, t + 256, ...; the smallest index wins) -- a pure function of
// the row, so every chunk of it shifts by the same K.  (Round 1 took row[0] and fell back to K = 0 when that was NaN
// or infinite -- a nodata corner -- which cancels catastrophically for |mean| / sd >~ 1e5: ADVICE r1.)
template <typename In> __device__ __forceinline__ double row_shift_cta(const In* row, int64_t R) {
  __shared__ int first_idx;
  if (threadIdx.x == 0) first_idx = 0x7fffffff;
  __syncthreads();
  const int lim = R < 4096 ? (int)R : 4096;
  for (int e = threadIdx.x; e < lim; e += blockDim.x) {
    if (fabs((double)row[e]) <= Lim<double>::maxv()) {
      atomicMin(&first_idx, e);
      break;
    }
  }
  __syncthreads();
  return first_idx == 0x7fffffff ? 0.0 : (double)row[first_idx];
}

constexpr int ROWS_THREADS = 256;

template <typename In, int VEC, uint32_t RG = RG_ALL>
__global__ void __launch_bounds__(ROWS_THREADS) k_rows_partial(const In* data, int64_t so, int64_t R,
                                                               int64_t chunk_len, int nchunks,
                                                               RowPartial* partials) {
  const int64_t o = blockIdx.y;
  const int c = blockIdx.x;
  const In* row = data + o * so;
  const double K = (RG & RG_VAR) ? row_shift_cta(row, R) : 0.0;
  const int64_t beg = (int64_t)c * chunk_len;
  const int64_t end = (beg + chunk_len < R) ? beg + chunk_len : R;
  RowPartial p;
  rp_init(p);
  for (int64_t e = beg + (int64_t)threadIdx.x * VEC; e < end; e += (int64_t)ROWS_THREADS * VEC) {
    if (e + VEC <= end) {
      const Vec<In, VEC> q = ld_vec<In, VEC>(row + e);
#pragma unroll
      for (int k = 0; k < VEC; ++k) rp_add<RG>(p, (double)q.v[k], K);
    } else {
      for (int64_t t = e; t < end; ++t) rp_add<RG>(p, (double)row[t], K);
    }
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const RowPartial q = rp_shfl_down(p, d);
    rp_merge(p, q);
  }
  __shared__ RowPartial warp_part[ROWS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_part[warp] = p;
  __syncthreads();
  if (threadIdx.x == 0) {
    RowPartial t = warp_part[0];
    for (int w = 1; w < ROWS_THREADS / 32; ++w) rp_merge(t, warp_part[w]);
    partials[o * nchunks + c] = t;
  }
}

// the results of one row from its merged partial (shared by the finishing kernel of the long-row path and the
// warp-per-row kernel)
template <typename A>
__device__ __forceinline__ void rows_finish(const RowPartial& t, int64_t o, int64_t R, uint32_t ops, int ignore_nodata,
                                            int ddof, const OutPtrs& outs) {
  const bool poisoned = !ignore_nodata && t.n < R;
  const A nanv = Lim<A>::nan();
  const double n = (double)t.n;
  if (ops & (1u << EOC_OP_SUM)) ((A*)outs.p[EOC_OP_SUM])[o] = poisoned ? nanv : (A)t.sum;
  if (ops & (1u << EOC_OP_MEAN)) ((A*)outs.p[EOC_OP_MEAN])[o] = poisoned ? nanv : (A)(t.sum / n);
  if (ops & (1u << EOC_OP_MIN)) ((A*)outs.p[EOC_OP_MIN])[o] = poisoned ? nanv : (A)t.mn;
  if (ops & (1u << EOC_OP_MAX)) ((A*)outs.p[EOC_OP_MAX])[o] = poisoned ? nanv : (A)t.mx;
  if (ops & (1u << EOC_OP_COUNT)) ((long long*)outs.p[EOC_OP_COUNT])[o] = t.nfin;
  if (ops & (1u << EOC_OP_NOT_NAN)) ((long long*)outs.p[EOC_OP_NOT_NAN])[o] = t.n;
  if (ops & (1u << EOC_OP_COUNT_ALL)) ((A*)outs.p[EOC_OP_COUNT_ALL])[o] = (A)R;
  if (ops & (1u << EOC_OP_PRODUCT)) ((A*)outs.p[EOC_OP_PRODUCT])[o] = poisoned ? nanv : (A)t.prod;
  if (ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) {
    const long long dof = t.n - ddof;
    double var = Lim<double>::nan();
    if (!poisoned && dof > 0) {
      const double m2 = t.s2 - t.s1 * t.s1 / n;
      var = (m2 > 0.0 ? m2 : (m2 == m2 ? 0.0 : m2)) / (double)dof;
    }
    if (ops & (1u << EOC_OP_VARIANCE)) ((A*)outs.p[EOC_OP_VARIANCE])[o] = (A)var;
    if (ops & (1u << EOC_OP_SD)) ((A*)outs.p[EOC_OP_SD])[o] = (A)sqrt(var);
  }
}

template <typename In, typename A>
__global__ void k_rows_final(const In* data, int64_t so, int64_t outer, int64_t R, int nchunks,
                             const RowPartial* partials, uint32_t ops, int ignore_nodata, int ddof,
                             OutPtrs outs) {
  const int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= outer) return;
  RowPartial t = partials[o * nchunks];
  for (int c = 1; c < nchunks; ++c) rp_merge(t, partials[o * nchunks + c]);
  rows_finish<A>(t, o, R, ops, ignore_nodata, ddof, outs);
}

// MANY short contiguous rows (reducing the last axis: x of a (t, bands, y, x) cube, t of a (y, x, t) cube): one WARP per
// row -- coalesced 128-bit loads, float64 partials, shuffle combine, no workspace -- where one thread per row (the column
// kernel with inner = 1) reads 4 KB-strided samples: mean over x of a (100, 4, 512, 1024) cube ran at 0.14 of the HBM roof.
template <typename In, typename A, int VEC, uint32_t RG>
__global__ void __launch_bounds__(256) k_rows_warp(const In* __restrict__ data, int64_t so, int64_t outer, int64_t R,
                                                   uint32_t ops, int ignore_nodata, int ddof, OutPtrs outs) {
  const int lane = threadIdx.x & 31;
  const int64_t o = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (o >= outer) return;
  const In* row = data + o * so;
  double K = 0.0;
  if constexpr ((RG & RG_VAR) != 0) {  // the first finite sample among the row's first 32
    const double v = lane < R ? (double)row[lane] : Lim<double>::nan();
    const unsigned m = __ballot_sync(0xffffffffu, fabs(v) <= Lim<double>::maxv());
    const double k0 = __shfl_sync(0xffffffffu, v, m ? __ffs(m) - 1 : 0);
    K = m ? k0 : 0.0;
  }
  RowPartial p;
  rp_init(p);
  const int64_t rv = R / VEC * VEC;
  for (int64_t e = (int64_t)lane * VEC; e < rv; e += 32 * VEC) {
    const Vec<In, VEC> q = ld_vec<In, VEC>(row + e);
#pragma unroll
    for (int k = 0; k < VEC; ++k) rp_add<RG>(p, (double)q.v[k], K);
  }
  for (int64_t e = rv + lane; e < R; e += 32) rp_add<RG>(p, (double)row[e], K);
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const RowPartial q = rp_shfl_down(p, d);
    rp_merge(p, q);
  }
  if (lane == 0) rows_finish<A>(p, o, R, ops, ignore_nodata, ddof, outs);
}

inline int rows_chunks(const eoc_view& in) {
  // aim at ~8 CTAs per SM over all rows, chunks of at least 16 Ki samples
  const int64_t want = (8LL * sm_count() + in.outer - 1) / in.outer;
  int64_t maxc = (in.R + 16383) / 16384;
  int64_t c = want < maxc ? want : maxc;
  if (c < 1) c = 1;
  if (c > 4096) c = 4096;
  return (int)c;
}

inline size_t rows_workspace_bytes(const eoc_view& in) {
  return sizeof(RowPartial) * (size_t)in.outer * (size_t)rows_chunks(in);
}

template <typename In>
int launch_rows_t(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                  RowPartial* partials, int nchunks, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  constexpr int VM = (sizeof(In) == 8) ? 2 : 4;
  int64_t chunk_len = (in.R + nchunks - 1) / nchunks;
  chunk_len = (chunk_len + 1023) / 1024 * 1024;  // keeps every chunk start vector-aligned
  const dim3 grid((unsigned)nchunks, (unsigned)in.outer);
  const bool al = ((uintptr_t)in.data % (sizeof(In) * VM) == 0) && (in.stride_outer % VM == 0);
#define EOC_ROWS_GO(RGV)                                                                                       \
  do {                                                                                                         \
    if (al)                                                                                                    \
      k_rows_partial<In, VM, RGV><<<grid, ROWS_THREADS, 0, s>>>((const In*)in.data, in.stride_outer, in.R,     \
                                                                chunk_len, nchunks, partials);                 \
    else                                                                                                       \
      k_rows_partial<In, 1, RGV><<<grid, ROWS_THREADS, 0, s>>>((const In*)in.data, in.stride_outer, in.R,      \
                                                               chunk_len, nchunks, partials);                  \
  } while (0)
  switch (row_groups_for(ops)) {
    case RG_SUM: EOC_ROWS_GO(RG_SUM); break;
    case RG_SUM | RG_VAR: EOC_ROWS_GO(RG_SUM | RG_VAR); break;
    case RG_MM: EOC_ROWS_GO(RG_MM); break;
    default: EOC_ROWS_GO(RG_ALL); break;
  }
#undef EOC_ROWS_GO
  EOC_LAUNCH_CHECK();
  const int threads = 128;
  k_rows_final<In, A><<<(unsigned)((in.outer + threads - 1) / threads), threads, 0, s>>>(
      (const In*)in.data, in.stride_outer, in.outer, in.R, nchunks, partials, ops, ignore_nodata, ddof,
      outs);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In>
int launch_rows_warp_t(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  constexpr int VM = (sizeof(In) == 8) ? 2 : 4;
  const int64_t blocks = (in.outer + 7) / 8;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const bool al = ((uintptr_t)in.data % (sizeof(In) * VM) == 0) && (in.stride_outer % VM == 0);
#define EOC_ROWSW_GO(RGV)                                                                                         \
  do {                                                                                                            \
    if (al)                                                                                                       \
      k_rows_warp<In, A, VM, RGV><<<(unsigned)blocks, 256, 0, s>>>((const In*)in.data, in.stride_outer, in.outer, \
                                                                  in.R, ops, ignore_nodata, ddof, outs);          \
    else                                                                                                          \
      k_rows_warp<In, A, 1, RGV><<<(unsigned)blocks, 256, 0, s>>>((const In*)in.data, in.stride_outer, in.outer,  \
                                                                 in.R, ops, ignore_nodata, ddof, outs);           \
  } while (0)
  switch (row_groups_for(ops)) {
    case RG_SUM: EOC_ROWSW_GO(RG_SUM); break;
    case RG_SUM | RG_VAR: EOC_ROWSW_GO(RG_SUM | RG_VAR); break;
    case RG_MM: EOC_ROWSW_GO(RG_MM); break;
    default: EOC_ROWSW_GO(RG_ALL); break;
  }
#undef EOC_ROWSW_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}
inline int launch_rows_warp(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof, const OutPtrs& outs,
                            cudaStream_t s) {
  switch (in.dtype) {
    case EOC_F32: return launch_rows_warp_t<float>(in, ops, ignore_nodata, ddof, outs, s);
    case EOC_F64: return launch_rows_warp_t<double>(in, ops, ignore_nodata, ddof, outs, s);
    case EOC_U8: return launch_rows_warp_t<uint8_t>(in, ops, ignore_nodata, ddof, outs, s);
    case EOC_I16: return launch_rows_warp_t<int16_t>(in, ops, ignore_nodata, ddof, outs, s);
    case EOC_U16: return launch_rows_warp_t<uint16_t>(in, ops, ignore_nodata, ddof, outs, s);
    case EOC_I32: return launch_rows_warp_t<int32_t>(in, ops, ignore_nodata, ddof, outs, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

inline int launch_rows(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof,
                       const OutPtrs& outs, void* workspace, size_t workspace_bytes, cudaStream_t s) {
  if (in.outer > 65535) EOC_FAIL(EOC_ERR_UNSUPPORTED, "row kernel: outer=%lld > 65535", (long long)in.outer);
  const int nchunks = rows_chunks(in);
  const size_t need = sizeof(RowPartial) * (size_t)in.outer * (size_t)nchunks;
  if (!workspace || workspace_bytes < need)
    EOC_FAIL(EOC_ERR_WORKSPACE, "row kernel needs %zu workspace bytes, got %zu", need, workspace_bytes);
  RowPartial* part = (RowPartial*)workspace;
  switch (in.dtype) {
    case EOC_F32: return launch_rows_t<float>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
    case EOC_F64: return launch_rows_t<double>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
    case EOC_U8: return launch_rows_t<uint8_t>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
    case EOC_I16: return launch_rows_t<int16_t>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
    case EOC_U16: return launch_rows_t<uint16_t>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
    case EOC_I32: return launch_rows_t<int32_t>(in, ops, ignore_nodata, ddof, outs, part, nchunks, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

}  // namespace eoc
