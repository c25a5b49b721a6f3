// Kernel C: per-column order statistics (median, linear-method quantiles) over R.
//
// k_select_regs   R <= N (N in {8,...,128}), the whole time series of a pixel lives in registers
//                 and is ordered by a straight-line Batcher network (generated/sort_networks.inc).
//                 * median only: NaN samples (and the N-R unused slots) are replaced ALTERNATELY by
//                   -inf / +inf.  With P pads in total, ceil(P/2) sit below and floor(P/2) above
//                   the n valid samples, so the median of the valid samples is at the fixed wires
//                   N/2-1, N/2 (n even) or N/2 (n odd): nan-median becomes a fixed-rank selection
//                   and the compiler prunes every comparator outside the cone of those two wires.
//                 * quantiles: NaN -> +inf, full sort, the sorted column is parked in shared
//                   memory (thread t owns bank t%32: conflict-free for any index) and the
//                   data-dependent ranks floor((n-1)q), +1 are fetched from there.  The full
//                   network for N = 100 is ~50 KB of straight-line code, more than the
//                   instruction cache holds: with warps drifting apart the kernel starved on
//                   instruction fetch (ncu r1: stall_no_instruction 4.3 per issue).  The quantile
//                   variant therefore runs 256-thread CTAs that re-converge at a CTA barrier every
//                   ~64 comparators (EOC_NET_SYNC), so the 8 warps of a CTA walk the code inside
//                   one cache window; two CTAs per SM alternate between their load and sort phases.
// k_select_smem   generic fallback for any dtype and R <= 512: per-thread bitonic sort of the
//                 column in shared memory (same conflict-free layout), runtime loops.
// Algorithmic bytes per column: R*sizeof(In) in, sizeof(A) (+ nq*sizeof(Q)) out.
#pragma once

#include "column_kernel.cuh"

namespace eoc {

struct SelectParams {
  int64_t outer, R, inner, so, sr;
  int ignore_nodata;
  void* median_out;  // A [outer, inner] or null
  int median_f64;    // 8/16-bit integer cubes run the float32 networks (exact) but report float64 like numpy
  int lerp_wrap16;   // int16 cubes: numpy's quantile interpolation forms b - a in int16 (wraps), see quantile_lerp
  int nq;
  double probs[EOC_SELECT_MAX_Q];
  void* quant_out;  // [outer, nq, inner]
  int quant_is_f64;
  float zero;  // 0.0f the compiler cannot see (keeps the pad flip on the FMA pipe, column_math.cuh)
  // eoc_sort: the whole sorted column, dense [outer, R, inner] in A (null: not wanted)
  void* sort_out;
  int sort_desc, sort_nan_first;
  ColumnParams fused;  // eoc_select_fused: streaming reducers of the same column (ops == 0: none)
};

// Streaming reducers of a register-resident column (the FUSED variants of the selection kernels), run BEFORE
// the sort.  Same arithmetic and order as the column kernel: sequential sum, two-pass variance.  Slots past R
// hold NaN and drop out of every statistic.  Almost all of it issues on the FMA pipe, which the ALU-bound
// sort leaves idle.  FUSED == 1: sum, mean, variance, sd, not-NaN count; FUSED == 2 adds min, max, product
// and the finite count.  The split keeps the common variant's straight-line code small: these kernels starve
// on instruction fetch (ncu r1d: stall_no_instruction 44 % of samples with every group compiled in).
// `reload(j)` re-reads sample j of the column from global memory: only the exact pass-2 term of a column
// whose mean is not finite takes it, as a rolled loop, so that rare path costs no unrolled code either.
template <int N, typename A, int FUSED, typename Reload>
__device__ __forceinline__ void fused_column_stats(const A* v, const ColumnParams& C, int64_t out_idx, bool active,
                                                   Reload reload) {
  ColStats<A, G_MINMAX | G_RARE> st;
  st.init();
  A cnt = (A)0;
#pragma unroll
  for (int j = 0; j < N; ++j) sum_step_fcount(v[j], st.sum, cnt);
  st.n = (int)cnt;
  if constexpr (FUSED >= 2) {
    if (C.ops & ((1u << EOC_OP_MIN) | (1u << EOC_OP_MAX))) {
#pragma unroll
      for (int j = 0; j < N; ++j) {
        st.mn = min_(st.mn, v[j]);
        st.mx = max_(st.mx, v[j]);
      }
    }
    if (C.ops & ((1u << EOC_OP_PRODUCT) | (1u << EOC_OP_COUNT))) {
#pragma unroll
      for (int j = 0; j < N; ++j) rare_step(v[j], st.prod, st.nfin);
    }
  }
  A sumsq = (A)0;
  if (C.ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) {
    const A avg = mean_of(st.sum, st.n);
#pragma unroll
    for (int j = 0; j < N; ++j) sumsq = add_rn(sumsq, sqdev_fast(v[j], avg));
    if (!sqdev_fast_ok(avg, st.n)) {
      sumsq = (A)0;
      for (int j = 0; j < (int)C.R; ++j) sumsq = add_rn(sumsq, sqdev(reload(j), avg));
    }
  }
  if (active) finalize_column<A, G_MINMAX | G_RARE>(st, (int)C.R, sumsq, C, out_idx, false);
}

// which fused variant an ops mask needs
inline int fused_level(uint32_t ops) {
  if (!ops) return 0;
  const uint32_t wide = (1u << EOC_OP_MIN) | (1u << EOC_OP_MAX) | (1u << EOC_OP_PRODUCT) | (1u << EOC_OP_COUNT);
  return (ops & wide) ? 2 : 1;
}

template <typename A> __device__ __forceinline__ void store_median(const SelectParams& P, int64_t idx, A m) {
  if (sizeof(A) == 4 && P.median_f64)
    ((double*)P.median_out)[idx] = (double)m;
  else
    ((A*)P.median_out)[idx] = m;
}

template <typename A>
__device__ __forceinline__ void store_quant(const SelectParams& P, int64_t o, int q, int64_t i, double v) {
  const int64_t idx = (o * P.nq + q) * P.inner + i;
  if (P.quant_is_f64)
    ((double*)P.quant_out)[idx] = v;
  else
    ((float*)P.quant_out)[idx] = (float)v;
}

// sort (arrays.py:560-584): the n valid samples of a parked column ascending (np.sort) or descending
// (-np.sort(-data)), the R - n NaNs behind them (nodata = True) or in front (nodata = False).  Row j of the
// output is written by every thread of the warp for its own column: coalesced.  Kept out of line so the
// median / quantile variants of the kernels keep their register allocation.
template <typename A>
__device__ __noinline__ void write_sorted(const A* col, int stride, int n, int R, int64_t inner, int desc,
                                          int nan_first, A* dst) {
  const int lead = nan_first ? R - n : 0;
  for (int j = 0; j < R; ++j) {
    const int k = j - lead;
    A y = Lim<A>::nan();
    if (k >= 0 && k < n) y = col[(desc ? n - 1 - k : k) * stride];
    dst[(int64_t)j * inner] = y;
  }
}

// a sorted column parked at col[k * stride]
template <typename A> struct ParkedColumn {
  const A* col;
  int stride;
  EOC_HD A operator()(int k) const { return col[k * stride]; }
};

// results from a parked sorted column whose first n entries are the valid samples
template <typename A>
__device__ __forceinline__ void emit_sorted(const SelectParams& P, int64_t o, int64_t i, int n, const A* col,
                                            int stride) {
  ParkedColumn<A> fetch;
  fetch.col = col;
  fetch.stride = stride;
  const bool poisoned = (!P.ignore_nodata && n < P.R) || n == 0;
  if (P.median_out) {
    A m = Lim<A>::nan();
    if (!poisoned) m = median_of_pair<A>(fetch((n - 1) >> 1), fetch(n >> 1));
    store_median<A>(P, o * P.inner + i, m);
  }
  for (int q = 0; q < P.nq; ++q) {
    double r = Lim<double>::nan();
    if (!poisoned) r = quantile_linear<A>(n, P.probs[q], fetch, P.lerp_wrap16 != 0);
    store_quant<A>(P, o, q, i, r);
  }
  if (P.sort_out)
    write_sorted<A>(col, stride, n, (int)P.R, P.inner, P.sort_desc, P.sort_nan_first,
                    (A*)P.sort_out + o * P.R * P.inner + i);
}

// registers: n32 = N samples in 32-bit registers + ~28 of bookkeeping
constexpr int select_min_blocks(int n32, bool wide) {
  return n32 <= 32 ? 8 : (n32 <= 100 ? 4 : (wide ? 2 : 3));
}

// FULL: R == N, no per-sample bounds predicate.  The address of sample j is col + j*sr_bytes with a
// 32-bit stride in bytes (one IMAD.WIDE.U32 on the FMA pipe per load; the ALU pipe is the bottleneck
// of this kernel and stays free of address arithmetic); strides >= 4 GiB take the shared-memory sorter.
// threads per CTA: 256 for the float32 quantile variant (lockstep CTAs), 128 otherwise
template <typename A, bool QUANT> constexpr int select_tpb() { return 128; }
template <typename A, bool QUANT> constexpr bool select_lockstep() { return QUANT && sizeof(A) == 4; }
// The fused variants keep ~10 more values live than the sort alone (sum, n, mean, sumsq, extrema); at the
// register cap of the plain kernel that spilled ~130 words per thread, so they run one CTA per SM fewer.
#ifndef EOC_FUSED_RELAX
#define EOC_FUSED_RELAX 0
#endif
template <typename A, int N, bool QUANT, int FUSED = 0> constexpr int select_blocks() {
  constexpr int b = select_min_blocks(N * (int)(sizeof(A) / 4), sizeof(A) == 8) * 128 / select_tpb<A, QUANT>();
  constexpr int f = (FUSED != 0 && b > 1 && N > 16 && EOC_FUSED_RELAX) ? b - 1 : b;
  return f < 1 ? 1 : f;
}

template <typename In, typename A, int N, bool QUANT, bool FULL, int FUSED = 0>
__global__ void __launch_bounds__((select_tpb<A, QUANT>()), (select_blocks<A, N, QUANT, FUSED>()))
k_select_regs(const In* __restrict__ data, const SelectParams P, const uint32_t sr_bytes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int64_t total = P.outer * P.inner;
  int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = item < total;
  // threads past the end redo the last column (branch-free network, loads stay in bounds) and skip the store
  item = active ? item : total - 1;
  const int64_t o = item / P.inner;
  const int64_t i = item - o * P.inner;
  const char* col = reinterpret_cast<const char*>(data + o * P.so + i);
  const int R = (int)P.R;

  A v[N];
#pragma unroll
  for (int j = 0; j < N; ++j) {
    const In* p = reinterpret_cast<const In*>(col + (uint64_t)((uint32_t)j) * (uint64_t)sr_bytes);
    if constexpr (FULL)
      v[j] = (A)__ldcs(p);
    else
      v[j] = (j < R) ? (A)__ldcs(p) : Lim<A>::nan();
  }

  if constexpr (FUSED != 0)
    fused_column_stats<N, A, FUSED>(v, P.fused, o * P.inner + i, active, [&](int j) {
      return (A)__ldg(reinterpret_cast<const In*>(col + (uint64_t)((uint32_t)j) * (uint64_t)sr_bytes));
    });

  if constexpr (!QUANT) {
    const A m = nanmedian_network<N, A>(v, R, P.ignore_nodata != 0, (A)P.zero);
    if (active) store_median<A>(P, o * P.inner + i, m);
  } else {
    A* tile = reinterpret_cast<A*>(smem_raw);  // [N][blockDim.x]
    const int W = blockDim.x, t = threadIdx.x;
    int n;
    if constexpr (sizeof(A) == 4) {
      // float32: NaN sorts last natively (cmpx<NANHI>), no canonicalisation pass; n by bisection below
      sort_nan_last<N, A, select_lockstep<A, QUANT>()>(v, (A)P.zero);
      n = 0;
    } else {
      n = sort_valid_first<N, A, select_lockstep<A, QUANT>()>(v);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) tile[j * W + t] = v[j];
    // each thread reads back only its own column: no barrier needed
    if constexpr (sizeof(A) == 4) n = first_nan_index(N, [&](int k) { return tile[k * W + t]; });
    if (active) emit_sorted<A>(P, o, i, n, tile + t, W);
  }
}

template <typename In, typename A>
__global__ void k_select_smem(const In* __restrict__ data, const SelectParams P, int npad) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  A* tile = reinterpret_cast<A*>(smem_raw);  // [npad][blockDim.x]
  const int W = blockDim.x, t = threadIdx.x;
  const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= P.outer * P.inner) return;  // columns are independent: no block-wide barrier below
  const int64_t o = item / P.inner;
  const int64_t i = item - o * P.inner;
  const In* col = data + o * P.so + i;
  const int R = (int)P.R;
  int n = 0;
  for (int j = 0; j < npad; ++j) {
    A x = Lim<A>::inf();
    if (j < R) {
      x = (A)__ldcs(col + (int64_t)j * P.sr);
      const bool nan = is_nan(x);
      n += nan ? 0 : 1;
      x = nan ? Lim<A>::inf() : x;
    }
    tile[j * W + t] = x;
  }
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int a = 0; a < npad; ++a) {
        const int b = a ^ j;
        if (b > a) {
          const A xa = tile[a * W + t], xb = tile[b * W + t];
          const bool asc = (a & k) == 0;
          const A lo = min_(xa, xb), hi = max_(xa, xb);
          tile[a * W + t] = asc ? lo : hi;
          tile[b * W + t] = asc ? hi : lo;
        }
      }
    }
  }
  emit_sorted<A>(P, o, i, n, tile + t, W);
}

}  // namespace eoc
