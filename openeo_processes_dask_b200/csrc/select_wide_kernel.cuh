// Kernel C-W: order statistics of LONG columns, 128 < R <= 512 (float32) -- a year of daily acquisitions
// (t = 365), BASELINE configs[3]'s t = 200.  One column no longer fits one thread's registers, so K = 2 or 4
// threads share it: thread (c, p) holds rows [128 p, 128 p + 128) of column c in registers, sorts them with the
// straight-line network of the narrow kernels (NaN last, mixed-pipe comparators), and the K sorted runs are
// merged by a bitonic merge whose cross-thread stages go through shared memory and whose in-thread stages
// are register networks again:
//   K = 2:  exchange with p^1 at mirrored rows (i <-> 127-i; lower part keeps the min, upper the max)
//           -> both halves are bitonic -> 7-stage in-register bitonic merge                      (sorted 256)
//   K = 4:  the same per pair, then across pairs: exchange with 3-p at mirrored rows, exchange with p^1 at equal
//           rows, in-register merge                                                              (sorted 512)
// A warp is 32 adjacent columns of ONE part, so every global load is a coalesced 128-byte row segment and every
// shared-memory access is conflict-free (row r of the exchange buffer is 64 or 32 consecutive floats).  The sorted
// column ends up parked in shared memory; part 0 finds n (first NaN, bisection) and emits median / quantiles /
// the sorted column exactly like the narrow kernels (emit_sorted).  Comparators per sample: 16 (K = 2), 21.5
// (K = 4) against 11 for the N = 100 network -- ALU-bound like them.  It replaces the per-thread bitonic sort
// in shared memory (k_select_smem: 54 GB/s at t = 200, 22 GB/s at t = 365) for float32 cubes.
// Generalised over the sample type: float64 columns keep 64 samples per thread (same register budget) and use
// K = 2, 4 or 8 threads for 64 < R <= 512; level L of the merge (runs of 2^(L-1) parts -> 2^L) is one mirrored
// exchange with part p ^ (2^L - 1), then equal-row exchanges with p ^ d for d = 2^(L-2) .. 1, then the in-register
// bitonic merge.
#pragma once

#include "select_launch.cuh"

namespace eoc {

// one stage of the in-register bitonic merge: compare-exchange at distance D
template <typename A, int NS, int D> __device__ __forceinline__ void bitonic_stage(A (&v)[NS], int one) {
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    if ((i & D) == 0) {
      // 3 of 5 comparators in the mixed-pipe form, like the sort networks (float32 only; see cmpx_k)
      if ((i % 5) < EOC_MIXQ_NUM * 5 / EOC_MIXQ_DEN)
        cmpx_k<true, true, A>(v[i], v[i + D], one);
      else
        cmpx_k<true, false, A>(v[i], v[i + D], one);
    }
  }
}

template <typename A, int NS> __device__ __forceinline__ void bitonic_merge_regs(A (&v)[NS], int one) {
  if constexpr (NS >= 128) bitonic_stage<A, NS, 64>(v, one);
  bitonic_stage<A, NS, 32>(v, one);
  __syncthreads();  // lockstep: keeps the CTA's warps inside one instruction-cache window
  bitonic_stage<A, NS, 16>(v, one);
  bitonic_stage<A, NS, 8>(v, one);
  __syncthreads();
  bitonic_stage<A, NS, 4>(v, one);
  bitonic_stage<A, NS, 2>(v, one);
  bitonic_stage<A, NS, 1>(v, one);
}

// cross-thread stage: my row j against the partner part's row (MIRROR ? NS-1-j : j); the lower part keeps the
// smaller sample, the upper part the larger one (NaN counts as largest)
template <bool MIRROR, typename A, int NS, int WC>
__device__ __forceinline__ void exchange(A (&v)[NS], A* X, int c, int p, int partner) {
#pragma unroll
  for (int j = 0; j < NS; ++j) X[(p * NS + j) * WC + c] = v[j];
  __syncthreads();
  const A* q = X + (partner * NS) * WC + c;
  if (p < partner) {
#pragma unroll
    for (int j = 0; j < NS; ++j) v[j] = min_(v[j], q[(MIRROR ? NS - 1 - j : j) * WC]);
  } else {
#pragma unroll
    for (int j = 0; j < NS; ++j) v[j] = max_nan_(v[j], q[(MIRROR ? NS - 1 - j : j) * WC]);
  }
  __syncthreads();  // everybody has read: the buffer may be overwritten
}

// 128-thread CTAs with 64 KB of exchange buffer for every K; float32: three of them per SM at 168 registers,
// float64: two at 255.  Measured at t = 365 (float32, K = 4): 3 x 128 threads 713 GB/s; one 256-thread CTA at 254
// registers 475; one 384-thread lockstep CTA at 168 registers 682 (t = 200, K = 2: 960 against 1435) -- the
// single big CTA shares the instruction fetch (ncu of the 3 x 128 form, profiles/r1i_median365_raw.csv:
// stall_no_instruction 3.6 per issue, ALU pipe 40 % busy) but has nothing to overlap its load phase with.
template <typename A> constexpr int wide_tpb() { return 128; }
template <typename In, typename A, int NS, int K>
__global__ void __launch_bounds__(wide_tpb<A>(), sizeof(A) == 8 ? 2 : 3)
k_select_wide(const In* __restrict__ data, const SelectParams P, const uint32_t sr_bytes) {
  constexpr int WC = wide_tpb<A>() / K;  // columns per CTA
  extern __shared__ __align__(16) unsigned char smem_raw[];
  A* X = reinterpret_cast<A*>(smem_raw);  // [K * NS][WC]
  const int t = threadIdx.x, c = t % WC, p = t / WC;
  const int64_t total = P.outer * P.inner;
  int64_t item = (int64_t)blockIdx.x * WC + c;
  const bool active = item < total;
  item = active ? item : total - 1;  // branch-free tail: redo the last column, skip the store
  const int64_t o = item / P.inner;
  const int64_t i = item - o * P.inner;
  const char* col = reinterpret_cast<const char*>(data + o * P.so + i);
  const int R = (int)P.R;

  A v[NS];
#pragma unroll
  for (int j = 0; j < NS; ++j) {
    const int row = p * NS + j;
    const In* src = reinterpret_cast<const In*>(col + (uint64_t)((uint32_t)row) * (uint64_t)sr_bytes);
    v[j] = (row < R) ? (A)__ldcs(src) : Lim<A>::nan();
  }
  sort_nan_last<NS, A, true>(v, (A)P.zero);  // canonical NaNs (float32), ascending, NaN last
  const int one = one_from_zero(P.zero);

#pragma unroll
  for (int L = 1; (1 << L) <= K; ++L) {
    exchange<true, A, NS, WC>(v, X, c, p, p ^ ((1 << L) - 1));
#pragma unroll
    for (int d = (1 << L) >> 2; d >= 1; d >>= 1) exchange<false, A, NS, WC>(v, X, c, p, p ^ d);
    bitonic_merge_regs<A, NS>(v, one);
  }

#pragma unroll
  for (int j = 0; j < NS; ++j) X[(p * NS + j) * WC + c] = v[j];
  __syncthreads();
  if (p == 0) {
    const A* mycol = X + c;
    const int n = first_nan_index(K * NS, [&](int r) { return mycol[r * WC]; });
    if (active) emit_sorted<A>(P, o, i, n, mycol, int(WC));
  }
}

template <typename In, typename A, int NS, int K>
int launch_wide_k(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int WC = wide_tpb<A>() / K;
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + WC - 1) / WC;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const size_t smem = (size_t)K * NS * WC * sizeof(A);
  auto kern = k_select_wide<In, A, NS, K>;
  EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, wide_tpb<A>(), smem, s>>>((const In*)in.data, P, (uint32_t)(in.stride_R * (int64_t)sizeof(In)));
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// float32 networks (float32 and 8/16-bit integer samples): 128 samples per thread
template <typename In> int launch_wide_t(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return in.R <= 256 ? launch_wide_k<In, float, 128, 2>(in, P, s) : launch_wide_k<In, float, 128, 4>(in, P, s);
}

}  // namespace eoc
