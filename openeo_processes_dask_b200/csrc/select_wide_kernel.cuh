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
#pragma once

#include "select_launch.cuh"

namespace eoc {

constexpr int WIDE_NS = 128;  // samples per thread
// columns per CTA: 128-thread CTAs with 64 KB of exchange buffer for both K, three of them per SM (a 256-thread
// CTA for K = 4 needs 254 registers per thread and runs alone on its SM: 475 GB/s at t = 365 instead of 713)
template <int K> constexpr int wide_wc() { return 128 / K; }

// one stage of the in-register bitonic merge: compare-exchange at distance D
template <int D> __device__ __forceinline__ void bitonic_stage(float (&v)[WIDE_NS], int one) {
#pragma unroll
  for (int i = 0; i < WIDE_NS; ++i) {
    if ((i & D) == 0) {
      constexpr int dummy = 0;
      (void)dummy;
      // every 3rd-of-5 comparator in the mixed-pipe form, like the sort networks (index: position in the stage)
      if ((i % 5) < EOC_MIXQ_NUM * 5 / EOC_MIXQ_DEN)
        cmpx_k<true, true, float>(v[i], v[i + D], one);
      else
        cmpx_k<true, false, float>(v[i], v[i + D], one);
    }
  }
}

__device__ __forceinline__ void bitonic_merge_regs(float (&v)[WIDE_NS], int one) {
  bitonic_stage<64>(v, one);
  bitonic_stage<32>(v, one);
  __syncthreads();  // lockstep: keeps the CTA's warps inside one instruction-cache window
  bitonic_stage<16>(v, one);
  bitonic_stage<8>(v, one);
  __syncthreads();
  bitonic_stage<4>(v, one);
  bitonic_stage<2>(v, one);
  bitonic_stage<1>(v, one);
}

// cross-thread stage: my row j against the partner part's row (MIRROR ? 127 - j : j); the lower part keeps the
// smaller sample, the upper part the larger one (NaN counts as largest)
template <bool MIRROR, int WIDE_WC>
__device__ __forceinline__ void exchange(float (&v)[WIDE_NS], float* X, int c, int p, int partner) {
#pragma unroll
  for (int j = 0; j < WIDE_NS; ++j) X[(p * WIDE_NS + j) * WIDE_WC + c] = v[j];
  __syncthreads();
  const float* q = X + (partner * WIDE_NS) * WIDE_WC + c;
  if (p < partner) {
#pragma unroll
    for (int j = 0; j < WIDE_NS; ++j) v[j] = min_(v[j], q[(MIRROR ? WIDE_NS - 1 - j : j) * WIDE_WC]);
  } else {
#pragma unroll
    for (int j = 0; j < WIDE_NS; ++j) v[j] = max_nan_(v[j], q[(MIRROR ? WIDE_NS - 1 - j : j) * WIDE_WC]);
  }
  __syncthreads();  // everybody has read: the buffer may be overwritten
}

template <typename In, int K>
__global__ void __launch_bounds__(128, 3)
k_select_wide(const In* __restrict__ data, const SelectParams P, const uint32_t sr_bytes) {
  constexpr int WIDE_WC = wide_wc<K>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* X = reinterpret_cast<float*>(smem_raw);  // [K * 128][WIDE_WC]
  const int t = threadIdx.x, c = t % WIDE_WC, p = t / WIDE_WC;
  const int64_t total = P.outer * P.inner;
  int64_t item = (int64_t)blockIdx.x * WIDE_WC + c;
  const bool active = item < total;
  item = active ? item : total - 1;  // branch-free tail: redo the last column, skip the store
  const int64_t o = item / P.inner;
  const int64_t i = item - o * P.inner;
  const char* col = reinterpret_cast<const char*>(data + o * P.so + i);
  const int R = (int)P.R;

  float v[WIDE_NS];
#pragma unroll
  for (int j = 0; j < WIDE_NS; ++j) {
    const int row = p * WIDE_NS + j;
    const In* src = reinterpret_cast<const In*>(col + (uint64_t)((uint32_t)row) * (uint64_t)sr_bytes);
    v[j] = (row < R) ? (float)__ldcs(src) : Lim<float>::nan();
  }
  sort_nan_last<WIDE_NS, float, true>(v, P.zero);  // canonical NaNs, ascending, NaN last
  const int one = one_from_zero(P.zero);

  exchange<true, WIDE_WC>(v, X, c, p, p ^ 1);
  bitonic_merge_regs(v, one);
  if constexpr (K == 4) {
    exchange<true, WIDE_WC>(v, X, c, p, 3 - p);
    exchange<false, WIDE_WC>(v, X, c, p, p ^ 1);
    bitonic_merge_regs(v, one);
  }

#pragma unroll
  for (int j = 0; j < WIDE_NS; ++j) X[(p * WIDE_NS + j) * WIDE_WC + c] = v[j];
  __syncthreads();
  if (p == 0) {
    const float* mycol = X + c;
    const int n = first_nan_index(K * WIDE_NS, [&](int r) { return mycol[r * WIDE_WC]; });
    if (active) emit_sorted<float>(P, o, i, n, mycol, int(WIDE_WC));
  }
}

template <typename In, int K> int launch_wide_k(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int WIDE_WC = wide_wc<K>();
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + WIDE_WC - 1) / WIDE_WC;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const size_t smem = (size_t)K * WIDE_NS * WIDE_WC * sizeof(float);
  auto kern = k_select_wide<In, K>;
  EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, WIDE_WC * K, smem, s>>>((const In*)in.data, P, (uint32_t)(in.stride_R * (int64_t)sizeof(In)));
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In> int launch_wide_t(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return in.R <= 2 * WIDE_NS ? launch_wide_k<In, 2>(in, P, s) : launch_wide_k<In, 4>(in, P, s);
}

}  // namespace eoc
