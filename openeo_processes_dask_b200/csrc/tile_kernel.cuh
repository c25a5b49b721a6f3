// Kernel T: TMA-staged (R x W) column tiles in shared memory, for the reducers that need the column
// twice (variance, sd: numpy's two-pass algorithm, numpy/lib/_nanfunctions_impl.py:1850-1892).
//
// A CTA is persistent and walks tiles of W adjacent columns.  One elected thread issues a single
// cp.async.bulk.tensor (TMA) per box of <= 256 rows x W columns; the bytes land in one of `stages`
// shared-memory buffers and complete an mbarrier.  Every thread then owns one column of the landed
// tile and runs BOTH passes out of shared memory (row r of column c sits at [r*W + c]: a warp reads
// 32 consecutive words, conflict-free), so the cube is read from HBM exactly once:
//   algorithmic bytes per column = R*sizeof(In) + sizeof(out) per selected reducer.
// While tile k is being reduced, tiles k+1 .. k+stages-1 are in flight; the buffer of tile k is
// re-armed for tile k+stages as soon as the CTA has passed it (__syncthreads).
// The arithmetic is ColStats / sqdev / finish_column of column_math.cuh, i.e. bit-identical to the
// column kernel and to the host model the CPU tests pin against numpy.
// A segmented reduce (aggregate_temporal) reads its member rows from the same resident tile.
#pragma once

#include <cuda.h>

#include "column_kernel.cuh"

namespace eoc {

struct TileParams {
  ColumnParams C;
  int W;       // columns per tile == threads per CTA
  int RB;      // rows per TMA box (<= 256)
  int nbox;    // boxes per tile: nbox * RB >= R
  int stages;  // shared-memory buffers
  uint32_t tile_bytes;  // nbox * RB * W * sizeof(In), multiple of 128
  int64_t tiles_per_outer, num_tiles;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// one box {W, RB, 1} of the 3-D tensor {inner, R, outer} -> shared memory; streaming data: evict-first in L2
__device__ __forceinline__ void tma_load_box(void* dst, const CUtensorMap* map, uint64_t* bar, int32_t c0,
                                             int32_t c1, int32_t c2) {
  constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(kEvictFirst)
      : "memory");
}

constexpr int TILE_MAX_STAGES = 4;

template <typename In, typename A, uint32_t G>
__global__ void __launch_bounds__(128) k_tile(const __grid_constant__ CUtensorMap tmap, const TileParams T) {
  extern __shared__ unsigned char smem_dyn[];
  // buffers need 128-byte alignment for TMA; barriers follow the buffers
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 127) & ~uintptr_t(127));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + (size_t)T.stages * T.tile_bytes);
  const ColumnParams& P = T.C;
  const int W = T.W, c = threadIdx.x;
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < T.num_tiles ? (T.num_tiles - first + step - 1) / step : 0;

  auto issue = [&](int64_t k) {  // thread 0 only
    const int s = (int)(k % T.stages);
    const int64_t t = first + k * step;
    const int64_t o = t / T.tiles_per_outer;
    const int64_t i0 = (t - o * T.tiles_per_outer) * W;
    mbar_expect_tx(&full[s], T.tile_bytes);
    unsigned char* dst = base + (size_t)s * T.tile_bytes;
    const uint32_t box_bytes = T.tile_bytes / T.nbox;
    for (int b = 0; b < T.nbox; ++b)
      tma_load_box(dst + (size_t)b * box_bytes, &tmap, &full[s], (int32_t)i0, b * T.RB, (int32_t)o);
  };

  if (c == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    for (int s = 0; s < T.stages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (c == 0)
    for (int64_t k = 0; k < T.stages && k < n_my; ++k) issue(k);

  const bool need_var = P.ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD));
  for (int64_t k = 0; k < n_my; ++k) {
    const int s = (int)(k % T.stages);
    const uint32_t parity = (uint32_t)((k / T.stages) & 1);
    const int64_t t = first + k * step;
    const int64_t o = t / T.tiles_per_outer;
    const int64_t i = (t - o * T.tiles_per_outer) * W + c;
    mbar_wait(&full[s], parity);
    if (i < P.inner) {
      const In* col = reinterpret_cast<const In*>(base + (size_t)s * T.tile_bytes) + c;
      for (int sg = 0; sg < P.nseg; ++sg) {
        const int beg = P.seg_offsets ? P.seg_offsets[sg] : 0;
        const int end = P.seg_offsets ? P.seg_offsets[sg + 1] : (int)P.R;
        ColStats<A, G> st;
        st.init();
        if (P.members) {
          for (int j = beg; j < end; ++j) st.add((A)col[P.members[j] * W]);
        } else {
          int j = beg;
          for (; j + 4 <= end; j += 4) {
            const A x0 = (A)col[(j + 0) * W], x1 = (A)col[(j + 1) * W], x2 = (A)col[(j + 2) * W],
                    x3 = (A)col[(j + 3) * W];
            st.add(x0);
            st.add(x1);
            st.add(x2);
            st.add(x3);
          }
          for (; j < end; ++j) st.add((A)col[j * W]);
        }
        A sumsq = (A)(-0.0);
        if (need_var) {
          const A avg = mean_of(st.sum, st.n);
          if (P.members) {
            for (int j = beg; j < end; ++j) sumsq = add_rn(sumsq, sqdev((A)col[P.members[j] * W], avg));
          } else {
            int j = beg;
            for (; j + 4 <= end; j += 4) {
              const A q0 = sqdev((A)col[(j + 0) * W], avg), q1 = sqdev((A)col[(j + 1) * W], avg),
                      q2 = sqdev((A)col[(j + 2) * W], avg), q3 = sqdev((A)col[(j + 3) * W], avg);
              sumsq = add_rn(add_rn(add_rn(add_rn(sumsq, q0), q1), q2), q3);
            }
            for (; j < end; ++j) sumsq = add_rn(sumsq, sqdev((A)col[j * W], avg));
          }
        }
        const bool empty_nan = (P.seg_offsets != nullptr) && (end == beg);
        finalize_column<A, G>(st, end - beg, sumsq, P, (o * P.nseg + sg) * P.inner + i, empty_nan);
      }
    }
    __syncthreads();  // every thread is done with buffer s
    if (c == 0 && k + T.stages < n_my) issue(k + T.stages);
  }
}

}  // namespace eoc
