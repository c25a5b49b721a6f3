// Kernel T: TMA-staged (R x W) column tiles in shared memory, for the reducers that need the column
// twice (variance, sd: numpy's two-pass algorithm, numpy/lib/_nanfunctions_impl.py:1850-1892).
//
// A CTA is persistent and walks tiles of W adjacent columns.  One elected thread issues a single
// cp.async.bulk.tensor (TMA) per box of <= 256 rows x W columns; the bytes land in one of `stages`
// shared-memory buffers and complete an mbarrier.  Every thread then owns CPT adjacent columns of the
// landed tile and runs BOTH passes out of shared memory (row r of column c sits at [r*W + c]: a warp
// reads consecutive words with one vector LDS per row, conflict-free), so the cube is read from HBM
// exactly once:
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
  int W;       // columns per tile (compile-time in the kernel; kept here for the host)
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

// pass 1 / pass 2 over one segment of a resident tile.  `row(j)` maps the j-th member to its row.
// CPT adjacent columns per thread are read with one vector LDS and reduced as independent chains.
template <typename In, typename A, uint32_t G, int W, int CPT, typename RowOf>
__device__ __forceinline__ void tile_segment(const In* col, int beg, int end, RowOf row, const ColumnParams& P,
                                             bool need_var, int64_t out_idx, int valid_cols, bool is_segmented) {
  ColStats<A, G> st[CPT];
#pragma unroll
  for (int k = 0; k < CPT; ++k) st[k].init();
  int j = beg;
  for (; j + 4 <= end; j += 4) {
    Vec<In, CPT> q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) q[u] = *reinterpret_cast<const Vec<In, CPT>*>(col + row(j + u) * W);
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < CPT; ++k) st[k].add((A)q[u].v[k]);
  }
  for (; j < end; ++j) {
    const Vec<In, CPT> q = *reinterpret_cast<const Vec<In, CPT>*>(col + row(j) * W);
#pragma unroll
    for (int k = 0; k < CPT; ++k) st[k].add((A)q.v[k]);
  }

  A sumsq[CPT];
#pragma unroll
  for (int k = 0; k < CPT; ++k) sumsq[k] = (A)0;
  if (need_var) {
    A avg[CPT];
    bool fast = true;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      avg[k] = mean_of(st[k].sum, st[k].n);
      fast = fast && sqdev_fast_ok(avg[k], st[k].n);
    }
    if (fast) {
      j = beg;
      for (; j + 4 <= end; j += 4) {
        Vec<In, CPT> q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) q[u] = *reinterpret_cast<const Vec<In, CPT>*>(col + row(j + u) * W);
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int k = 0; k < CPT; ++k) sumsq[k] = add_rn(sumsq[k], sqdev_fast((A)q[u].v[k], avg[k]));
      }
      for (; j < end; ++j) {
        const Vec<In, CPT> q = *reinterpret_cast<const Vec<In, CPT>*>(col + row(j) * W);
#pragma unroll
        for (int k = 0; k < CPT; ++k) sumsq[k] = add_rn(sumsq[k], sqdev_fast((A)q.v[k], avg[k]));
      }
    } else {  // a non-finite mean next to valid samples: the exact (masking) term
      for (j = beg; j < end; ++j) {
        const Vec<In, CPT> q = *reinterpret_cast<const Vec<In, CPT>*>(col + row(j) * W);
#pragma unroll
        for (int k = 0; k < CPT; ++k) sumsq[k] = add_rn(sumsq[k], sqdev((A)q.v[k], avg[k]));
      }
    }
  }
  const bool empty_nan = is_segmented && (end == beg);
#pragma unroll
  for (int k = 0; k < CPT; ++k)
    if (k < valid_cols) finalize_column<A, G>(st[k], end - beg, sumsq[k], P, out_idx + k, empty_nan);
}

// W columns per tile, CPT columns per thread: W / CPT threads per CTA.
template <typename In, typename A, uint32_t G, int W, int CPT>
__global__ void __launch_bounds__(W / CPT) k_tile(const __grid_constant__ CUtensorMap tmap, const TileParams T) {
  // buffers need 128-byte alignment for TMA; barriers follow the buffers.  (No pointer arithmetic through
  // integers here: the compiler must keep seeing a shared-memory address to emit LDS instead of generic LD.)
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  unsigned char* base = smem_dyn;
  uint64_t* full = reinterpret_cast<uint64_t*>(base + (size_t)T.stages * T.tile_bytes);
  const ColumnParams& P = T.C;
  const int tid = threadIdx.x;
  const int c = tid * CPT;
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

  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    for (int s = 0; s < T.stages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0)
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
      const int64_t left = P.inner - i;
      const int valid = left < CPT ? (int)left : CPT;
      const In* col = reinterpret_cast<const In*>(base + (size_t)s * T.tile_bytes) + c;
      if (!P.seg_offsets) {
        tile_segment<In, A, G, W, CPT>(col, 0, (int)P.R, [](int j) { return j; }, P, need_var,
                                       o * P.inner + i, valid, false);
      } else {
        for (int sg = 0; sg < P.nseg; ++sg) {
          const int beg = P.seg_offsets[sg], end = P.seg_offsets[sg + 1];
          const int64_t out_idx = (o * P.nseg + sg) * P.inner + i;
          if (P.members) {
            const int32_t* mem = P.members;
            tile_segment<In, A, G, W, CPT>(col, beg, end, [mem](int j) { return mem[j]; }, P, need_var, out_idx,
                                           valid, true);
          } else {
            tile_segment<In, A, G, W, CPT>(col, beg, end, [](int j) { return j; }, P, need_var, out_idx, valid,
                                           true);
          }
        }
      }
    }
    __syncthreads();  // every thread is done with buffer s
    if (tid == 0 && k + T.stages < n_my) issue(k + T.stages);
  }
}

}  // namespace eoc
