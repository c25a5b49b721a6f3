// Launch template of k_select_tile, shared by select_tile.cu (plain) and select_tile_fused.cu (fused).
#pragma once

#include <cstdlib>

#include "select_tile_kernel.cuh"
#include "tma_host.cuh"

namespace eoc {

template <int N, bool QUANT, int FUSED, int W = SELT_W, bool MASKED = false>
int launch_select_tile_n(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk = nullptr) {
  CUtensorMap map;
  if (int rc = tma_make_map(in, W < 256 ? W : 256, (int)in.R, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  T.tile_bytes = (uint32_t)(in.R * W * sizeof(float));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = 0;
  if (const char* e = getenv("EOC_SELT_STAGGER")) T.stagger_ns = (uint32_t)atoi(e);
  T.mask = MASKED ? (const uint8_t*)mk->data : nullptr;
  T.mso = MASKED ? mk->stride_outer : 0;
  T.msr = MASKED ? mk->stride_R : 0;
  T.mf_mul = 0xFFC00000u;
  const size_t smem = (size_t)(QUANT ? N : in.R) * W * sizeof(float) + (MASKED ? (size_t)in.R * W : 0);
  int per_sm = 0;
#define EOC_SELT_GO(FULLV, LOCKV)                                                                       \
  do {                                                                                                  \
    auto kern = k_select_tile<N, QUANT, FULLV, FUSED, LOCKV, W, MASKED>;                                \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W, smem));                    \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W, smem, s>>>(map, T);                                                       \
  } while (0)
  if constexpr (FUSED < 2 && (W == SELT_W || MASKED)) {
    if (in.R == N)
      EOC_SELT_GO(true, true);
    else
      EOC_SELT_GO(false, true);
  } else {
    EOC_SELT_GO(false, true);
  }
#undef EOC_SELT_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <bool QUANT, int FUSED>
int launch_select_tile_q(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 64) return launch_select_tile_n<64, QUANT, FUSED>(in, P, s);
  if (in.R <= 100) return launch_select_tile_n<100, QUANT, FUSED>(in, P, s);
  return launch_select_tile_n<128, QUANT, FUSED>(in, P, s);
}

// Which calls take the TMA-staged kernel: float32, 32 < R <= 128, TMA-addressable, enough columns to keep a
// persistent grid busy.  Measured on B200 (T=100, 30 % NaN, GB/s input), TMA-staged 256-column tiles vs
// per-thread loads: quantiles 2684 vs 2607, median 3556 vs 3485, quantiles+sd 2122 vs 1858, median+mean
// 3017 vs 2817.  EOC_SELECT_TMA=0 disables it, =1 forces it for small cubes too (tests / tuning).
inline bool select_tile_applies(const eoc_view& in, bool quant, bool fused) {
  (void)quant;
  (void)fused;
  const char* e = getenv("EOC_SELECT_TMA");
  const int env = (!e || !e[0]) ? -1 : (e[0] == '0' ? 0 : 1);
  if (env == 0) return false;
  if (in.dtype != EOC_F32 || in.R <= 32 || in.R > 128) return false;
  if (!tma_view_ok(in)) return false;
  if (env == 1) return true;
  return in.outer * in.inner >= 65536 && tma_tiles_fill(in, 128);
}

// select_tile.cu / select_tile_fused.cu / select_tile_masked.cu
int launch_select_tile(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s);
int launch_select_tile_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, bool quant, cudaStream_t s);
int launch_select_tile_fused(const eoc_view& in, const SelectParams& P, bool quant, int level, cudaStream_t s);

}  // namespace eoc
