// Launch template of k_select_split, shared by select_split.cu (plain), select_split_fused.cu and
// select_split_masked.cu (separate translation units: ptxas works on them in parallel).
#pragma once

#include <cstdlib>

#include "select_split_kernel.cuh"
#include "tma_host.cuh"

namespace eoc {

inline int split_env(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && e[0]) ? atoi(e) : dflt;
}

template <int NH, int FUSED, bool MASKED, int W, bool PAR = false>
int launch_select_split_w(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk) {
  constexpr int BW = W < 256 ? W : 256;
  constexpr int NT = PAR ? 2 * W : W;
  constexpr int MINB = NT >= 256 ? 2 : 4;
  CUtensorMap map;
  if (int rc = tma_make_map(in, BW, (int)in.R, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  T.tile_bytes = (uint32_t)(in.R * W * sizeof(float));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = (uint32_t)split_env("EOC_SELT_STAGGER", 0);
  T.mask = MASKED ? (const uint8_t*)mk->data : nullptr;
  T.mso = MASKED ? mk->stride_outer : 0;
  T.msr = MASKED ? mk->stride_R : 0;
  T.mf_mul = 0xFFC00000u;
  const size_t smem = (size_t)2 * NH * W * sizeof(float) + (MASKED ? (size_t)in.R * W : 0);
  int per_sm = 0;
#define EOC_SPLIT_GO(FULLV)                                                                             \
  do {                                                                                                  \
    auto kern = k_select_split<NH, FULLV, FUSED, MASKED, W, MINB, false, 2, PAR>;                       \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem));                   \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, NT, smem, s>>>(map, T);                                                      \
  } while (0)
  if (in.R == 2 * NH)
    EOC_SPLIT_GO(true);
  else
    EOC_SPLIT_GO(false);
#undef EOC_SPLIT_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// Tile width = threads per CTA.  Measured on B200 (T = 100, 30 % NaN, 1024-row C2 strip, input GB/s), 128 / 256
// columns: median 4118 / 3974, quantiles 3330 / 3193, quantiles + sd 2430 / 2367; CTA barriers inside the
// network (the lockstep of k_select_tile) cost 10-13 % here -- the body is small enough for the instruction cache.
template <int NH, int FUSED, bool MASKED>
int launch_select_split_n(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk = nullptr) {
  if constexpr (!MASKED) {
    if (split_env("EOC_SPLIT_W", 128) == 256) return launch_select_split_w<NH, FUSED, MASKED, 256>(in, P, s, mk);
  }
  // (one thread per (column, run) -- the layout of the long-column variants, select_split_wide.cuh -- measured
  // SLOWER here: median 4117 -> 3594, quantiles 3329 -> 2977 GB/s at t = 100; 16 warps per SM are resident anyway)
  return launch_select_split_w<NH, FUSED, MASKED, 128>(in, P, s, mk);
}

// run length for a column of R samples: the smallest network with 2 * NH >= R
template <int FUSED, bool MASKED>
int launch_select_split_r(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk = nullptr) {
  if (in.R <= 50) return launch_select_split_n<25, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 64) return launch_select_split_n<32, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 100) return launch_select_split_n<50, FUSED, MASKED>(in, P, s, mk);
  return launch_select_split_n<64, FUSED, MASKED>(in, P, s, mk);
}

// Which calls take the split kernel, among those k_select_tile takes (float32, 32 < R <= 128, TMA-addressable,
// enough columns for a persistent grid): at most EOC_SPLIT_MAX_Q probabilities, no sorted output (eoc_sort needs
// the column as ONE run).  The plain median (no quantiles, no fused reducers, no mask) of 64 < R <= 128 samples goes
// to k_select_quad (four runs, select_quad.cu: 4980 GB/s at t = 100); for R <= 64 it stays on the pruned median cone
// of k_select_tile, which is ahead of the two-run form there (4366 vs 4118 GB/s at t = 100).
// EOC_SELECT_SPLIT=0 keeps the round-1 kernels everywhere, =1 sends the plain median here too (A/B runs, tests).
inline bool select_split_applies(const eoc_view& in, int nq, bool want_sorted, bool plain_median) {
  (void)in;
  if (want_sorted || nq > EOC_SPLIT_MAX_Q) return false;
  const int env = split_env("EOC_SELECT_SPLIT", -1);
  if (env == 0) return false;
  if (plain_median && env != 1) return false;
  return true;
}

// select_split.cu / select_split_fused.cu / select_split_masked.cu
int launch_select_split(const eoc_view& in, const SelectParams& P, cudaStream_t s);
int launch_select_split_fused(const eoc_view& in, const SelectParams& P, int level, cudaStream_t s);
int launch_select_split_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s);
// select_quad*.cu: four runs of 20 / 25 / 32 samples (64 < R <= 128), see select_quad_kernel.cuh.  EOC_SELECT_QUAD=0
// keeps the round-1 / two-run kernels (A/B runs, tests); EOC_QUAD_QUANT / EOC_QUAD_FUSED / EOC_QUAD_MASKED = 0 / 1
// switch single uses.
int launch_select_quad(const eoc_view& in, const SelectParams& P, cudaStream_t s);
int launch_select_quad_fused(const eoc_view& in, const SelectParams& P, int level, cudaStream_t s);
int launch_select_quad_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s);
int launch_select_quad_int(const eoc_view& in, const SelectParams& P, cudaStream_t s);
int launch_select_quad_masked_int(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s);
int launch_select_quad64(const eoc_view& in, const SelectParams& P, cudaStream_t s);  // select_quad64.cu: float64 columns
int launch_select_multi64(const eoc_view& in, const SelectParams& P, cudaStream_t s);  // select_multi64.cu: long float64 columns
// plain float32 median / quantiles of 17 .. 32 samples (runs of 6 / 8) on a TMA-addressable view with enough columns.
// OFF unless EOC_QUAD_SHORT=1: the register kernel is ahead there (measured on B200, median / quantiles, GB/s:
// t = 17: 2843 / 1728 vs 2802 / 1942; t = 24: 4017 / 2324 vs 3914 / 2731; t = 32: 4505 / 2744 vs 5367 / 3847 --
// the per-tile overheads of the staged kernel weigh more on short columns).
inline bool select_quad_short_applies(const eoc_view& in, int nq, bool want_sorted) {
  if (split_env("EOC_QUAD_SHORT", 0) != 1) return false;
  if (want_sorted || nq > EOC_SPLIT_MAX_Q || in.dtype != EOC_F32 || in.R <= 16 || in.R > 32) return false;
  if (split_env("EOC_SELECT_QUAD", 1) != 1) return false;
  if (nq > 0 && split_env("EOC_QUAD_QUANT", 1) != 1) return false;
  const char* e = getenv("EOC_SELECT_TMA");
  if (e && e[0] == '0') return false;
  if (!tma_view_ok(in)) return false;
  if (e && e[0] == '1') return true;
  return in.outer * in.inner >= 65536 && tma_tiles_fill(in, 128);
}
enum QuadUse { QUAD_PLAIN, QUAD_FUSED, QUAD_MASKED };
inline bool select_quad_applies(const eoc_view& in, int nq, bool want_sorted, QuadUse use = QUAD_PLAIN) {
  // EOC_QUAD_MIN_R: the shortest column the four-run kernels take (32 < R <= 64 joined late in round 2; the plain
  // variant also takes 16 < R <= 32, see select_quad_short_applies)
  if (want_sorted || nq > split_env("EOC_QUAD_MAX_Q", EOC_SPLIT_MAX_Q) || in.R <= split_env("EOC_QUAD_MIN_R", 32) || in.R > 128)
    return false;
  if (split_env("EOC_SELECT_QUAD", 1) != 1) return false;
  if (use == QUAD_FUSED) return split_env("EOC_QUAD_FUSED", 1) == 1;
  if (use == QUAD_MASKED) return split_env("EOC_QUAD_MASKED", 1) == 1;
  if (nq > 0) return split_env("EOC_QUAD_QUANT", 1) == 1;
  return true;
}

}  // namespace eoc
