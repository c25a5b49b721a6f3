// Launch template of k_select_quad, shared by select_quad.cu (plain), select_quad_fused.cu and
// select_quad_masked.cu (separate translation units: ptxas works on them in parallel).
#pragma once

#include "select_quad_kernel.cuh"
#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int W, int FUSED, bool MASKED, int G = 1, typename In = float>
int launch_quad_w(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk) {
  constexpr int KR = 4;
  constexpr int MINB = W * G >= 256 ? (MASKED ? 1 : (W >= 256 ? 2 : 4)) : (MASKED ? 3 : 4);
  CUtensorMap map;
  if (int rc = tma_make_map(in, W, NH, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  const int nbox = (int)((in.R + NH - 1) / NH);  // runs that hold samples: one full box each
  T.tile_bytes = (uint32_t)((size_t)nbox * NH * W * sizeof(In));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = 0;
  T.mask = MASKED ? (const uint8_t*)mk->data : nullptr;
  T.mso = MASKED ? mk->stride_outer : 0;
  T.msr = MASKED ? mk->stride_R : 0;
  T.mf_mul = 0xFFC00000u;
  const size_t smem = (size_t)KR * (NH + 1) * W * sizeof(float) + (MASKED ? (size_t)in.R * W : 0);
  int per_sm = 0;
#define EOC_QUAD_GO(FULLV)                                                                              \
  do {                                                                                                  \
    auto kern = k_select_quad<NH, FULLV, W, MINB, FUSED, MASKED, G, In>;                                       \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * G, smem));                \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W * G, smem, s>>>(map, T);                                                   \
  } while (0)
  if (in.R == KR * NH)
    EOC_QUAD_GO(true);
  else
    EOC_QUAD_GO(false);
#undef EOC_QUAD_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// run length for a column of R samples: the smallest raw network with 4 * NH >= R.  Tile width 128 (measured on
// B200, median t = 100: 5064 GB/s against 4478 with 64-column tiles -- 256-byte rows make poorer DRAM bursts).
template <int FUSED, bool MASKED>
int launch_quad_r(const eoc_view& in, const SelectParams& P, cudaStream_t s, const eoc_view* mk = nullptr) {
  if constexpr (FUSED == 0 && !MASKED) {
    const int g = split_env("EOC_QUAD_G", 1);
    if (g == 2 && split_env("EOC_QUAD_W", 128) == 64) {
      if (in.R <= 80) return launch_quad_w<20, 64, 0, false, 2>(in, P, s, mk);
      if (in.R <= 100) return launch_quad_w<25, 64, 0, false, 2>(in, P, s, mk);
      return launch_quad_w<32, 64, 0, false, 2>(in, P, s, mk);
    }
    if (g == 2) {
      if (in.R <= 80) return launch_quad_w<20, 128, 0, false, 2>(in, P, s, mk);
      if (in.R <= 100) return launch_quad_w<25, 128, 0, false, 2>(in, P, s, mk);
      return launch_quad_w<32, 128, 0, false, 2>(in, P, s, mk);
    }
    if (split_env("EOC_QUAD_W", 128) == 256) {
      if (in.R <= 80) return launch_quad_w<20, 256, 0, false>(in, P, s, mk);
      if (in.R <= 100) return launch_quad_w<25, 256, 0, false>(in, P, s, mk);
      return launch_quad_w<32, 256, 0, false>(in, P, s, mk);
    }
    if (split_env("EOC_QUAD_W", 128) == 64) {
      if (in.R <= 80) return launch_quad_w<20, 64, 0, false>(in, P, s, mk);
      if (in.R <= 100) return launch_quad_w<25, 64, 0, false>(in, P, s, mk);
      return launch_quad_w<32, 64, 0, false>(in, P, s, mk);
    }
  }
  // (run lengths 13 / 14: a year of weekly composites is 52 - 53 samples; with 16-sample runs the quantile variants
  // lost to the two-run kernel at t = 50: 2734 vs 3191 GB/s)
  if constexpr (FUSED == 0 && !MASKED) {  // 17 .. 32 samples: plain median / quantiles only
    if (in.R <= 24) return launch_quad_w<6, 128, 0, false>(in, P, s, mk);
    if (in.R <= 32) return launch_quad_w<8, 128, 0, false>(in, P, s, mk);
  }
  if (in.R <= 48) return launch_quad_w<12, 128, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 52) return launch_quad_w<13, 128, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 56) return launch_quad_w<14, 128, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 64) return launch_quad_w<16, 128, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 80) return launch_quad_w<20, 128, FUSED, MASKED>(in, P, s, mk);
  if (in.R <= 100) return launch_quad_w<25, 128, FUSED, MASKED>(in, P, s, mk);
  return launch_quad_w<32, 128, FUSED, MASKED>(in, P, s, mk);
}

// raw 8 / 16-bit integer cubes under a mask (mask -> median / quantiles of digital numbers in one pass)
template <typename In> int launch_quad_int_masked_r(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 48) return launch_quad_w<12, 128, 0, true, 1, In>(in, P, s, &mk);
  if (in.R <= 64) return launch_quad_w<16, 128, 0, true, 1, In>(in, P, s, &mk);
  if (in.R <= 80) return launch_quad_w<20, 128, 0, true, 1, In>(in, P, s, &mk);
  if (in.R <= 100) return launch_quad_w<25, 128, 0, true, 1, In>(in, P, s, &mk);
  return launch_quad_w<32, 128, 0, true, 1, In>(in, P, s, &mk);
}

// raw 8 / 16-bit integer cubes (select_quad_int.cu): float64 results, see the kernel
template <typename In> int launch_quad_int_r(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 48) return launch_quad_w<12, 128, 0, false, 1, In>(in, P, s, nullptr);
  if (in.R <= 52) return launch_quad_w<13, 128, 0, false, 1, In>(in, P, s, nullptr);
  if (in.R <= 56) return launch_quad_w<14, 128, 0, false, 1, In>(in, P, s, nullptr);
  if (in.R <= 64) return launch_quad_w<16, 128, 0, false, 1, In>(in, P, s, nullptr);
  if (in.R <= 80) return launch_quad_w<20, 128, 0, false, 1, In>(in, P, s, nullptr);
  if (in.R <= 100) return launch_quad_w<25, 128, 0, false, 1, In>(in, P, s, nullptr);
  return launch_quad_w<32, 128, 0, false, 1, In>(in, P, s, nullptr);
}

}  // namespace eoc
