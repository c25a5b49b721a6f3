// Median / quantiles of long float32 columns (128 < R <= 512) as eight sorted runs.  See select_multi_kernel.cuh.
#include "select_multi_kernel.cuh"
#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int G, int W, typename In = float>
static int launch_multi_w(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int KR = 8;
  constexpr int MINB = (W * G >= 512 || NH > 64) ? 1 : (W * G <= 128 ? (NH <= 32 ? 4 : 2) : (NH <= 32 ? 3 : 2));
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
  T.mask = nullptr;
  T.mso = T.msr = 0;
  T.mf_mul = 0xFFC00000u;
  const size_t smem = (size_t)KR * (NH + 1) * W * sizeof(float);
  int per_sm = 0;
#define EOC_MULTI_GO(FULLV)                                                                             \
  do {                                                                                                  \
    auto kern = k_select_multi<NH, G, W, MINB, FULLV, In>;                                                  \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * G, smem));                \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W * G, smem, s>>>(map, T);                                                   \
  } while (0)
  if (in.R == KR * NH)
    EOC_MULTI_GO(true);
  else
    EOC_MULTI_GO(false);
#undef EOC_MULTI_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// run length for a column of R samples: the smallest raw network with 8 * NH >= R
int launch_select_multi(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int w = split_env("EOC_MULTI_W", 0), g = split_env("EOC_MULTI_G", 0);  // tuning knobs (A/B runs)
#define EOC_MULTI_PICK(NH, GD, WD)                                                       \
  do {                                                                                   \
    if (w == 32 && g == 8) return launch_multi_w<NH, 8, 32>(in, P, s);                   \
    if (w == 32 && g == 4) return launch_multi_w<NH, 4, 32>(in, P, s);                   \
    if (w == 64 && g == 8) return launch_multi_w<NH, 8, 64>(in, P, s);                   \
    if (w == 64 && g == 4) return launch_multi_w<NH, 4, 64>(in, P, s);                   \
    if (w == 64 && g == 2) return launch_multi_w<NH, 2, 64>(in, P, s);                   \
    if (w == 64 && g == 1) return launch_multi_w<NH, 1, 64>(in, P, s);                   \
    if constexpr (NH <= 50) {                                                            \
      if (w == 128 && g == 2) return launch_multi_w<NH, 2, 128>(in, P, s);               \
      if (w == 128 && g == 1) return launch_multi_w<NH, 1, 128>(in, P, s);               \
    }                                                                                    \
    return launch_multi_w<NH, GD, WD>(in, P, s);                                         \
  } while (0)
  // threads per column G and tile width W measured on B200 (median / quantiles, input GB/s, 30 % NaN; the K-run split
  // kernel of select_split_wide.cuh in brackets): t = 160: (2, 64) 3747 / 2507; t = 200: (2, 64) 3915 / 2728, (4, 64)
  // 3397 / 2559 [2920 / 2646]; t = 256: (4, 64) 3639 / 2805, (2, 64) 3642 / 2679 [2269 / 2072]; t = 365: (4, 64)
  // 2768 / 2342, (8, 64) 2078 / 1814 [2474 / 2288]; t = 400: (4, 64) 3109 / 2657 [2521 / 2250]; t = 512: (8, 64)
  // 2492 / 2215, (4, 64) 2385 / 2091 [1849 / 1709]; 128-column tiles and one thread per column are never ahead.
  if (in.R <= 160) EOC_MULTI_PICK(20, 2, 64);
  if (in.R <= 200) EOC_MULTI_PICK(25, 2, 64);
  if (in.R <= 256) EOC_MULTI_PICK(32, 4, 64);
  if (in.R <= 320) EOC_MULTI_PICK(40, 4, 64);
  if (in.R <= 400) EOC_MULTI_PICK(50, 4, 64);
  if (in.R <= 512) EOC_MULTI_PICK(64, 8, 64);
  // 513 .. 1024 samples (two to three years of daily scenes): runs of 128 in 32-column tiles (132 KB), one thread per
  // run; before this the any-length general path took them at 11 GB/s
  return launch_multi_w<128, 8, 32>(in, P, s);
#undef EOC_MULTI_PICK
}

// raw 8 / 16-bit integer cubes: the measured (G, W) of the float32 table, float64 results
template <typename In> static int launch_multi_int_t(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 160) return launch_multi_w<20, 2, 64, In>(in, P, s);
  if (in.R <= 200) return launch_multi_w<25, 2, 64, In>(in, P, s);
  if (in.R <= 256) return launch_multi_w<32, 4, 64, In>(in, P, s);
  if (in.R <= 320) return launch_multi_w<40, 4, 64, In>(in, P, s);
  if (in.R <= 400) return launch_multi_w<50, 4, 64, In>(in, P, s);
  if (in.R <= 512) return launch_multi_w<64, 8, 64, In>(in, P, s);
  return launch_multi_w<128, 8, 32, In>(in, P, s);  // 513 .. 1024 samples
}
int launch_select_multi_int(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  switch (in.dtype) {
    case EOC_U16: return launch_multi_int_t<uint16_t>(in, P, s);
    case EOC_I16: return launch_multi_int_t<int16_t>(in, P, s);
    case EOC_U8: return launch_multi_int_t<uint8_t>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "launch_select_multi_int: dtype %d", in.dtype);
}

}  // namespace eoc
