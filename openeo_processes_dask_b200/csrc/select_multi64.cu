// Median / quantiles of long float64 columns (128 < R <= 512) as eight sorted runs.  See select_multi64_kernel.cuh.
#include "select_multi64_kernel.cuh"
#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int G, int W>
static int launch_multi64_w(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int KR = 8;
  constexpr int MINB = 1;
  CUtensorMap map;
  if (int rc = tma_make_map(in, W, NH, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  const int nbox = (int)((in.R + NH - 1) / NH);
  T.tile_bytes = (uint32_t)((size_t)nbox * NH * W * sizeof(double));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = 0;
  T.mask = nullptr;
  T.mso = T.msr = 0;
  T.mf_mul = 0;
  const size_t smem = (size_t)KR * (NH + 1) * W * sizeof(double) + (size_t)G * W * sizeof(int);
  if (smem > 227 * 1024) EOC_FAIL(EOC_ERR_UNSUPPORTED, "float64 tile of %zu bytes does not fit shared memory", smem);
  int per_sm = 0;
#define EOC_M64_GO(FULLV)                                                                               \
  do {                                                                                                  \
    auto kern = k_select_multi64<NH, G, W, MINB, FULLV>;                                                \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * G, smem));                \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W * G, smem, s>>>(map, T);                                                   \
  } while (0)
  if (in.R == KR * NH)
    EOC_M64_GO(true);
  else
    EOC_M64_GO(false);
#undef EOC_M64_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// 8-byte samples: 32-column tiles keep the tile of the longest columns (8 x 65 rows) at 133 KB
int launch_select_multi64(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  // threads per column, measured on B200 (median / quantiles, GB/s of float64 input; the K-thread bitonic kernel
  // before: t = 200: 189 / 187, t = 365: 127 / 126): t = 200 (64-column tiles): G = 1: 2891 / 1959, 2: 2702 / 2118,
  // 4: 2718 / 2236; t = 365 (32-column tiles): G = 1: 735 / 641, 2: 1295 / 1124, 4: 1987 / 1840
  const int g = split_env("EOC_MULTI64_G", 0);
#define EOC_M64_PICK(NH, W, GD)                                           \
  do {                                                                    \
    if (g == 1) return launch_multi64_w<NH, 1, W>(in, P, s);              \
    if (g == 2) return launch_multi64_w<NH, 2, W>(in, P, s);              \
    if (g == 4) return launch_multi64_w<NH, 4, W>(in, P, s);              \
    return launch_multi64_w<NH, GD, W>(in, P, s);                         \
  } while (0)
  if (in.R <= 160) EOC_M64_PICK(20, 64, 2);
  if (in.R <= 200) EOC_M64_PICK(25, 64, 2);
  if (in.R <= 256) EOC_M64_PICK(32, 64, 2);
  if (in.R <= 320) EOC_M64_PICK(40, 32, 4);
  if (in.R <= 400) EOC_M64_PICK(50, 32, 4);
  EOC_M64_PICK(64, 32, 4);
#undef EOC_M64_PICK
}

}  // namespace eoc
