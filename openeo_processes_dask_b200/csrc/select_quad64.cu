// Median / quantiles of float64 columns of 32 < R <= 128 samples on the four-run scheme.  See select_quad64_kernel.cuh.
#include "select_quad64_kernel.cuh"
#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int W, int G>
static int launch_quad64_w(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int KR = 4;
  constexpr int MINB = 2;
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
  int per_sm = 0;
#define EOC_Q64_GO(FULLV)                                                                               \
  do {                                                                                                  \
    auto kern = k_select_quad64<NH, FULLV, W, MINB, G>;                                                 \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * G, smem));                \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W * G, smem, s>>>(map, T);                                                   \
  } while (0)
  if (in.R == KR * NH)
    EOC_Q64_GO(true);
  else
    EOC_Q64_GO(false);
#undef EOC_Q64_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

int launch_select_quad64(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  // threads per column, measured on B200 (t = 100, median / quantiles, GB/s of float64 input): 1: 3785 / 2991,
  // 2: 3669 / 2832, 4: 3050 / 2108
  const int g = split_env("EOC_QUAD64_G", 1);
#define EOC_Q64_PICK(NH)                                                  \
  do {                                                                    \
    if (g == 2) return launch_quad64_w<NH, 64, 2>(in, P, s);              \
    if (g == 4) return launch_quad64_w<NH, 64, 4>(in, P, s);              \
    return launch_quad64_w<NH, 64, 1>(in, P, s);                          \
  } while (0)
  if (in.R <= 48) EOC_Q64_PICK(12);
  if (in.R <= 64) EOC_Q64_PICK(16);
  if (in.R <= 80) EOC_Q64_PICK(20);
  if (in.R <= 100) EOC_Q64_PICK(25);
  EOC_Q64_PICK(32);
#undef EOC_Q64_PICK
}

}  // namespace eoc
