// Plain median of float32 columns of 64 < R <= 128 samples: four short sorted runs + rank elimination.
// See select_quad_kernel.cuh.
#include "select_quad_kernel.cuh"
#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int W>
static int launch_quad_w(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr int KR = 4;
  constexpr int MINB = W >= 256 ? 2 : 4;
  CUtensorMap map;
  if (int rc = tma_make_map(in, W, NH, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  const int nbox = (int)((in.R + NH - 1) / NH);  // runs that hold samples: one full box each
  T.tile_bytes = (uint32_t)((size_t)nbox * NH * W * sizeof(float));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = 0;
  T.mask = nullptr;
  T.mso = T.msr = 0;
  T.mf_mul = 0xFFC00000u;
  const size_t smem = (size_t)KR * (NH + 1) * W * sizeof(float);
  int per_sm = 0;
#define EOC_QUAD_GO(FULLV)                                                                              \
  do {                                                                                                  \
    auto kern = k_select_quad<NH, FULLV, W, MINB>;                                                      \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W, smem));                    \
    if (per_sm < 1) per_sm = 1;                                                                         \
    int64_t grid = (int64_t)sm_count() * per_sm;                                                        \
    if (grid > T.num_tiles) grid = T.num_tiles;                                                         \
    kern<<<(unsigned)grid, W, smem, s>>>(map, T);                                                       \
  } while (0)
  if (in.R == KR * NH)
    EOC_QUAD_GO(true);
  else
    EOC_QUAD_GO(false);
#undef EOC_QUAD_GO
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// run length for a column of R samples: the smallest raw network with 4 * NH >= R
int launch_select_quad(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const bool w64 = split_env("EOC_QUAD_W", 128) == 64;
  if (in.R <= 80) return w64 ? launch_quad_w<20, 64>(in, P, s) : launch_quad_w<20, 128>(in, P, s);
  if (in.R <= 100) return w64 ? launch_quad_w<25, 64>(in, P, s) : launch_quad_w<25, 128>(in, P, s);
  return w64 ? launch_quad_w<32, 64>(in, P, s) : launch_quad_w<32, 128>(in, P, s);
}

}  // namespace eoc
