// TMA-staged register selection kernel with the mask fused into the register fill (eoc_masked_select).
// 128-column tiles: data tile + uint8 mask tile = 64 KB at R = 100, three CTAs per SM.
#include "select_tile_launch.cuh"

namespace eoc {

namespace {
constexpr int MASKED_W = 128;

template <bool QUANT>
int launch_q(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 64) return launch_select_tile_n<64, QUANT, 0, MASKED_W, true>(in, P, s, &mk);
  if (in.R <= 100) return launch_select_tile_n<100, QUANT, 0, MASKED_W, true>(in, P, s, &mk);
  return launch_select_tile_n<128, QUANT, 0, MASKED_W, true>(in, P, s, &mk);
}
}  // namespace

int launch_select_tile_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, bool quant,
                              cudaStream_t s) {
  return quant ? launch_q<true>(in, mk, P, s) : launch_q<false>(in, mk, P, s);
}

}  // namespace eoc
