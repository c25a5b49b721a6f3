// TMA-staged register selection kernel, fused variants (order statistics + streaming reducers).
#include <cstdlib>

#include "select_tile_launch.cuh"

namespace eoc {

int launch_select_tile_fused(const eoc_view& in, const SelectParams& P, bool quant, int level, cudaStream_t s) {
  if (level >= 2) return quant ? launch_select_tile_q<true, 2>(in, P, s) : launch_select_tile_q<false, 2>(in, P, s);
  return quant ? launch_select_tile_q<true, 1>(in, P, s) : launch_select_tile_q<false, 1>(in, P, s);
}

}  // namespace eoc
