// TMA-staged register selection kernel, fused variants (order statistics + streaming reducers).
#include <cstdlib>

#include "select_tile_launch.cuh"

namespace eoc {

int launch_select_tile_fused(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return quant ? launch_select_tile_q<true, true>(in, P, s) : launch_select_tile_q<false, true>(in, P, s);
}

}  // namespace eoc
