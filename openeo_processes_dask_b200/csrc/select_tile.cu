// TMA-staged register selection kernel, plain variants (median / quantiles).  See select_tile_kernel.cuh.
#include <cstdlib>

#include "select_tile_launch.cuh"

namespace eoc {

int launch_select_tile(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return quant ? launch_select_tile_q<true, 0>(in, P, s) : launch_select_tile_q<false, 0>(in, P, s);
}

}  // namespace eoc
