// Median / quantiles of float32 columns of 64 < R <= 128 samples: four short sorted runs + rank elimination.
// See select_quad_kernel.cuh.
#include "select_quad_launch.cuh"

namespace eoc {

int launch_select_quad(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return launch_quad_r<0, false>(in, P, s);
}

}  // namespace eoc
