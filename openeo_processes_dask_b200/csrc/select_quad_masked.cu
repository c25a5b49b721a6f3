// k_select_quad with the mask applied in the register fill (mask -> median / quantiles in one pass).
#include "select_quad_launch.cuh"

namespace eoc {

int launch_select_quad_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s) {
  return launch_quad_r<0, true>(in, P, s, &mk);
}

}  // namespace eoc
