// k_select_quad with the streaming reducers of eoc_select_fused (quantiles + sd of one read: BASELINE config 3).
#include "select_quad_launch.cuh"

namespace eoc {

int launch_select_quad_fused(const eoc_view& in, const SelectParams& P, int level, cudaStream_t s) {
  if (level >= 2) return launch_quad_r<2, false>(in, P, s);
  return launch_quad_r<1, false>(in, P, s);
}

}  // namespace eoc
