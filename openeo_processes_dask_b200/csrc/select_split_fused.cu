// Split selection kernel, fused variants (order statistics + streaming reducers of the same column in one pass).
#include "select_split_launch.cuh"

namespace eoc {

int launch_select_split_fused(const eoc_view& in, const SelectParams& P, int level, cudaStream_t s) {
  return level >= 2 ? launch_select_split_r<2, false>(in, P, s) : launch_select_split_r<1, false>(in, P, s);
}

}  // namespace eoc
