// Split selection kernel, plain variants (median / up to EOC_SPLIT_MAX_Q quantiles).  See select_split_kernel.cuh.
#include "select_split_launch.cuh"

namespace eoc {

int launch_select_split(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return launch_select_split_r<0, false>(in, P, s);
}

}  // namespace eoc
