// Split selection kernel for long columns, 64-column tiles.  See select_split_wide.cuh.
#include "select_split_wide.cuh"

namespace eoc {

int launch_select_split_wide64(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return launch_select_split_wide_w<64>(in, P, s);
}

}  // namespace eoc
