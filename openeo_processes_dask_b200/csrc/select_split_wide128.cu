// Split selection kernel for long columns, 128-column tiles.  See select_split_wide.cuh.
#include "select_split_wide.cuh"

namespace eoc {

int launch_select_split_wide128(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return launch_select_split_wide_w<128>(in, P, s);
}

}  // namespace eoc
