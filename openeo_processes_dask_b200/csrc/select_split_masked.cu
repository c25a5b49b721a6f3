// Split selection kernel with the mask fused into the register fill (eoc_masked_select): 128-column tiles, data
// tile + uint8 mask tile = 64 KB at R = 100.
#include "select_split_launch.cuh"

namespace eoc {

int launch_select_split_masked(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s) {
  return launch_select_split_r<0, true>(in, P, s, &mk);
}

}  // namespace eoc
