// Fused float32 variants, level 2 (every streaming reducer).  See select_fused.cu.
#include "select_launch.cuh"

namespace eoc {

int launch_select_fused2_f32(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return quant ? launch_regs_n<float, float, true, 2>(in, P, s) : launch_regs_n<float, float, false, 2>(in, P, s);
}

}  // namespace eoc
