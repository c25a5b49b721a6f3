// Fused float32 variants of the register selection kernel (eoc_select_fused): order statistics plus
// streaming reducers of the same register-resident column.  See select_kernel.cuh.
#include "select_launch.cuh"

namespace eoc {

int launch_select_fused_f32(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return quant ? launch_regs_n<float, float, true, true>(in, P, s)
               : launch_regs_n<float, float, false, true>(in, P, s);
}

}  // namespace eoc
