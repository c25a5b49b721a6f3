// Fused float32 variants of the register selection kernel (eoc_select_fused): order statistics plus
// streaming reducers of the same register-resident column.  See select_kernel.cuh.  Level 1 (sum / mean /
// variance / sd) here, level 2 (adds min / max / product / count) in select_fused2.cu.
#include "select_launch.cuh"

namespace eoc {

int launch_select_fused2_f32(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s);

int launch_select_fused_f32(const eoc_view& in, const SelectParams& P, bool quant, int level, cudaStream_t s) {
  if (level >= 2) return launch_select_fused2_f32(in, P, quant, s);
  return quant ? launch_regs_n<float, float, true, 1>(in, P, s) : launch_regs_n<float, float, false, 1>(in, P, s);
}

}  // namespace eoc
