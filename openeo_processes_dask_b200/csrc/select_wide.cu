// Wide selection kernel (128 < R <= 512), float32 instantiation.  See select_wide_kernel.cuh.
#include "select_wide_kernel.cuh"

namespace eoc {

bool select_wide_applies(const eoc_view& in) {
  const bool small_int = in.dtype == EOC_U8 || in.dtype == EOC_I16 || in.dtype == EOC_U16;
  return (in.dtype == EOC_F32 || small_int) && in.R > 128 && in.R <= 512 &&
         (uint64_t)in.stride_R * dtype_size(in.dtype) < (1ull << 32);
}

int launch_select_wide(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  return launch_wide_t<float>(in, P, s);
}

}  // namespace eoc
