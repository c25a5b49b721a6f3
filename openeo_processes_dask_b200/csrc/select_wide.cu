// Wide selection kernel (128 < R <= 512), float32 instantiation.  See select_wide_kernel.cuh.
// measured on B200 (t = 200 / 365): 1 of 2 comparators in the mixed-pipe form 1420 / 855 GB/s, 3 of 5 (the default
// of the narrow kernels) 1384 / 704, none 1184 / 767 -- these kernels wait on instruction fetch, code size counts
#define EOC_MIXQ_NUM 1
#define EOC_MIXQ_DEN 2
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

// float64 columns, 64 < R <= 512: 64 samples per thread, 2 / 4 / 8 threads per column
bool select_wide64_applies(const eoc_view& in) {
  return in.dtype == EOC_F64 && in.R > 64 && in.R <= 512 && (uint64_t)in.stride_R * 8 < (1ull << 32);
}

int launch_select_wide64(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 128) return launch_wide_k<double, double, 64, 2>(in, P, s);
  if (in.R <= 256) return launch_wide_k<double, double, 64, 4>(in, P, s);
  return launch_wide_k<double, double, 64, 8>(in, P, s);
}

}  // namespace eoc
