// k_select_quad on raw uint16 / int16 / uint8 cubes (median / quantiles of 64 < R <= 128 samples, float64 results).
#include "select_quad_launch.cuh"

namespace eoc {

int launch_select_quad_int(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  switch (in.dtype) {
    case EOC_U16: return launch_quad_int_r<uint16_t>(in, P, s);
    case EOC_I16: return launch_quad_int_r<int16_t>(in, P, s);
    case EOC_U8: return launch_quad_int_r<uint8_t>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "launch_select_quad_int: dtype %d", in.dtype);
}

int launch_select_quad_masked_int(const eoc_view& in, const eoc_view& mk, const SelectParams& P, cudaStream_t s) {
  switch (in.dtype) {
    case EOC_U16: return launch_quad_int_masked_r<uint16_t>(in, mk, P, s);
    case EOC_I16: return launch_quad_int_masked_r<int16_t>(in, mk, P, s);
    case EOC_U8: return launch_quad_int_masked_r<uint8_t>(in, mk, P, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "launch_select_quad_masked_int: dtype %d", in.dtype);
}

}  // namespace eoc
