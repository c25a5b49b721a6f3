// int16_t instantiation of the integer selection path.  See select_int.cuh.
#include "select_int.cuh"

namespace eoc {

int launch_select_int_i16(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return launch_select_int_t<int16_t>(in, P, quant, s);
}

}  // namespace eoc
