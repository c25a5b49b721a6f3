// Deterministic synthetic samples (device twin of oracle/synth.py); see eoc_synth_fill in eoc.h.
#include "column_math.cuh"
#include "eoc_common.cuh"

namespace eoc {
namespace {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

template <typename T>
__global__ void __launch_bounds__(256) k_synth(T* out, int64_t planes, int64_t plane_stride,
                                               int64_t first_index, int64_t strip_len, uint64_t base,
                                               uint32_t nan_threshold) {
  const int64_t total = planes * strip_len;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = e / strip_len;
    const int64_t j = e - p * strip_len;
    const uint64_t idx = (uint64_t)(p * plane_stride + first_index + j);
    const uint64_t h = splitmix64(base + idx);
    T v = (T)((h >> 40) % 10000ull);
    if constexpr (std::is_floating_point<T>::value) {
      if ((uint32_t)(h & 0xFFFFFFull) < nan_threshold) v = Lim<T>::nan();
    }
    out[e] = v;
  }
}

template <typename T>
int run(void* out, int64_t planes, int64_t plane_stride, int64_t first_index, int64_t strip_len,
        uint64_t seed, uint32_t thr, cudaStream_t s) {
  const int64_t total = planes * strip_len;
  if (total == 0) return EOC_OK;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 32;
  if (blocks > cap) blocks = cap;
  k_synth<T><<<(unsigned)blocks, 256, 0, s>>>((T*)out, planes, plane_stride, first_index, strip_len,
                                               seed * 0x9E3779B97F4A7C15ull, thr);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace

int launch_synth(void* out, int dtype, int64_t planes, int64_t plane_stride, int64_t first_index,
                 int64_t strip_len, uint64_t seed, uint32_t nan_threshold, cudaStream_t s) {
  if (!out || planes < 0 || strip_len < 0) EOC_FAIL(EOC_ERR_INVALID, "bad synth_fill arguments");
  switch (dtype) {
    case EOC_F32: return run<float>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
    case EOC_F64: return run<double>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
    case EOC_U8: return run<uint8_t>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
    case EOC_I16: return run<int16_t>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
    case EOC_U16: return run<uint16_t>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
    case EOC_I32: return run<int32_t>(out, planes, plane_stride, first_index, strip_len, seed, nan_threshold, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", dtype);
}

}  // namespace eoc
