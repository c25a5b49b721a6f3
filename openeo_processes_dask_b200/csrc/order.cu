// eoc_order / eoc_take -- `order` and `rearrange` along a dimension (reference arrays.py:498-555).
//
// order: np.argsort(data, kind="mergesort", axis): the permutation that sorts every column ascending, STABLE (equal
// samples keep their index order; -0.0 == +0.0) with the NaNs last, themselves in index order.  A sorting network
// is not stable, but ordering (value, index) PAIRS lexicographically is a total order, so any correct sorter gives
// the stable permutation.  Generic kernel: one thread per column, (value, index) pairs in shared memory
// ([slot][thread]: conflict-free), run-time bitonic network over the next power of two; any dtype, R <= 512.
// Shared-memory bound (~18 K instructions per 100-sample column): a functional row, not a tuned one.
// take: out[o, j, i] = in[o, idx[j], i] -- np.take(data, order, axis) of rearrange, a gather of planes.
#include <type_traits>

#include "column_kernel.cuh"

namespace eoc {
namespace {

template <typename A> __device__ __forceinline__ bool pair_less(A va, int ia, A vb, int ib) {
  const bool na = va != va, nb = vb != vb;
  if (na != nb) return nb;           // a number sorts before a NaN
  if (!na && va != vb) return va < vb;
  return ia < ib;                    // equal values (or two NaNs): index order
}

template <typename In, typename A>
__global__ void k_order_smem(const In* __restrict__ data, int64_t outer, int64_t R, int64_t inner, int64_t so,
                             int64_t sr, long long* __restrict__ out, int npad) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int W = blockDim.x, t = threadIdx.x;
  A* val = reinterpret_cast<A*>(smem_raw);                         // [npad][W]
  int* idx = reinterpret_cast<int*>(smem_raw + (size_t)npad * W * sizeof(A));  // [npad][W]
  const int64_t item = (int64_t)blockIdx.x * W + t;
  if (item >= outer * inner) return;  // columns are independent: no block-wide barrier below
  const int64_t o = item / inner, i = item - o * inner;
  const In* col = data + o * so + i;
  for (int j = 0; j < npad; ++j) {
    // pads: NaN with an index past every real one -> they end up behind the real NaNs
    val[j * W + t] = j < R ? (A)col[(int64_t)j * sr] : Lim<A>::nan();
    idx[j * W + t] = j;
  }
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int a = 0; a < npad; ++a) {
        const int b = a ^ j;
        if (b > a) {
          const A va = val[a * W + t], vb = val[b * W + t];
          const int ia = idx[a * W + t], ib = idx[b * W + t];
          const bool asc = (a & k) == 0;
          const bool swap = asc ? pair_less(vb, ib, va, ia) : pair_less(va, ia, vb, ib);
          if (swap) {
            val[a * W + t] = vb;
            val[b * W + t] = va;
            idx[a * W + t] = ib;
            idx[b * W + t] = ia;
          }
        }
      }
    }
  }
  long long* dst = out + o * R * inner + i;
  for (int j = 0; j < R; ++j) dst[(int64_t)j * inner] = idx[j * W + t];
}

template <typename In> int launch_order_t(const eoc_view& in, long long* out, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  int npad = 2;
  while (npad < in.R) npad <<= 1;
  int threads = 128;
  const size_t per = sizeof(A) + sizeof(int);
  while (threads > 32 && (size_t)npad * threads * per > 100 * 1024) threads >>= 1;
  const size_t smem = (size_t)npad * threads * per;
  if (smem > 227 * 1024) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_order: R=%lld does not fit in shared memory", (long long)in.R);
  const int64_t items = in.outer * in.inner;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_order: cube too large for one launch");
  auto kern = k_order_smem<In, A>;
  if (smem > 48 * 1024) EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, in.outer, in.R, in.inner, in.stride_outer,
                                               in.stride_R, out, npad);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename T>
__global__ void __launch_bounds__(256) k_take(const T* __restrict__ data, int64_t outer, int64_t n_idx, int64_t inner,
                                              int64_t so, int64_t sr, const long long* __restrict__ index, int64_t R,
                                              T* __restrict__ out) {
  const int64_t total = outer * n_idx * inner;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += step) {
    const int64_t i = e % inner, j = (e / inner) % n_idx, o = e / (inner * n_idx);
    long long r = index[j];
    r = r < 0 ? r + R : r;  // numpy's negative indices
    out[e] = data[o * so + r * sr + i];
  }
}

template <typename T> int launch_take_t(const eoc_view& in, const long long* index, int64_t n_idx, void* out, cudaStream_t s) {
  const int64_t total = in.outer * n_idx * in.inner;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  k_take<T><<<(unsigned)blocks, 256, 0, s>>>((const T*)in.data, in.outer, n_idx, in.inner, in.stride_outer, in.stride_R,
                                             index, in.R, (T*)out);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" int eoc_order(const eoc_view* in, void* out, void* stream) {
  if (int rc = check_view(in, "eoc_order")) return rc;
  if (!out) EOC_FAIL(EOC_ERR_INVALID, "eoc_order: out is NULL");
  if (in->R < 1) EOC_FAIL(EOC_ERR_INVALID, "eoc_order: R must be >= 1");
  if (in->R > EOC_SELECT_MAX_R)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_order: R=%lld exceeds EOC_SELECT_MAX_R=%d", (long long)in->R, EOC_SELECT_MAX_R);
  if (in->outer * in->inner == 0) return EOC_OK;
  cudaStream_t s = (cudaStream_t)stream;
  switch (in->dtype) {
    case EOC_F32: return launch_order_t<float>(*in, (long long*)out, s);
    case EOC_F64: return launch_order_t<double>(*in, (long long*)out, s);
    case EOC_U8: return launch_order_t<uint8_t>(*in, (long long*)out, s);
    case EOC_I16: return launch_order_t<int16_t>(*in, (long long*)out, s);
    case EOC_U16: return launch_order_t<uint16_t>(*in, (long long*)out, s);
    case EOC_I32: return launch_order_t<int32_t>(*in, (long long*)out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in->dtype);
}

extern "C" int eoc_take(const eoc_view* in, const int64_t* index, int64_t n_index, void* out, void* stream) {
  if (int rc = check_view(in, "eoc_take")) return rc;
  if (n_index < 0 || (n_index > 0 && (!index || !out))) EOC_FAIL(EOC_ERR_INVALID, "eoc_take: index / out is NULL");
  if (in->outer * n_index * in->inner == 0) return EOC_OK;
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype_size(in->dtype)) {  // a gather moves bytes: the element width is all that matters
    case 1: return launch_take_t<uint8_t>(*in, (const long long*)index, n_index, out, s);
    case 2: return launch_take_t<uint16_t>(*in, (const long long*)index, n_index, out, s);
    case 4: return launch_take_t<uint32_t>(*in, (const long long*)index, n_index, out, s);
    case 8: return launch_take_t<unsigned long long>(*in, (const long long*)index, n_index, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in->dtype);
}
