// eoc_select_general -- order statistics and sort for columns of ANY length (R > EOC_SELECT_MAX_R = 512: several
// years of daily scenes, a spatial dimension).  np.nanmedian / np.nanquantile / np.sort take any length
// (math.py:184-188, 341-387, arrays.py:560-584); the register / shared-memory kernels stop at 512.
//
// The slower general path: every column is copied into a caller-provided workspace laid out [slot][column]
// (adjacent threads = adjacent columns: every access of the sort is coalesced) and sorted IN PLACE by Batcher's
// merge exchange (Knuth TAOCP 5.2.2 Algorithm M), which works for any R without padding because every comparator
// moves the smaller key to the lower slot; NaN counts as the largest key, so the valid samples end up first.
// Then emit_sorted() -- the code of the fast kernels -- fetches the middle pair / quantile neighbours or writes
// the sorted column.  About (R / 4) lg^2 R compare-exchanges per column, each two loads and up to two stores: the
// library walks the cube in batches of columns that fit the workspace, and a workspace of <= 64 MB keeps a batch
// inside the 126 MB L2 so the ~lg^2 R / 2 sweeps do not go to HBM.
#include <type_traits>

#include "select_launch.cuh"

namespace eoc {
namespace {

template <typename A> __device__ __forceinline__ bool key_greater(A a, A b) {
  const bool na = a != a, nb = b != b;
  return na ? !nb : (!nb && a > b);  // NaN is larger than every number; two NaNs are equal
}

template <typename In, typename A>
__global__ void __launch_bounds__(256) k_select_general(const In* __restrict__ data, const SelectParams P, A* __restrict__ ws,
                                                        int64_t col0, int64_t ncols) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncols) return;  // columns are independent
  const int64_t item = col0 + c;
  const int64_t o = item / P.inner, i = item - o * P.inner;
  const In* col = data + o * P.so + i;
  const int R = (int)P.R;
  A* w = ws + c;  // slot j at w[j * ncols]
  int n = 0;
  for (int j = 0; j < R; ++j) {
    const A x = (A)__ldcs(col + (int64_t)j * P.sr);
    n += (x != x) ? 0 : 1;
    w[(int64_t)j * ncols] = x;
  }
  int t = 0;
  while ((1 << t) < R) ++t;
  for (int p = t > 0 ? 1 << (t - 1) : 0; p > 0; p >>= 1) {
    int q = 1 << (t - 1), r = 0, d = p;
    while (true) {
      for (int a = 0; a + d < R; ++a) {
        if ((a & p) == r) {
          const A x = w[(int64_t)a * ncols], y = w[(int64_t)(a + d) * ncols];
          if (key_greater(x, y)) {
            w[(int64_t)a * ncols] = y;
            w[(int64_t)(a + d) * ncols] = x;
          }
        }
      }
      if (q == p) break;
      d = q - p;
      q >>= 1;
      r = p;
    }
  }
  emit_sorted<A>(P, o, i, n, w, (int)ncols);
}

template <typename In>
int launch_general_t(const eoc_view& in, const SelectParams& P, void* workspace, size_t workspace_bytes, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  const int64_t total = in.outer * in.inner;
  int64_t batch = (int64_t)(workspace_bytes / ((size_t)in.R * sizeof(A)));
  if (batch > total) batch = total;
  if (batch > 0x7fffffffLL / in.R) batch = 0x7fffffffLL / in.R;  // emit_sorted indexes slot * stride in an int
  if (batch >= 256) batch -= batch % 32;
  if (batch < 1)
    EOC_FAIL(EOC_ERR_WORKSPACE, "eoc_select_general needs at least %zu workspace bytes (one column), got %zu",
             (size_t)in.R * sizeof(A), workspace_bytes);
  for (int64_t c0 = 0; c0 < total; c0 += batch) {
    const int64_t nc = (total - c0 < batch) ? total - c0 : batch;
    k_select_general<In, A><<<(unsigned)((nc + 255) / 256), 256, 0, s>>>((const In*)in.data, P, (A*)workspace, c0, nc);
    EOC_LAUNCH_CHECK();
  }
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" size_t eoc_select_general_workspace_bytes(const eoc_view* in) {
  if (!in || dtype_size(in->dtype) == 0 || in->R < 1) return 0;
  const size_t es = in->dtype == EOC_F32 ? 4 : 8;
  const size_t all = (size_t)in->outer * (size_t)in->inner * (size_t)in->R * es;
  const size_t l2_friendly = (size_t)64 << 20;
  const size_t one_warp = (size_t)in->R * es * 32;
  size_t want = all < l2_friendly ? all : l2_friendly;
  return want > one_warp ? want : one_warp;
}

extern "C" int eoc_select_general(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                                  void* quant_out, int quant_dtype, void* sort_out, int sort_descending,
                                  int sort_nan_first, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_view(in, "eoc_select_general")) return rc;
  if (nq < 0 || nq > EOC_SELECT_MAX_Q) EOC_FAIL(EOC_ERR_INVALID, "nq=%d outside [0, %d]", nq, EOC_SELECT_MAX_Q);
  if (nq > 0 && (!probs || !quant_out)) EOC_FAIL(EOC_ERR_INVALID, "nq > 0 needs probs and quant_out");
  if (nq == 0 && !median_out && !sort_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute");
  if (quant_dtype != EOC_F64 && quant_dtype != EOC_F32) EOC_FAIL(EOC_ERR_INVALID, "quant_dtype must be f32 or f64");
  if (in->R < 1) EOC_FAIL(EOC_ERR_INVALID, "R must be >= 1 (empty axes are handled by the host layer)");
  if (in->R > (1 << 24)) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_select_general: R=%lld exceeds 2^24", (long long)in->R);
  if (in->outer * in->inner == 0) return EOC_OK;
  if (!workspace) EOC_FAIL(EOC_ERR_WORKSPACE, "eoc_select_general: workspace is NULL (see eoc_select_general_workspace_bytes)");
  SelectParams P;
  P.outer = in->outer;
  P.R = in->R;
  P.inner = in->inner;
  P.so = in->stride_outer;
  P.sr = in->stride_R;
  P.ignore_nodata = ignore_nodata;
  P.median_out = median_out;
  P.median_f64 = 0;
  P.lerp_wrap16 = in->dtype == EOC_I16;
  P.nq = nq;
  for (int q = 0; q < EOC_SELECT_MAX_Q; ++q) P.probs[q] = q < nq ? probs[q] : 0.0;
  P.quant_out = quant_out;
  P.quant_is_f64 = quant_dtype == EOC_F64;
  P.zero = 0.0f;
  P.sort_out = sort_out;
  P.sort_desc = sort_descending;
  P.sort_nan_first = sort_nan_first;
  P.fused = ColumnParams{};
  cudaStream_t s = (cudaStream_t)stream;
  switch (in->dtype) {
    case EOC_F32: return launch_general_t<float>(*in, P, workspace, workspace_bytes, s);
    case EOC_F64: return launch_general_t<double>(*in, P, workspace, workspace_bytes, s);
    case EOC_U8: return launch_general_t<uint8_t>(*in, P, workspace, workspace_bytes, s);
    case EOC_I16: return launch_general_t<int16_t>(*in, P, workspace, workspace_bytes, s);
    case EOC_U16: return launch_general_t<uint16_t>(*in, P, workspace, workspace_bytes, s);
    case EOC_I32: return launch_general_t<int32_t>(*in, P, workspace, workspace_bytes, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in->dtype);
}
