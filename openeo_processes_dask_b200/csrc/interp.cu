// array_interpolate_linear along R (reference arrays.py:405-458): every run of non-finite
// samples inside a column is replaced by the straight line between its finite neighbours, runs at either
// end become NaN (np.interp(..., left=nan, right=nan)), columns with fewer than two finite samples or none
// missing are returned as they are.
//
// A CTA stages an (R x 128) tile of columns in shared memory with coalesced loads (any dtype / alignment),
// thread c then walks column c in place -- row r at tile[r*128 + c], conflict-free -- and the tile goes back
// with coalesced stores: R*s bytes in, R*s bytes out per column, each crossing HBM once.  The arithmetic is
// np.interp's (numpy/_core/src/multiarray/compiled_base.c: slope = (fp[j+1]-fp[j]) / (xp[j+1]-xp[j]),
// slope*(x - xp[j]) + fp[j]) in float64 without contraction, cast back to the sample dtype on assignment.
#include <type_traits>

#include "column_launch.cuh"

namespace eoc {

namespace {

constexpr int INTERP_W = 128;

template <typename A> __device__ __forceinline__ bool finite_(A v) { return abs_(v) <= Lim<A>::maxv(); }

// a quiet NaN carrying a small integer in its mantissa, and back
template <typename A> __device__ __forceinline__ A nan_code(int k);
template <> __device__ __forceinline__ float nan_code<float>(int k) { return __uint_as_float(0x7FC00000u | (uint32_t)k); }
template <> __device__ __forceinline__ double nan_code<double>(int k) {
  return __longlong_as_double((long long)(0x7FF8000000000000ull | (uint64_t)k));
}
template <typename A> __device__ __forceinline__ int nan_payload(A v);
template <> __device__ __forceinline__ int nan_payload<float>(float v) { return (int)(__float_as_uint(v) & 0x3FFFFFu); }
template <> __device__ __forceinline__ int nan_payload<double>(double v) {
  return (int)((uint64_t)__double_as_longlong(v) & 0x3FFFFFull);
}

template <typename In, typename A>
__global__ void __launch_bounds__(INTERP_W) k_interp(const In* __restrict__ data, int64_t so, int64_t sr, int64_t outer,
                                                     int R, int64_t inner, const double* __restrict__ xs,
                                                     A* __restrict__ out, int64_t tiles_per_outer) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  A* tile = reinterpret_cast<A*>(smem_raw);                                   // [R][W]
  double* xl = reinterpret_cast<double*>(smem_raw + (size_t)R * INTERP_W * sizeof(A));  // [R] label positions
  const int t = threadIdx.x;
  const int64_t o = blockIdx.x / tiles_per_outer;
  const int64_t i = (blockIdx.x - o * tiles_per_outer) * INTERP_W + t;
  const bool active = i < inner;
  const In* src = data + o * so + (active ? i : 0);
  for (int r = t; r < R; r += INTERP_W) xl[r] = xs ? xs[r] : (double)r;
  int cnt = 0;
  constexpr int U = 10;  // rows requested before any is consumed (one load in flight per thread is latency-bound)
  int r0 = 0;
  for (; r0 + U <= R; r0 += U) {
    In q[U];
#pragma unroll
    for (int u = 0; u < U; ++u) q[u] = active ? __ldcs(src + (int64_t)(r0 + u) * sr) : (In)0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const A v = (A)q[u];
      tile[(r0 + u) * INTERP_W + t] = v;
      cnt += finite_(v) ? 1 : 0;
    }
  }
  for (; r0 < R; ++r0) {
    const A v = active ? (A)__ldcs(src + (int64_t)r0 * sr) : (A)0;
    tile[r0 * INTERP_W + t] = v;
    cnt += finite_(v) ? 1 : 0;
  }
  __syncthreads();  // xl is complete (the tile column is private to its thread)
  A* col = tile + t;
  if (cnt >= 2 && cnt < R) {
    // Backward pass: every non-finite slot receives the index of the next finite sample, packed into a NaN
    // payload (index + 1; 0 = none), so that the forward pass is the same instruction stream for every lane
    // of the warp whatever the gap pattern (a per-gap inner loop diverged badly: 447 GB/s).
    int nx = 0;
    for (int r = R - 1; r >= 0; --r) {
      const A v = col[r * INTERP_W];
      if (finite_(v))
        nx = r + 1;
      else
        col[r * INTERP_W] = nan_code<A>(nx);
    }
    int prev = -1;
    double f0 = 0.0, x0 = 0.0;
    for (int r = 0; r < R; ++r) {
      const A v = col[r * INTERP_W];
      if (finite_(v)) {
        prev = r;
        f0 = (double)v;
        x0 = xl[r];
      } else {
        const int next = nan_payload<A>(v) - 1;
        A y = Lim<A>::nan();
        if (prev >= 0 && next >= 0) {
          const double slope = div_rn(sub_rn((double)col[next * INTERP_W], f0), sub_rn(xl[next], x0));
          y = (A)add_rn(mul_rn(slope, sub_rn(xl[r], x0)), f0);
        }
        col[r * INTERP_W] = y;
      }
    }
  }
  if (active) {
    A* dst = out + o * R * inner + i;
    for (int r = 0; r < R; ++r) __stcs(dst + (int64_t)r * inner, col[r * INTERP_W]);
  }
}

// The same algorithm for columns too long for a shared-memory tile (R > ~420 float32 / ~210 float64 samples):
// the dense OUTPUT is the working buffer -- adjacent threads own adjacent columns, so every access is coalesced --
// copy, backward pass (index of the next finite sample into the NaN payload of every gap slot; 22 bits: R < 4 Mi),
// forward pass.  Three sweeps over the output instead of one: the slower general path.
template <typename In, typename A>
__global__ void __launch_bounds__(256) k_interp_global(const In* __restrict__ data, int64_t so, int64_t sr, int64_t outer,
                                                       int R, int64_t inner, const double* __restrict__ xs,
                                                       A* __restrict__ out) {
  const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= outer * inner) return;
  const int64_t o = item / inner, i = item - o * inner;
  const In* src = data + o * so + i;
  A* col = out + o * R * inner + i;  // row r at col[r * inner]
  int cnt = 0;
  for (int r = 0; r < R; ++r) {
    const A v = (A)__ldcs(src + (int64_t)r * sr);
    col[(int64_t)r * inner] = v;
    cnt += finite_(v) ? 1 : 0;
  }
  if (cnt < 2 || cnt == R) return;
  int nx = 0;
  for (int r = R - 1; r >= 0; --r) {
    const A v = col[(int64_t)r * inner];
    if (finite_(v))
      nx = r + 1;
    else
      col[(int64_t)r * inner] = nan_code<A>(nx);
  }
  int prev = -1;
  double f0 = 0.0, x0 = 0.0;
  for (int r = 0; r < R; ++r) {
    const A v = col[(int64_t)r * inner];
    if (finite_(v)) {
      prev = r;
      f0 = (double)v;
      x0 = xs ? xs[r] : (double)r;
    } else {
      const int next = nan_payload<A>(v) - 1;
      A y = Lim<A>::nan();
      if (prev >= 0 && next >= 0) {
        const double xn = xs ? xs[next] : (double)next, xr = xs ? xs[r] : (double)r;
        const double slope = div_rn(sub_rn((double)col[(int64_t)next * inner], f0), sub_rn(xn, x0));
        y = (A)add_rn(mul_rn(slope, sub_rn(xr, x0)), f0);
      }
      col[(int64_t)r * inner] = y;
    }
  }
}

template <typename In>
int launch_interp_t(const eoc_view& in, const double* xs, void* out, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  const size_t smem = (size_t)in.R * INTERP_W * sizeof(A) + (size_t)in.R * sizeof(double);
  if (smem > 220 * 1024) {
    if (in.R >= (1 << 22)) EOC_FAIL(EOC_ERR_UNSUPPORTED, "interpolation along an axis of %lld samples", (long long)in.R);
    const int64_t items = in.outer * in.inner;
    const int64_t blocks = (items + 255) / 256;
    if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
    k_interp_global<In, A><<<(unsigned)blocks, 256, 0, s>>>((const In*)in.data, in.stride_outer, in.stride_R, in.outer,
                                                            (int)in.R, in.inner, xs, (A*)out);
    EOC_LAUNCH_CHECK();
    return EOC_OK;
  }
  const int64_t tiles_per_outer = (in.inner + INTERP_W - 1) / INTERP_W;
  const int64_t blocks = tiles_per_outer * in.outer;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  auto kern = k_interp<In, A>;
  EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, INTERP_W, smem, s>>>((const In*)in.data, in.stride_outer, in.stride_R, in.outer, (int)in.R,
                                                 in.inner, xs, (A*)out, tiles_per_outer);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace

int launch_interpolate_linear(const eoc_view& in, const double* xs, void* out, cudaStream_t s) {
  if (in.outer * in.R * in.inner == 0) return EOC_OK;
  if (!out) EOC_FAIL(EOC_ERR_INVALID, "eoc_interpolate_linear: out is NULL");
  if (in.R > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld too large", (long long)in.R);
  switch (in.dtype) {
    case EOC_F32: return launch_interp_t<float>(in, xs, out, s);
    case EOC_F64: return launch_interp_t<double>(in, xs, out, s);
    default: EOC_FAIL(EOC_ERR_UNSUPPORTED, "interpolation is defined for floating-point samples (integers have no gaps)");
  }
}

}  // namespace eoc
