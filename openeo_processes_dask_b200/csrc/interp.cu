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
#include <cstdlib>
#include <type_traits>

#include "column_launch.cuh"
#include "tile_kernel.cuh"
#include "tma_host.cuh"

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

// double from a small non-negative integer without a conversion instruction: 2^52 + k is exact, minus 2^52
__device__ __forceinline__ double small_int_to_double(int k) {
  return __dsub_rn(__hiloint2double(0x43300000, k), 4503599627370496.0);
}

// the gaps of one column of a shared-memory tile, in place (row r at col[r * INTERP_W]); xl = positions, or with IDX
// the reciprocal table RN(1 / d)
// COUNT (the TMA kernel, which has no load loop to count in): the backward pass counts the finite samples itself and
// keeps an infinity's sign in payload bits 20 / 21, so a column that turns out to have fewer than two finite samples
// -- it must be returned as it is -- gets its infinities back (rare; the row index uses bits 0 .. 19: R <= 430 here).
template <typename A, bool IDX, bool COUNT = false>
__device__ __forceinline__ void fill_gaps(A* col, int R, const double* xl) {
  // Backward pass: every non-finite slot receives the index of the next finite sample, packed into a NaN
  // payload (index + 1; 0 = none), so that the forward pass is the same instruction stream for every lane
  // of the warp whatever the gap pattern (a per-gap inner loop diverged badly: 447 GB/s).
  int nx = 0, cnt = 0;
#pragma unroll 4
  for (int r = R - 1; r >= 0; --r) {
    const A v = col[r * INTERP_W];
    const bool fin = finite_(v);
    int code = nx;
    if constexpr (COUNT) {
      cnt += fin ? 1 : 0;
      code |= (v == Lim<A>::inf()) ? (1 << 20) : 0;
      code |= (v == -Lim<A>::inf()) ? (1 << 21) : 0;
    }
    nx = fin ? r + 1 : nx;
    if (!fin) col[r * INTERP_W] = nan_code<A>(code);
  }
  if constexpr (COUNT) {
    if (cnt == R) return;  // nothing missing: nothing was written
    if (cnt < 2) {         // returned as it is: give the infinities back
      for (int r = 0; r < R; ++r) {
        const A v = col[r * INTERP_W];
        if (!finite_(v)) {
          const int pl = nan_payload<A>(v);
          col[r * INTERP_W] = (pl & (1 << 20)) ? Lim<A>::inf() : ((pl & (1 << 21)) ? -Lim<A>::inf() : Lim<A>::nan());
        }
      }
      return;
    }
  }
  int prev = -1;
  double f0 = 0.0, x0 = 0.0;
  if constexpr (IDX) {
#pragma unroll 4
    for (int r = 0; r < R; ++r) {
      const A v = col[r * INTERP_W];
      const bool fin = finite_(v);
      const int next = (nan_payload<A>(v) & 0xFFFFF) - 1;
      const bool ok = !fin && prev >= 0 && next >= 0;
      const int nidx = ok ? next : r;
      const double dy = sub_rn((double)col[nidx * INTERP_W], f0);
      const int d = ok ? nidx - prev : 1;
      const double rcp = xl[d], b = small_int_to_double(d);
      const double q = mul_rn(dy, rcp);
      double slope = __fma_rn(__fma_rn(-b, q, dy), rcp, q);
      if constexpr (sizeof(A) == 8) {
        // float64 cubes: differences near the ends of the double range leave the theorem's premises
        const double ady = fabs(dy);
        if (!(ady >= 1e-270 && ady <= 1e270) && dy != 0.0) slope = div_rn(dy, b);
      }
      const A y = (A)add_rn(mul_rn(slope, small_int_to_double(ok ? r - prev : 0)), f0);
      if (!fin) col[r * INTERP_W] = ok ? y : Lim<A>::nan();
      prev = fin ? r : prev;
      f0 = fin ? (double)v : f0;
    }
  } else {
    for (int r = 0; r < R; ++r) {
      const A v = col[r * INTERP_W];
      if (finite_(v)) {
        prev = r;
        f0 = (double)v;
        x0 = xl[r];
      } else {
        const int next = (nan_payload<A>(v) & 0xFFFFF) - 1;
        A y = Lim<A>::nan();
        if (prev >= 0 && next >= 0) {
          const double slope = div_rn(sub_rn((double)col[next * INTERP_W], f0), sub_rn(xl[next], x0));
          y = (A)add_rn(mul_rn(slope, sub_rn(xl[r], x0)), f0);
        }
        col[r * INTERP_W] = y;
      }
    }
  }
}

// IDX: the sample positions are the indices 0 .. R-1 (what apply_dimension passes: no labels).  The position
// differences are then small integers, and the division of np.interp's slope -- ~35 instructions of double-precision
// Newton iteration per row of every warp that holds a gap slot, which bounded the first version at 0.27 of the HBM
// roof (ncu r1f) -- becomes three: with y = RN(1 / b) from a per-CTA table, q = RN(a y), r = a - b q (exact, one
// FMA), q' = RN(q + r y) IS the correctly rounded a / b (Markstein's theorem; checked against exact rational
// arithmetic for 4 x 10^5 differences of float32 / float64 samples over b = 1 .. 511, and bit for bit against the
// reference's outputs by the golden tests).  The forward pass is branch-free: every row computes the gap value with
// clamped indices and selects.
template <typename In, typename A, bool IDX = false>
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
  // IDX: xl[d] = RN(1 / d), the reciprocal table (xl[0] unused); else the label positions
  for (int r = t; r < R; r += INTERP_W) xl[r] = IDX ? __ddiv_rn(1.0, (double)(r > 0 ? r : 1)) : (xs ? xs[r] : (double)r);
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
  if (cnt >= 2 && cnt < R) fill_gaps<A, IDX>(col, R, xl);
  if (active) {
    A* dst = out + o * R * inner + i;
    for (int r = 0; r < R; ++r) __stcs(dst + (int64_t)r * inner, col[r * INTERP_W]);
  }
}

// TMA in, TMA out (round 2): the first version's load and store phases -- 100 strided LDG / STG per thread, ten in
// flight -- left the kernel latency-bound (issue slots 48 % busy, DRAM 21 %).  Here a persistent CTA receives the
// (R x 128) tile by cp.async.bulk.tensor like the selection kernels, every thread fills the gaps of its column in
// place, and ONE thread sends the tile back with a bulk tensor STORE (clipped to the tensor, so a ragged last tile
// needs no predicates); the next tile's load is issued once that store has finished reading shared memory.
// Several CTAs per SM alternate their load / compute / store phases.
template <typename A, bool IDX>
__global__ void __launch_bounds__(INTERP_W)
k_interp_tma(const __grid_constant__ CUtensorMap tin, const __grid_constant__ CUtensorMap tout, int R, int box_r, int nbox,
             int64_t inner, int64_t tiles_per_outer, int64_t num_tiles, const double* __restrict__ xs) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  A* tile = reinterpret_cast<A*>(smem_raw);  // [nbox * box_r][W], the first R rows are samples
  double* xl = reinterpret_cast<double*>(smem_raw + (size_t)nbox * box_r * INTERP_W * sizeof(A));
  __shared__ uint64_t full;
  const int t = threadIdx.x;
  const uint32_t tile_bytes = (uint32_t)((size_t)nbox * box_r * INTERP_W * sizeof(A));
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < num_tiles ? (num_tiles - first + step - 1) / step : 0;
  auto issue = [&](int64_t k) {  // thread 0 only
    const int64_t tl = first + k * step;
    const int64_t o = tl / tiles_per_outer, i0 = (tl - o * tiles_per_outer) * INTERP_W;
    mbar_expect_tx(&full, tile_bytes);
    for (int b = 0; b < nbox; ++b)
      tma_load_box(tile + (size_t)b * box_r * INTERP_W, &tin, &full, (int32_t)i0, b * box_r, (int32_t)o);
  };
  if (t == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tin)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tout)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int r = t; r < R; r += INTERP_W) xl[r] = IDX ? __ddiv_rn(1.0, (double)(r > 0 ? r : 1)) : xs[r];
  __syncthreads();
  if (t == 0 && n_my > 0) issue(0);
  A* col = tile + t;
  for (int64_t k = 0; k < n_my; ++k) {
    const int64_t tl = first + k * step;
    const int64_t o = tl / tiles_per_outer, i0 = (tl - o * tiles_per_outer) * INTERP_W;
    mbar_wait(&full, (uint32_t)(k & 1));
    fill_gaps<A, IDX, true>(col, R, xl);
    // the fills went through the generic proxy; order them before the bulk store (async proxy) reads the tile
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (t == 0) {
      for (int b = 0; b < nbox; ++b) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                         reinterpret_cast<uint64_t>(&tout)),
                     "r"(smem_u32(tile + (size_t)b * box_r * INTERP_W)), "r"((int32_t)i0), "r"(b * box_r), "r"((int32_t)o)
                     : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the store has read the tile: it may be refilled
      if (k + 1 < n_my) issue(k + 1);
    }
  }
  if (t == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
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
  // columns too long for a shared-memory tile, or an `inner` so short that the 128-column tiles of ONE outer slice would
  // be mostly empty (the fixture layout (x, y, t, bands): inner = 4): one thread per column over global memory
  const char* te = getenv("EOC_INTERP_TMA");
  const bool short_inner = !(te && te[0]) && !tma_tiles_fill(in, INTERP_W);
  if (smem > 220 * 1024 || short_inner) {
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
  // TMA in / TMA out when both tensors are addressable by it (16-byte aligned bases and row pitches) and the cube is
  // large enough for a persistent grid; EOC_INTERP_TMA=0 / 1 forces the choice (tests, A/B runs)
  {
    eoc_view ov = in;
    ov.data = out;
    ov.stride_outer = in.R * in.inner;
    ov.stride_R = in.inner;
    const char* e = getenv("EOC_INTERP_TMA");
    const bool want = e && e[0] ? e[0] == '1' : blocks >= 4 * (int64_t)sm_count();
    if (want && !(e && e[0] == '0') && tma_view_ok(in) && tma_view_ok(ov)) {
      const int nbox = (int)((in.R + 255) / 256);
      const int box_r = (int)((in.R + nbox - 1) / nbox);
      const size_t tsmem = (size_t)nbox * box_r * INTERP_W * sizeof(A) + (size_t)in.R * sizeof(double);
      if (tsmem <= 220 * 1024) {
        CUtensorMap tin, tout;
        if (int rc = tma_make_map(in, INTERP_W, box_r, &tin)) return rc;
        if (int rc = tma_make_map(ov, INTERP_W, box_r, &tout)) return rc;
        auto tk = xs ? k_interp_tma<A, false> : k_interp_tma<A, true>;
        EOC_CUDA(cudaFuncSetAttribute(tk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
        int per_sm = 0;
        EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tk, INTERP_W, tsmem));
        if (per_sm < 1) per_sm = 1;
        int64_t grid = (int64_t)sm_count() * per_sm;
        if (grid > blocks) grid = blocks;
        tk<<<(unsigned)grid, INTERP_W, tsmem, s>>>(tin, tout, (int)in.R, box_r, nbox, in.inner, tiles_per_outer, blocks, xs);
        EOC_LAUNCH_CHECK();
        return EOC_OK;
      }
    }
  }
  auto kern = xs ? k_interp<In, A, false> : k_interp<In, A, true>;
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
