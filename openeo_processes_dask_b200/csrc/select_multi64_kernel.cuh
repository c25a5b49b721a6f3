// Kernel C-M64 (round 2, late): median / quantiles of long FLOAT64 columns (128 < R <= 512) -- k_select_multi's eight
// sentinel runs with k_select_quad64's arithmetic (NaN counted and padded with +inf in the register fill, one-DSETP
// comparator, 64-bit rank elimination).
#pragma once

#include "select_quad64_kernel.cuh"

namespace eoc {

template <int NH, int RS> struct MultiCursor8x64 {
  static constexpr uint32_t S = (uint32_t)(NH + 1) * RS;  // bytes between runs
  uint32_t a[8];                                           // heads
  uint32_t col;
  int pos;
  __device__ __forceinline__ void reset(uint32_t c) {
    col = c;
#pragma unroll
    for (int r = 0; r < 8; ++r) a[r] = c + r * S;
    pos = 0;
  }
  __device__ __forceinline__ void advance(int k) {
    const uint32_t e0 = col + (uint32_t)NH * RS;  // sentinel of run 0; run r: e0 + r * S
    int rem = k - pos;
    pos = k;
#pragma unroll 1
    while (rem > 0) {
      const int r1 = rem - 1;
      const int q = r1 >> 3;  // d - 1, d = ceil(rem / 8)
      const uint32_t qs = (uint32_t)q * RS, ds = qs + RS;
      asm(
          "{\n\t.reg .pred p, o;\n\t.reg .u32 c0, c1, c2, c3, c4, c5, c6, c7;\n\t.reg .f64 v0, v1, v2, v3, v4, v5, v6, v7, m, q0, q1, q2, q3, q4, q5, q6, q7;\n\t"
          "add.u32 c0, %0, %8;\n\t"
          "add.u32 c1, %1, %8;\n\t"
          "add.u32 c2, %2, %8;\n\t"
          "add.u32 c3, %3, %8;\n\t"
          "add.u32 c4, %4, %8;\n\t"
          "add.u32 c5, %5, %8;\n\t"
          "add.u32 c6, %6, %8;\n\t"
          "add.u32 c7, %7, %8;\n\t"
          "min.u32 c0, c0, %10;\n\t"
          "min.u32 c1, c1, %11;\n\t"
          "min.u32 c2, c2, %12;\n\t"
          "min.u32 c3, c3, %13;\n\t"
          "min.u32 c4, c4, %14;\n\t"
          "min.u32 c5, c5, %15;\n\t"
          "min.u32 c6, c6, %16;\n\t"
          "min.u32 c7, c7, %17;\n\t"
          "ld.shared.f64 v0, [c0];\n\t"
          "ld.shared.f64 v1, [c1];\n\t"
          "ld.shared.f64 v2, [c2];\n\t"
          "ld.shared.f64 v3, [c3];\n\t"
          "ld.shared.f64 v4, [c4];\n\t"
          "ld.shared.f64 v5, [c5];\n\t"
          "ld.shared.f64 v6, [c6];\n\t"
          "ld.shared.f64 v7, [c7];\n\t"
          "min.f64 q0, v0, v1;\n\t"
          "min.f64 q1, v2, v3;\n\t"
          "min.f64 q2, v4, v5;\n\t"
          "min.f64 q3, v6, v7;\n\t"
          "min.f64 q4, q0, q1;\n\t"
          "min.f64 q5, q2, q3;\n\t"
          "min.f64 m, q4, q5;\n\t"
          "setp.eq.b64 o, v0, m;\n\t"
          "@o add.u32 %0, %0, %9;\n\t"
          "setp.eq.and.b64 p, v1, m, !o;\n\t"
          "@p add.u32 %1, %1, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b64 p, v2, m, !o;\n\t"
          "@p add.u32 %2, %2, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b64 p, v3, m, !o;\n\t"
          "@p add.u32 %3, %3, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b64 p, v4, m, !o;\n\t"
          "@p add.u32 %4, %4, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b64 p, v5, m, !o;\n\t"
          "@p add.u32 %5, %5, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b64 p, v6, m, !o;\n\t"
          "@p add.u32 %6, %6, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "@!o add.u32 %7, %7, %9;\n\t}"
          : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7])
          : "r"(qs), "r"(ds), "r"(e0), "r"(e0 + S), "r"(e0 + 2 * S), "r"(e0 + 3 * S), "r"(e0 + 4 * S), "r"(e0 + 5 * S),
            "r"(e0 + 6 * S), "r"(e0 + 7 * S));
      rem = r1 - q;
    }
  }
  __device__ __forceinline__ void heads(double& xk, double& xk1) const {
    double m, mn;
    asm(
        "{\n\t.reg .pred p, o;\n\t.reg .u32 hn;\n\t.reg .f64 v0, v1, v2, v3, v4, v5, v6, v7, nx, q0, q1, q2, q3, q4, q5, q6, q7;\n\t"
        "ld.shared.f64 v0, [%2];\n\t"
        "ld.shared.f64 v1, [%3];\n\t"
        "ld.shared.f64 v2, [%4];\n\t"
        "ld.shared.f64 v3, [%5];\n\t"
        "ld.shared.f64 v4, [%6];\n\t"
        "ld.shared.f64 v5, [%7];\n\t"
        "ld.shared.f64 v6, [%8];\n\t"
        "ld.shared.f64 v7, [%9];\n\t"
        "min.f64 q0, v0, v1;\n\t"
        "min.f64 q1, v2, v3;\n\t"
        "min.f64 q2, v4, v5;\n\t"
        "min.f64 q3, v6, v7;\n\t"
        "min.f64 q4, q0, q1;\n\t"
        "min.f64 q5, q2, q3;\n\t"
        "min.f64 %0, q4, q5;\n\t"
        "mov.u32 hn, %9;\n\t"
        "setp.eq.b64 o, v0, %0;\n\t"
        "@o mov.u32 hn, %2;\n\t"
        "@o mov.f64 v0, 0d7FFFFFFFFFFFFFFF;\n\t"
        "setp.eq.and.b64 p, v1, %0, !o;\n\t"
        "@p mov.u32 hn, %3;\n\t"
        "@p mov.f64 v1, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b64 p, v2, %0, !o;\n\t"
        "@p mov.u32 hn, %4;\n\t"
        "@p mov.f64 v2, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b64 p, v3, %0, !o;\n\t"
        "@p mov.u32 hn, %5;\n\t"
        "@p mov.f64 v3, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b64 p, v4, %0, !o;\n\t"
        "@p mov.u32 hn, %6;\n\t"
        "@p mov.f64 v4, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b64 p, v5, %0, !o;\n\t"
        "@p mov.u32 hn, %7;\n\t"
        "@p mov.f64 v5, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b64 p, v6, %0, !o;\n\t"
        "@p mov.u32 hn, %8;\n\t"
        "@p mov.f64 v6, 0d7FFFFFFFFFFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "@!o mov.f64 v7, 0d7FFFFFFFFFFFFFFF;\n\t"
        "ld.shared.f64 nx, [hn+%10];\n\t"
        "min.f64 q0, v0, v1;\n\t"
        "min.f64 q1, v2, v3;\n\t"
        "min.f64 q2, v4, v5;\n\t"
        "min.f64 q3, v6, v7;\n\t"
        "min.f64 q4, q0, q1;\n\t"
        "min.f64 q5, q2, q3;\n\t"
        "min.f64 q6, q4, q5;\n\t"
        "min.f64 %1, q6, nx;\n\t}"
        : "=&d"(m), "=&d"(mn)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "n"(RS));
    xk = m;
    xk1 = mn;
  }
};

template <int NH, int G, int W, int MINB, bool FULL>
__global__ void __launch_bounds__(W * G, MINB)
k_select_multi64(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = double;
  constexpr int KR = 8;
  constexpr int S = NH + 1;
  static_assert(KR % G == 0, "G divides the number of runs");
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  A* buf = reinterpret_cast<A*>(smem_dyn);  // [KR * S][W]
  int* nanc = reinterpret_cast<int*>(smem_dyn + (size_t)KR * S * W * sizeof(A));  // [G][W]
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int tid = threadIdx.x;
  const int t = tid % W;
  const int g = tid / W;
  const int R = (int)P.R;
  A* mycol = buf + t;
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < T.num_tiles ? (T.num_tiles - first + step - 1) / step : 0;

  auto coords = [&](int64_t k, int64_t& o, int64_t& i0) {
    const int64_t tl = first + k * step;
    o = tl / T.tiles_per_outer;
    i0 = (tl - o * T.tiles_per_outer) * W;
  };
  auto issue = [&](int64_t k) {
    int64_t o, i0;
    coords(k, o, i0);
    mbar_expect_tx(&full, T.tile_bytes);
#pragma unroll
    for (int r = 0; r < KR; ++r)
      if (r * NH < R) tma_load_box(buf + r * S * W, &tmap, &full, (int32_t)i0, r * NH, (int32_t)o);
  };

  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int r = g; r < KR; r += G) mycol[(r * S + NH) * W] = Lim<A>::nan();
  __syncthreads();
  if (tid == 0 && n_my > 0) issue(0);

  for (int64_t k = 0; k < n_my; ++k) {
    int64_t o, i0;
    coords(k, o, i0);
    const int64_t i = i0 + t;
    const bool active = i < P.inner;
    if (tid == 0 && k + 1 < n_my) {
      int64_t o2, i2;
      coords(k + 1, o2, i2);
#pragma unroll
      for (int r = 0; r < KR; ++r)
        if (r * NH < R) tma_prefetch_box(&tmap, (int32_t)i2, r * NH, (int32_t)o2);
    }
    mbar_wait(&full, (uint32_t)(k & 1));

    int mynan = 0;
#pragma unroll 1
    for (int h = g; h < KR; h += G) {
      A* run = mycol + h * S * W;
      const int left = R - h * NH;
      DblKey v[NH];
#pragma unroll
      for (int j = 0; j < NH; ++j) {
        const A x = (FULL || j < left) ? run[j * W] : Lim<A>::inf();
        const bool isn = x != x;
        mynan += isn ? 1 : 0;
        v[j].v = isn ? Lim<A>::inf() : x;
      }
      sortraw<NH, DblKey>(v, 1);
#pragma unroll
      for (int j = 0; j < NH; ++j) run[j * W] = v[j].v;
    }
    nanc[g * W + t] = mynan;
    __syncthreads();

    if (active && (g == 0 || g < P.nq + (P.median_out ? 1 : 0))) {
      int n = R;
#pragma unroll
      for (int gg = 0; gg < G; ++gg) n -= nanc[gg * W + t];
      const uint32_t col = smem_u32(mycol);
      const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
      MultiCursor8x64<NH, W * 8> cur;
      cur.reset(col);
      if (g == 0 && P.median_out) {
        A m = Lim<A>::nan();
        if (!poisoned) {
          A lo, hi;
          cur.advance((n - 1) >> 1);
          cur.heads(lo, hi);
          m = (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
        }
        ((A*)P.median_out)[o * P.inner + i] = m;
      }
      const int q0 = P.median_out ? g - 1 : g, qstep = P.median_out ? (G > 1 ? G - 1 : 1) : G;
      for (int q = (G == 1 ? 0 : q0); q >= 0 && q < P.nq; q += (G == 1 ? 1 : qstep)) {
        double r = Lim<double>::nan();
        if (!poisoned) {
          int ip, in_;
          double gamma;
          quantile_ranks(n, P.probs[q], ip, in_, gamma);
          if (ip < cur.pos) cur.reset(col);
          A x0, x1;
          cur.advance(ip);
          cur.heads(x0, x1);
          r = quantile_lerp<A>(x0, in_ == ip ? x0 : x1, gamma, false);
        }
        store_quant<A>(P, o, q, i, r);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0 && k + 1 < n_my) issue(k + 1);
  }
}

}  // namespace eoc
