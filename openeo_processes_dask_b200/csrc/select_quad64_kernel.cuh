// Kernel C-Q64 (round 2, late): median / quantiles of FLOAT64 columns of 32 < R <= 128 samples -- k_select_quad's scheme
// for doubles.  numpy computes in float64 whenever a cube went through default-dtype arithmetic, and the only float64
// paths so far were the register kernel (R <= 64) and the K-thread bitonic kernel with its multi-instruction float64
// min / max (255 GB/s at t = 100: 0.04 of the HBM roof, the slowest kernel of the repository).
//   * NaN is counted and replaced by +inf in the register fill, so the network needs no NaN semantics and a comparator
//     is ONE DSETP and four 32-bit selects (p = a > b; lo = p ? b : a; hi = p ? a : b); the valid count is R minus the
//     NaN count, and a pad that lands among the n smallest can only be a real +inf's twin (same value);
//   * four runs of NH doubles parked behind NaN sentinels, the rank elimination of QuadCursor on ld.shared.f64 /
//     min.f64 / 64-bit pattern compares;
//   * optionally G = 2 / 4 threads per column (thread (c, g) sorts the runs g, g + G); one thread per column measured
//     fastest (t = 100: median 3785 / 3669 / 3050 GB/s for G = 1 / 2 / 4).
// Algorithmic bytes per column: R * 8 in, 8 (+ nq * 8) out.
#pragma once

#include "select_quad_kernel.cuh"

namespace eoc {

// the sample type of the float64 raw networks: a comparator that assumes NaN-free operands
struct DblKey {
  double v;
};
template <> EOC_HD void cmpx<true, DblKey>(DblKey& a, DblKey& b) {
  const bool p = a.v > b.v;
  const double lo = p ? b.v : a.v;
  const double hi = p ? a.v : b.v;
  a.v = lo;
  b.v = hi;
}

template <int NH, int RS> struct QuadCursor64 {
  static constexpr uint32_t S = (uint32_t)(NH + 1) * RS;  // bytes between runs
  uint32_t a0, a1, a2, a3;                                 // heads
  uint32_t col;
  int pos;
  __device__ __forceinline__ void reset(uint32_t c) {
    col = c;
    a0 = c;
    a1 = c + S;
    a2 = c + 2 * S;
    a3 = c + 3 * S;
    pos = 0;
  }
  __device__ __forceinline__ void advance(int k) {
    const uint32_t e0 = col + (uint32_t)NH * RS, e1 = e0 + S, e2 = e0 + 2 * S, e3 = e0 + 3 * S;  // sentinels
    int rem = k - pos;
    pos = k;
#pragma unroll 1
    while (rem > 0) {
      const int r1 = rem - 1;
      const int q = r1 >> 2;  // d - 1
      const uint32_t qs = (uint32_t)q * RS, ds = qs + RS;
      asm("{\n\t.reg .pred p0, p1, p2, p01, p012;\n\t.reg .u32 c0, c1, c2, c3;\n\t.reg .f64 v0, v1, v2, v3, m, m2;\n\t"
          "add.u32 c0, %0, %4;\n\t"
          "add.u32 c1, %1, %4;\n\t"
          "add.u32 c2, %2, %4;\n\t"
          "add.u32 c3, %3, %4;\n\t"
          "min.u32 c0, c0, %6;\n\t"
          "min.u32 c1, c1, %7;\n\t"
          "min.u32 c2, c2, %8;\n\t"
          "min.u32 c3, c3, %9;\n\t"
          "ld.shared.f64 v0, [c0];\n\t"
          "ld.shared.f64 v1, [c1];\n\t"
          "ld.shared.f64 v2, [c2];\n\t"
          "ld.shared.f64 v3, [c3];\n\t"
          "min.f64 m2, v2, v3;\n\t"
          "min.f64 m, v0, v1;\n\t"
          "min.f64 m, m, m2;\n\t"
          "setp.eq.b64 p0, v0, m;\n\t"
          "setp.eq.and.b64 p1, v1, m, !p0;\n\t"
          "or.pred p01, p0, p1;\n\t"
          "setp.eq.and.b64 p2, v2, m, !p01;\n\t"
          "or.pred p012, p01, p2;\n\t"
          "@p0 add.u32 %0, %0, %5;\n\t"
          "@p1 add.u32 %1, %1, %5;\n\t"
          "@p2 add.u32 %2, %2, %5;\n\t"
          "@!p012 add.u32 %3, %3, %5;\n\t}"
          : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3)
          : "r"(qs), "r"(ds), "r"(e0), "r"(e1), "r"(e2), "r"(e3));
      rem = r1 - q;
    }
  }
  // x_pos = the smallest head; x_{pos+1} = the smallest once that one is removed (NaN when nothing is left)
  __device__ __forceinline__ void heads(double& xk, double& xk1) const {
    double m, m1;
    asm("{\n\t.reg .pred p0, p1, p2, p01, p012;\n\t.reg .u32 hn;\n\t.reg .f64 v0, v1, v2, v3, m2, nx;\n\t"
        "ld.shared.f64 v0, [%2];\n\t"
        "ld.shared.f64 v1, [%3];\n\t"
        "ld.shared.f64 v2, [%4];\n\t"
        "ld.shared.f64 v3, [%5];\n\t"
        "min.f64 m2, v2, v3;\n\t"
        "min.f64 %0, v0, v1;\n\t"
        "min.f64 %0, %0, m2;\n\t"
        "setp.eq.b64 p0, v0, %0;\n\t"
        "setp.eq.and.b64 p1, v1, %0, !p0;\n\t"
        "or.pred p01, p0, p1;\n\t"
        "setp.eq.and.b64 p2, v2, %0, !p01;\n\t"
        "or.pred p012, p01, p2;\n\t"
        "mov.u32 hn, %5;\n\t"
        "@p0 mov.u32 hn, %2;\n\t"
        "@p1 mov.u32 hn, %3;\n\t"
        "@p2 mov.u32 hn, %4;\n\t"
        "ld.shared.f64 nx, [hn+%6];\n\t"
        "@p0 mov.f64 v0, nx;\n\t"
        "@p1 mov.f64 v1, nx;\n\t"
        "@p2 mov.f64 v2, nx;\n\t"
        "@!p012 mov.f64 v3, nx;\n\t"
        "min.f64 m2, v2, v3;\n\t"
        "min.f64 %1, v0, v1;\n\t"
        "min.f64 %1, %1, m2;\n\t}"
        : "=&d"(m), "=&d"(m1)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "n"(RS));
    xk = m;
    xk1 = m1;
  }
};

template <int NH, bool FULL, int W, int MINB, int G>
__global__ void __launch_bounds__(W * G, MINB)
k_select_quad64(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = double;
  constexpr int KR = 4;
  constexpr int S = NH + 1;  // rows per run in shared memory: NH samples + the sentinel
  static_assert(KR % G == 0, "G divides the number of runs");
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  A* buf = reinterpret_cast<A*>(smem_dyn);  // [KR * S][W]
  int* nanc = reinterpret_cast<int*>(smem_dyn + (size_t)KR * S * W * sizeof(A));  // [G][W] NaN counts of the threads' runs
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int tid = threadIdx.x;
  const int t = G > 1 ? tid % W : tid;
  const int g = G > 1 ? tid / W : 0;
  const int R = (int)P.R;
  A* mycol = buf + t;
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < T.num_tiles ? (T.num_tiles - first + step - 1) / step : 0;

  auto coords = [&](int64_t k, int64_t& o, int64_t& i0) {
    const int64_t tl = first + k * step;
    o = tl / T.tiles_per_outer;
    i0 = (tl - o * T.tiles_per_outer) * W;
  };
  auto issue = [&](int64_t k) {  // thread 0 only: one box of NH rows per run (rows past R are zero-filled)
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
        const A x = (FULL || j < left) ? run[j * W] : Lim<A>::inf();  // an unused slot is a pad that is not counted
        const bool isn = x != x;
        mynan += isn ? 1 : 0;
        v[j].v = isn ? Lim<A>::inf() : x;
      }
      sortraw<NH, DblKey>(v, 1);
#pragma unroll
      for (int j = 0; j < NH; ++j) run[j * W] = v[j].v;
    }
    if constexpr (G > 1) {
      nanc[g * W + t] = mynan;
      __syncthreads();  // the other runs of the column were sorted by other threads
    }

    if (g == 0 && active) {
      int n = R - mynan;
      if constexpr (G > 1) {
#pragma unroll
        for (int gg = 1; gg < G; ++gg) n -= nanc[gg * W + t];
      }
      const uint32_t col = smem_u32(mycol);
      const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
      QuadCursor64<NH, W * 8> cur;
      cur.reset(col);
      if (P.median_out) {
        A m = Lim<A>::nan();
        if (!poisoned) {
          A lo, hi;
          cur.advance((n - 1) >> 1);
          cur.heads(lo, hi);
          m = (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
        }
        ((A*)P.median_out)[o * P.inner + i] = m;
      }
      for (int q = 0; q < P.nq; ++q) {
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
