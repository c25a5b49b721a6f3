// Kernel C-M (round 2, late): median / quantiles of LONG float32 columns (128 < R <= 512: a year of daily scenes) as
// EIGHT sorted runs with NaN sentinels -- k_select_quad's scheme with K = 8 and several threads per column.
//
// The K-run split kernel (select_split_wide.cuh) sorts 2 .. 4 runs of 100 / 128 samples: 2 S(100) = 2208 comparators
// for R = 200, 3 S(128) = 4413 for R = 365, and 128-165 registers.  Eight short runs cost 8 S(25) = 1120 (R <= 200),
// 8 S(32) = 1528 (R <= 256), 8 S(50) = 3224 (R <= 400), 8 S(64) = 4344 (R <= 512), sorted by the raw networks (no pad
// substitution, no canonicalisation pass), and the ranks come from the same sentinel-based elimination as in
// k_select_quad, eight candidates per step (MultiCursor8: ~50 instructions per step, ~25 steps for the median of 200).
// G threads share a column: thread (c, g) sorts the runs g, g + G, ... of column c (a warp = 32 adjacent columns of
// one g: conflict-free, and G times the resident warps the shared memory per column would allow), thread (c, 0)
// then finds the valid counts and the ranks.
// Algorithmic bytes per column: R * 4 in, 4 (median) + nq * 8 (quantiles) out.
#pragma once

#include "select_quad_kernel.cuh"

namespace eoc {

template <int NH, int RS> struct MultiCursor8 {
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
          "{\n\t.reg .pred p, o;\n\t.reg .u32 c0, c1, c2, c3, c4, c5, c6, c7;\n\t.reg .f32 v0, v1, v2, v3, v4, v5, v6, v7, m, m1, m2;\n\t"
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
          "ld.shared.f32 v0, [c0];\n\t"
          "ld.shared.f32 v1, [c1];\n\t"
          "ld.shared.f32 v2, [c2];\n\t"
          "ld.shared.f32 v3, [c3];\n\t"
          "ld.shared.f32 v4, [c4];\n\t"
          "ld.shared.f32 v5, [c5];\n\t"
          "ld.shared.f32 v6, [c6];\n\t"
          "ld.shared.f32 v7, [c7];\n\t"
          "min.f32 m1, v0, v1, v2;\n\t"
          "min.f32 m2, v3, v4, v5;\n\t"
          "min.f32 m1, m1, v6, v7;\n\t"
          "min.f32 m, m1, m2;\n\t"
          "setp.eq.b32 o, v0, m;\n\t"
          "@o add.u32 %0, %0, %9;\n\t"
          "setp.eq.and.b32 p, v1, m, !o;\n\t"
          "@p add.u32 %1, %1, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b32 p, v2, m, !o;\n\t"
          "@p add.u32 %2, %2, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b32 p, v3, m, !o;\n\t"
          "@p add.u32 %3, %3, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b32 p, v4, m, !o;\n\t"
          "@p add.u32 %4, %4, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b32 p, v5, m, !o;\n\t"
          "@p add.u32 %5, %5, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "setp.eq.and.b32 p, v6, m, !o;\n\t"
          "@p add.u32 %6, %6, %9;\n\t"
          "or.pred o, o, p;\n\t"
          "@!o add.u32 %7, %7, %9;\n\t}"
          : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7])
          : "r"(qs), "r"(ds), "r"(e0), "r"(e0 + S), "r"(e0 + 2 * S), "r"(e0 + 3 * S), "r"(e0 + 4 * S), "r"(e0 + 5 * S),
            "r"(e0 + 6 * S), "r"(e0 + 7 * S));
      rem = r1 - q;
    }
  }
  // x_pos = the smallest head; x_{pos+1} = the smallest once that one is removed (NaN when nothing is left)
  __device__ __forceinline__ void heads(float& xk, float& xk1) const {
    float m, mn;
    asm(
        "{\n\t.reg .pred p, o;\n\t.reg .u32 hn;\n\t.reg .f32 v0, v1, v2, v3, v4, v5, v6, v7, m1, m2, nx;\n\t"
        "ld.shared.f32 v0, [%2];\n\t"
        "ld.shared.f32 v1, [%3];\n\t"
        "ld.shared.f32 v2, [%4];\n\t"
        "ld.shared.f32 v3, [%5];\n\t"
        "ld.shared.f32 v4, [%6];\n\t"
        "ld.shared.f32 v5, [%7];\n\t"
        "ld.shared.f32 v6, [%8];\n\t"
        "ld.shared.f32 v7, [%9];\n\t"
        "min.f32 m1, v0, v1, v2;\n\t"
        "min.f32 m2, v3, v4, v5;\n\t"
        "min.f32 m1, m1, v6, v7;\n\t"
        "min.f32 %0, m1, m2;\n\t"
        "mov.u32 hn, %9;\n\t"
        "setp.eq.b32 o, v0, %0;\n\t"
        "@o mov.u32 hn, %2;\n\t"
        "@o mov.f32 v0, 0f7FFFFFFF;\n\t"
        "setp.eq.and.b32 p, v1, %0, !o;\n\t"
        "@p mov.u32 hn, %3;\n\t"
        "@p mov.f32 v1, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b32 p, v2, %0, !o;\n\t"
        "@p mov.u32 hn, %4;\n\t"
        "@p mov.f32 v2, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b32 p, v3, %0, !o;\n\t"
        "@p mov.u32 hn, %5;\n\t"
        "@p mov.f32 v3, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b32 p, v4, %0, !o;\n\t"
        "@p mov.u32 hn, %6;\n\t"
        "@p mov.f32 v4, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b32 p, v5, %0, !o;\n\t"
        "@p mov.u32 hn, %7;\n\t"
        "@p mov.f32 v5, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "setp.eq.and.b32 p, v6, %0, !o;\n\t"
        "@p mov.u32 hn, %8;\n\t"
        "@p mov.f32 v6, 0f7FFFFFFF;\n\t"
        "or.pred o, o, p;\n\t"
        "@!o mov.f32 v7, 0f7FFFFFFF;\n\t"
        "ld.shared.f32 nx, [hn+%10];\n\t"
        "min.f32 m1, v0, v1, v2;\n\t"
        "min.f32 m2, v3, v4, v5;\n\t"
        "min.f32 m1, m1, v6, v7;\n\t"
        "min.f32 m2, m2, nx;\n\t"
        "min.f32 %1, m1, m2;\n\t}"
        : "=&f"(m), "=&f"(mn)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "n"(RS));
    xk = m;
    xk1 = mn;
  }
};

// In != float: raw uint16 / int16 / uint8 samples, landing zone + exact conversion + barrier as in k_select_quad.
template <int NH, int G, int W, int MINB, bool FULL, typename In = float>
__global__ void __launch_bounds__(W * G, MINB)
k_select_multi(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = float;
  constexpr int KR = 8;
  constexpr int S = NH + 1;  // rows per run in shared memory: NH samples + the sentinel
  constexpr bool RAWINT = !std::is_same<In, float>::value;
  constexpr int LAND = RAWINT ? (S * W * (int)sizeof(A) - NH * W * (int)sizeof(In)) / 128 * 128 : 0;  // byte offset of a run's TMA box
  static_assert(LAND % 128 == 0, "the landing zone of a TMA box is 128-byte aligned");
  static_assert(KR % G == 0, "G divides the number of runs");
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  A* buf = reinterpret_cast<A*>(smem_dyn);  // [KR * S][W]
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int tid = threadIdx.x;
  const int t = tid % W;  // column of the tile
  const int g = tid / W;  // this thread sorts the runs g, g + G, ...
  const int R = (int)P.R;
  A* mycol = buf + t;  // row j of this thread's column: mycol[j * W]
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
      if (r * NH < R)
        tma_load_box(reinterpret_cast<unsigned char*>(buf + r * S * W) + LAND, &tmap, &full, (int32_t)i0, r * NH, (int32_t)o);
  };

  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // the sentinels: written once, never touched by TMA (its boxes end one row short) nor by the park
  for (int r = g; r < KR; r += G) mycol[(r * S + NH) * W] = Lim<A>::nan();
  __syncthreads();
  if (tid == 0 && n_my > 0) issue(0);
  const int one = one_from_zero(P.zero);

  for (int64_t k = 0; k < n_my; ++k) {
    int64_t o, i0;
    coords(k, o, i0);
    const int64_t i = i0 + t;
    const bool active = i < P.inner;
    if (tid == 0 && k + 1 < n_my) {  // warm L2 with the next tile; its TMA is issued once this tile is done with the buffer
      int64_t o2, i2;
      coords(k + 1, o2, i2);
#pragma unroll
      for (int r = 0; r < KR; ++r)
        if (r * NH < R) tma_prefetch_box(&tmap, (int32_t)i2, r * NH, (int32_t)o2);
    }
    mbar_wait(&full, (uint32_t)(k & 1));

#pragma unroll 1
    for (int h = g; h < KR; h += G) {
      A* run = mycol + h * S * W;
      const int left = R - h * NH;  // rows of this run that hold samples (FULL: all NH)
      A v[NH];
#pragma unroll
      for (int j = 0; j < NH; ++j) {
        if constexpr (RAWINT) {
          const In* land = reinterpret_cast<const In*>(reinterpret_cast<const unsigned char*>(buf + h * S * W) + LAND) + t;
          v[j] = (FULL || j < left) ? quad_sample_to_float<In>(land[j * W]) : Lim<A>::nan();
        } else {
          v[j] = (FULL || j < left) ? run[j * W] : Lim<A>::nan();
        }
      }
      if constexpr (RAWINT) __syncthreads();  // every thread holds its run: the park may overwrite the landing zones
      sortraw<NH, A>(v, one);
#pragma unroll
      for (int j = 0; j < NH; ++j) run[j * W] = v[j];
      if constexpr (RAWINT) run[NH * W] = Lim<A>::nan();  // the landing zone covered the sentinel
    }
    if constexpr (G > 1) __syncthreads();  // the other runs of the column were sorted by other threads

    // thread (c, 0) finds the median; the probabilities are dealt out over the G threads of the column (each with a
    // cursor of its own, from the start: different warps, so the ranks are searched side by side)
    if (g == 0 || g < P.nq + (P.median_out ? 1 : 0)) {
      const uint32_t col = smem_u32(mycol);
      int n = 0;
      if constexpr (RAWINT) {
        n = R;  // integers hold no NaN
      } else {
#pragma unroll
        for (int r = 0; r < KR; ++r) n += quad_valid_prefix_dev<NH, W * 4>(col, r);
      }
      if (active) {
        const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
        MultiCursor8<NH, W * 4> cur;
        cur.reset(col);
        if (g == 0 && P.median_out) {
          A m = Lim<A>::nan();
          if (!poisoned) {
            A lo, hi;
            cur.advance((n - 1) >> 1);
            cur.heads(lo, hi);
            m = (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
          }
          store_median<A>(P, o * P.inner + i, m);
        }
        // with a median: thread g >= 1 takes probability g - 1, g - 1 + (G - 1), ...; without: thread g takes g, g + G, ...
        const int q0 = P.median_out ? g - 1 : g, qstep = P.median_out ? G - 1 : G;
        for (int q = q0; q >= 0 && q < P.nq; q += qstep) {
          double r = Lim<double>::nan();
          if (!poisoned) {
            int ip, in_;
            double gamma;
            quantile_ranks(n, P.probs[q], ip, in_, gamma);
            if (ip < cur.pos) cur.reset(col);  // probabilities not ascending: start over
            A x0, x1;
            cur.advance(ip);
            cur.heads(x0, x1);
            r = quantile_lerp<A>(x0, in_ == ip ? x0 : x1, gamma, P.lerp_wrap16 != 0);
          }
          store_quant<A>(P, o, q, i, r);
        }
      }
    }
    // the park wrote the buffer through the generic proxy; order those writes before the TMA (async proxy)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();  // all rank fetches done: the buffer may be overwritten by the next tile
    if (tid == 0 && k + 1 < n_my) issue(k + 1);
  }
}

}  // namespace eoc
