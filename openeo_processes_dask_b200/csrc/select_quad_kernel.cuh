// Kernel C-Q (round 2, late): median / quantiles of a float32 column of 64 < R <= 128 samples as FOUR short sorted runs.
//
// ncu on the pruned median cone of k_select_tile (profiles/r2_median_raw.csv): 3016 instructions per 100-sample
// column, ALU pipe 86 % / issue slots 82 % busy -- instruction-bound at 0.67 of the HBM peak.  Fewer instructions
// per column is the only lever, so the algorithm changes once more:
//
//   * the column is sorted as 4 runs of NH = 25 (R <= 100) or 32 samples by ONE raw network body (sortraw_NH,
//     140 / 191 comparators) in a rolled loop: 4 S(25) = 560 comparators instead of 898 (cone) or 806 (two runs);
//   * no pad substitution and no canonicalisation pass: the first comparator a wire meets is the pure NaN-native
//     one (min / max.NaN return a sample or the canonical NaN), the later ones may take the mixed-pipe form;
//   * each run is parked back into shared memory with a NaN SENTINEL row behind it (the TMA tile is brought in
//     as one box per run, NH + 1 rows apart), so the rank elimination over the four parked runs
//     (select_quad_runs, column_math.cuh) needs no per-run bounds: ~33 instructions per step, ~10 steps;
//   * 25 live samples: 48-56 registers, four 128-thread CTAs per SM alternate their load and sort phases;
//   * quantiles (<= EOC_SPLIT_MAX_Q probabilities): the same elimination as a cursor that moves on from rank to rank.
//
// Arithmetic: cmpx / median_of_pair of column_math.cuh, pinned against numpy by tests/hostmodel
// (test_quad_selection_bit_exact).  Algorithmic bytes per column: R * 4 in, 4 out.
#pragma once

#include <type_traits>

#include "select_tile_kernel.cuh"

namespace eoc {

// Device spelling of quad_valid_prefix / select_quad_runs (column_math.cuh, the host model) on 32-bit shared-memory
// byte ADDRESSES: a step of the elimination is 4 x (add, min against the sentinel's address, ld.shared), one
// FMNMX + FMNMX3, three bit-pattern compares chained through their predicates (the winner is the run whose
// candidate IS the minimum bit for bit: FMNMX orders -0.0 < +0.0 and returns one of its operands) and four
// predicated adds -- 30 instructions where the compiler's rendering of the generic code took 45.
// `col` = address of row 0 of the thread's column, RS = bytes between rows (W * 4).
template <int NH, int RS> __device__ __forceinline__ int quad_valid_prefix_dev(uint32_t col, int r) {
  const uint32_t base = col + (uint32_t)(r * (NH + 1)) * RS;
  const uint32_t end = base + (uint32_t)NH * RS;  // the sentinel
  uint32_t p = base;                               // address of the first element not known to be valid
#pragma unroll
  for (int s = (NH > 64 ? 64 : (NH > 32 ? 32 : (NH > 16 ? 16 : 8))); s >= 1; s >>= 1) {
    asm("{\n\t.reg .pred q;\n\t.reg .u32 c;\n\t.reg .f32 x;\n\t"
        "add.u32 c, %0, %1;\n\t"
        "min.u32 c, c, %2;\n\t"
        "ld.shared.f32 x, [c];\n\t"
        "setp.eq.f32 q, x, x;\n\t"
        "@q add.u32 %0, %0, %3;\n\t}"
        : "+r"(p)
        : "r"((uint32_t)(s - 1) * RS), "r"(end), "r"((uint32_t)s * RS));
  }
  if ((NH & (NH - 1)) == 0) {  // a full run of 2^m samples is one more than the steps can reach
    asm("{\n\t.reg .pred q;\n\t.reg .u32 c;\n\t.reg .f32 x;\n\t"
        "min.u32 c, %0, %1;\n\t"
        "ld.shared.f32 x, [c];\n\t"
        "setp.eq.f32 q, x, x;\n\t"
        "@q add.u32 %0, %0, %2;\n\t}"
        : "+r"(p)
        : "r"(end), "r"((uint32_t)RS));
  }
  return (int)((p - base) / RS);
}

// The elimination as a CURSOR over the four parked runs: `pos` elements of the union have been passed, the heads
// a0..a3 point at the first element of each run that has not.  advance(k) moves on to rank k >= pos, so the
// ranks of ascending probabilities (and x_k, x_{k+1} of one probability) share their steps.
template <int NH, int RS> struct QuadCursor {
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
      asm("{\n\t.reg .pred p0, p1, p2, p01, p012;\n\t.reg .u32 c0, c1, c2, c3;\n\t.reg .f32 v0, v1, v2, v3, m, m2;\n\t"
          "add.u32 c0, %0, %4;\n\t"
          "add.u32 c1, %1, %4;\n\t"
          "add.u32 c2, %2, %4;\n\t"
          "add.u32 c3, %3, %4;\n\t"
          "min.u32 c0, c0, %6;\n\t"
          "min.u32 c1, c1, %7;\n\t"
          "min.u32 c2, c2, %8;\n\t"
          "min.u32 c3, c3, %9;\n\t"
          "ld.shared.f32 v0, [c0];\n\t"
          "ld.shared.f32 v1, [c1];\n\t"
          "ld.shared.f32 v2, [c2];\n\t"
          "ld.shared.f32 v3, [c3];\n\t"
          "min.f32 m2, v2, v3;\n\t"
          "min.f32 m, v0, v1, m2;\n\t"
          "setp.eq.b32 p0, v0, m;\n\t"
          "setp.eq.and.b32 p1, v1, m, !p0;\n\t"
          "or.pred p01, p0, p1;\n\t"
          "setp.eq.and.b32 p2, v2, m, !p01;\n\t"
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
  __device__ __forceinline__ void heads(float& xk, float& xk1) const {
    float m, m1;
    asm("{\n\t.reg .pred p0, p1, p2, p01, p012;\n\t.reg .u32 hn;\n\t.reg .f32 v0, v1, v2, v3, m2, nx;\n\t"
        "ld.shared.f32 v0, [%2];\n\t"
        "ld.shared.f32 v1, [%3];\n\t"
        "ld.shared.f32 v2, [%4];\n\t"
        "ld.shared.f32 v3, [%5];\n\t"
        "min.f32 m2, v2, v3;\n\t"
        "min.f32 %0, v0, v1, m2;\n\t"
        "setp.eq.b32 p0, v0, %0;\n\t"
        "setp.eq.and.b32 p1, v1, %0, !p0;\n\t"
        "or.pred p01, p0, p1;\n\t"
        "setp.eq.and.b32 p2, v2, %0, !p01;\n\t"
        "or.pred p012, p01, p2;\n\t"
        "mov.u32 hn, %5;\n\t"
        "@p0 mov.u32 hn, %2;\n\t"
        "@p1 mov.u32 hn, %3;\n\t"
        "@p2 mov.u32 hn, %4;\n\t"
        "ld.shared.f32 nx, [hn+%6];\n\t"
        "@p0 mov.f32 v0, nx;\n\t"
        "@p1 mov.f32 v1, nx;\n\t"
        "@p2 mov.f32 v2, nx;\n\t"
        "@!p012 mov.f32 v3, nx;\n\t"
        "min.f32 m2, v2, v3;\n\t"
        "min.f32 %1, v0, v1, m2;\n\t}"
        : "=&f"(m), "=&f"(m1)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "n"(RS));
    xk = m;
    xk1 = m1;
  }
};

// FUSED (eoc_select_fused: mean / sd / ... of the same column, BASELINE config 3): streaming reducers run BEFORE the
// sort as rolled loops over the tile rows in shared memory -- sequential in r like numpy, both variance passes out
// of the resident tile.  MASKED (mask -> median / quantiles in one pass): the uint8 mask tile is staged next to the
// data tile by cp.async and a set byte becomes NaN in the register fill on the FMA pipe (byte * 0xFFC00000 +
// 0x80000000 is -0.0f or NaN, then x + m), exactly as in k_select_tile / k_select_split.
// exact float32 value of a raw 8 / 16-bit sample without a conversion instruction: 2^23 + u is exact, minus 2^23
template <typename In> __device__ __forceinline__ float quad_sample_to_float(In x);
template <> __device__ __forceinline__ float quad_sample_to_float<float>(float x) { return x; }
template <> __device__ __forceinline__ float quad_sample_to_float<uint16_t>(uint16_t x) {
  return __fsub_rn(__uint_as_float(0x4B000000u | (uint32_t)x), 8388608.0f);
}
template <> __device__ __forceinline__ float quad_sample_to_float<uint8_t>(uint8_t x) {
  return __fsub_rn(__uint_as_float(0x4B000000u | (uint32_t)x), 8388608.0f);
}
template <> __device__ __forceinline__ float quad_sample_to_float<int16_t>(int16_t x) {
  return __fsub_rn(__uint_as_float(0x4B000000u | ((uint32_t)(uint16_t)x ^ 0x8000u)), 8388608.0f + 32768.0f);
}

// In != float (raw uint16 / int16 / uint8 digital numbers, the dtype Sentinel-2 cubes arrive in): the TMA box of a run
// lands in the UPPER part of the run's own float rows (NH * W * sizeof(In) bytes ending where the sentinel row ends), the
// register fill converts on the way (exact), and a CTA barrier separates "every thread has its run in registers" from the
// park that overwrites the landing zone with float rows; the sentinel is rewritten behind each park.  No NaN exists, so
// the valid count is R.  Results leave as float64 like numpy's (SelectParams::median_f64).
// G > 1 (plain variant only): G threads share a column -- thread (c, g) sorts the runs g, g + G, ... (a warp = 32
// adjacent columns of one g), thread (c, 0) then searches -- as in k_select_multi.  Measured SLOWER at t = 100 (G = 2:
// median 4811 vs 4969 GB/s, three quantiles dealt over the two threads 3514 vs 3740): kept as a knob (EOC_QUAD_G).
template <int NH, bool FULL, int W, int MINB, int FUSED = 0, bool MASKED = false, int G = 1, typename In = float>
__global__ void __launch_bounds__(W * G, MINB)
k_select_quad(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = float;
  constexpr int KR = 4;
  constexpr int S = NH + 1;  // rows per run in shared memory: NH samples + the sentinel
  constexpr bool RAWINT = !std::is_same<In, float>::value;
  constexpr int LAND = RAWINT ? (S * W * (int)sizeof(A) - NH * W * (int)sizeof(In)) / 128 * 128 : 0;  // byte offset of a run's TMA box
  static_assert(LAND % 128 == 0, "the landing zone of a TMA box is 128-byte aligned");
  static_assert(!MASKED || FUSED == 0, "masked variant: no fused reducers");
  static_assert(G == 1 || (FUSED == 0 && !MASKED), "several threads per column: plain variant only");
  static_assert(!RAWINT || (FUSED == 0 && G == 1), "raw integer samples: plain and masked variants only");
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  A* buf = reinterpret_cast<A*>(smem_dyn);  // [KR * S][W]
  uint8_t* mbuf = smem_dyn + (size_t)KR * S * W * sizeof(A);  // MASKED: [R][W] bytes behind the data tile
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int tid = threadIdx.x;
  const int t = G > 1 ? tid % W : tid;  // column of the tile
  const int g = G > 1 ? tid / W : 0;    // this thread sorts the runs g, g + G, ...
  const int R = (int)P.R;
  A* mycol = buf + t;  // row j of this thread's column: mycol[j * W]
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < T.num_tiles ? (T.num_tiles - first + step - 1) / step : 0;

  auto coords = [&](int64_t k, int64_t& o, int64_t& i0) {
    const int64_t tl = first + k * step;
    if constexpr (MASKED) {
      // pixel-tile-major: the `outer` slices (bands) of one pixel tile are neighbours in tile order, so the CTAs
      // running at the same time share the (t, y, x) mask tile in L2 instead of re-reading it once per band
      const int64_t it = tl / P.outer;
      o = tl - it * P.outer;
      i0 = it * W;
    } else {
      o = tl / T.tiles_per_outer;
      i0 = (tl - o * T.tiles_per_outer) * W;
    }
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
  auto issue_mask = [&](int64_t k) {  // every thread: 4-byte chunk (t % (W/4)) of the rows t / (W/4) + 4 m
    int64_t o, i0;
    coords(k, o, i0);
    constexpr int CPR = W / 4;
    const int c4 = (t % CPR) * 4;
    if (i0 + c4 < P.inner) {
      const uint8_t* src = T.mask + o * T.mso + (int64_t)(t / CPR) * T.msr + i0 + c4;
      uint32_t dst = smem_u32(mbuf + (t / CPR) * W + c4);
      for (int r = t / CPR; r < R; r += W / CPR) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
        src += (W / CPR) * T.msr;
        dst += (W / CPR) * W;
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  // sample j of run h as the reducers see it (MASKED: a set mask byte reads as NaN)
  auto sample = [&](int h, int j) -> A {
    if constexpr (MASKED)
      return __fadd_rn(mycol[(h * S + j) * W],
                       __uint_as_float((uint32_t)mbuf[(h * NH + j) * W + t] * T.mf_mul + 0x80000000u));
    else
      return mycol[(h * S + j) * W];
  };

  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // the sentinels: written once, never touched by TMA (its boxes end one row short) nor by the park
  if (g == 0) {
#pragma unroll
    for (int r = 0; r < KR; ++r) mycol[(r * S + NH) * W] = Lim<A>::nan();
  }
  __syncthreads();
  if (tid == 0 && n_my > 0) issue(0);
  if (MASKED && n_my > 0) issue_mask(0);
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
    if constexpr (MASKED) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();  // the mask bytes other threads copied are visible
    }

    // FUSED: pass 1 (sum, count, extrema) is a rolled loop over the resident tile, sequential in r like numpy; pass 2
    // of the variance (sum of squared deviations from the mean, in r order too) rides on the register fill of the
    // sort below -- the samples are in registers there anyway, in their original order -- so the tile is read twice,
    // not three times.  A slot past R holds NaN and adds fmax(NaN^2, 0) = 0.
    ColStats<A, G_MINMAX | G_RARE> st;
    A sumsq = (A)0, avg = (A)0;
    bool var_in_fill = false;
    if constexpr (FUSED != 0) {
      const ColumnParams& C = P.fused;
      st.init();
      A cnt = (A)0;
#pragma unroll 1
      for (int h = 0; h < KR; ++h) {
        const int rows = FULL ? NH : (R - h * NH < NH ? R - h * NH : NH);
#pragma unroll 5
        for (int j = 0; j < rows; ++j) {
          const A x = sample(h, j);
          sum_step_fcount(x, st.sum, cnt);
          if constexpr (FUSED >= 2) {
            st.mn = min_(st.mn, x);
            st.mx = max_(st.mx, x);
            rare_step(x, st.prod, st.nfin);
          }
        }
      }
      st.n = (int)cnt;
      if (C.ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) {
        avg = mean_of(st.sum, st.n);
        var_in_fill = sqdev_fast_ok(avg, st.n);
        if (!var_in_fill) {  // non-finite mean: the exact masking form, rare, rolled
          for (int h = 0; h < KR; ++h)
            for (int j = 0; j < NH && h * NH + j < R; ++j) sumsq = add_rn(sumsq, sqdev(sample(h, j), avg));
        }
      }
    }

#pragma unroll 1
    for (int h = g; h < KR; h += G) {
      A* run = mycol + h * S * W;
      const int left = R - h * NH;  // rows of this run that hold samples (FULL: all NH)
      A v[NH];
#pragma unroll
      for (int j = 0; j < NH; ++j) {
        if constexpr (MASKED) {
          const uint8_t* mrun = mbuf + h * NH * W + t;
          if (FULL || j < left) {
            A xv;
            if constexpr (RAWINT) {  // raw digital numbers under a cloud mask: convert, then the mask's NaN
              const In* land = reinterpret_cast<const In*>(reinterpret_cast<const unsigned char*>(buf + h * S * W) + LAND) + t;
              xv = quad_sample_to_float<In>(land[j * W]);
            } else {
              xv = run[j * W];
            }
            v[j] = __fadd_rn(xv, __uint_as_float((uint32_t)mrun[j * W] * T.mf_mul + 0x80000000u));
          } else {
            v[j] = Lim<A>::nan();
          }
        } else if constexpr (RAWINT) {
          const In* land = reinterpret_cast<const In*>(reinterpret_cast<const unsigned char*>(buf + h * S * W) + LAND) + t;
          v[j] = (FULL || j < left) ? quad_sample_to_float<In>(land[j * W]) : Lim<A>::nan();
        } else {
          v[j] = (FULL || j < left) ? run[j * W] : Lim<A>::nan();
        }
      }
      if constexpr (RAWINT) __syncthreads();  // every thread holds run h: the park may overwrite the landing zone
      if constexpr (FUSED != 0) {
        if (var_in_fill) {
#pragma unroll
          for (int j = 0; j < NH; ++j) sumsq = add_rn(sumsq, sqdev_fast(v[j], avg));
        }
      }
      sortraw<NH, A>(v, one);
#pragma unroll
      for (int j = 0; j < NH; ++j) run[j * W] = v[j];
      if constexpr (RAWINT) run[NH * W] = Lim<A>::nan();  // the landing zone covered the sentinel
    }
    if constexpr (FUSED != 0) {
      if (active) finalize_column<A, G_MINMAX | G_RARE>(st, R, sumsq, P.fused, o * P.inner + i, false);
    }

    // each thread reads back only its own column: no barrier needed before the fetches -- unless other threads
    // sorted the other runs of it
    if constexpr (G > 1) __syncthreads();
    const uint32_t col = smem_u32(mycol);
    // G > 1: thread (c, 0) takes the median, the probabilities are dealt out in contiguous blocks over the column's
    // threads (ascending probabilities of one thread share their elimination steps); every thread with work finds
    // the valid count itself
    const int qb = (P.nq + G - 1) / G;
    const int q_lo = g * qb, q_hi = (q_lo + qb < P.nq) ? q_lo + qb : P.nq;
    const bool has_work = (g == 0 && P.median_out != nullptr) || q_lo < q_hi;
    int n = 0;
    if constexpr (RAWINT && !MASKED) {
      n = R;  // integers hold no NaN
    } else if (has_work) {
#pragma unroll
      for (int r = 0; r < KR; ++r) n += quad_valid_prefix_dev<NH, W * 4>(col, r);
    }
    if (active && has_work) {
      const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
      QuadCursor<NH, W * 4> cur;
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
      for (int q = q_lo; q < q_hi; ++q) {
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
    // the park wrote the buffer through the generic proxy; order those writes before the TMA (async proxy)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();  // all rank fetches done: the buffer may be overwritten by the next tile
    if (tid == 0 && k + 1 < n_my) issue(k + 1);
    if (MASKED && k + 1 < n_my) issue_mask(k + 1);
  }
}

}  // namespace eoc
