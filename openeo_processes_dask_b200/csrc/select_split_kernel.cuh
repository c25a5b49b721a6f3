// Kernel C-S (round 2): TMA-staged selection with the column sorted as TWO runs.
//
// ncu on k_select_tile (profiles/r1i_median_raw.csv, r1i_quantiles_raw.csv) showed both variants bound by the
// number of instructions per column -- 3016 for the median of 100 samples: 898 comparators of the pruned
// Batcher cone, 400 instructions of NaN pad substitution, and a straight-line body of 40-70 KB that made the
// warps wait on instruction fetch unless the whole CTA ran in lockstep.  This kernel changes the algorithm:
//
//   * the R <= 2*NH samples of a column are sorted as two runs of NH samples, each with the NaN-native
//     comparator (NaN = nodata / unused slot sorts last), by ONE network body executed twice in a rolled loop:
//     2 * S(NH) comparators (806 for NH = 50) instead of S(2 NH) = 1104 (quantiles), NO pad substitution, and
//     ~17 KB of code instead of ~50 KB (no lockstep barriers needed: stall_no_instruction 0.08-0.19 per issue);
//   * each run is parked back into the rows of the TMA tile it came from (thread t owns column t of the tile:
//     conflict-free), so only NH samples are ever live in registers (72-95 registers instead of 128+);
//   * the order statistics numpy asks for -- the middle pair, or the neighbours of (n-1)q -- are ranks of the
//     UNION of the two runs: select_two_runs() (column_math.cuh) finds them with a 6-step partition search over
//     the parked runs, the same code for the median and for quantiles;
//   * fused streaming reducers (eoc_select_fused: mean / sd / ... of the same column, BASELINE config 3) run
//     BEFORE the sort as rolled loops over the tile rows in shared memory -- sequential in r like numpy, both
//     variance passes out of the resident tile, no unrolled 100-step blocks and no reload from global memory;
//   * MASKED (mask -> median in one pass): the uint8 mask tile is staged next to the data tile and a set byte
//     becomes NaN in the register fill on the FMA pipe (byte * 0xFFC00000 + 0x80000000 is -0.0f or NaN, then
//     x + m), exactly as in k_select_tile.
//
// A CTA is persistent (grid = SMs x resident CTAs), W threads walk tiles of W adjacent columns x R rows brought
// in by one cp.async.bulk.tensor.3d per <= 256-column box; several CTAs per SM alternate their load and sort
// phases.  Arithmetic: cmpx / quantile_lerp / ColStats of column_math.cuh, i.e. what tests/hostmodel pins
// against numpy bit for bit (test_split_selection_bit_exact).
// Algorithmic bytes per column: R * 4 in, 4 (median) + nq * 8 (quantiles) + the fused outputs out.
#pragma once

#include "select_tile_kernel.cuh"

namespace eoc {

#ifndef EOC_SPLIT_MAX_Q
#define EOC_SPLIT_MAX_Q 4  // one rank search per probability: past this the full sort is cheaper
#endif

// KR = number of runs: 2 for columns of up to 128 samples; 3 or 4 runs of 100 / 128 samples cover 129 .. 512
// samples (a year of daily scenes) with the same sort body and the K-run rank elimination of column_math.cuh
// (select_k_runs) in place of the two-run partition search.  Rows beyond 256 arrive as a second TMA box.
// PAR: the KR runs of a column are sorted by KR THREADS at the same time (thread t: column t % W, run t / W, so a
// warp works on one run of 32 adjacent columns), then one thread per column does the rank search.  Same shared
// memory per column, KR times the resident warps: long columns leave room for only ~4 warps per SM otherwise
// (ncu r2 median365: issue slots 35 % busy, 1.3 no-instruction stalls per issue with one warp per scheduler).
template <int NH, bool FULL, int FUSED, bool MASKED, int W, int MINB, bool LOCK, int KR = 2, bool PAR = false>
__global__ void __launch_bounds__(PAR ? W * KR : W, MINB)
k_select_split(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = float;
  static_assert(KR >= 2 && KR <= 4, "2 .. 4 runs");
  static_assert(KR == 2 || !MASKED, "the masked variant covers two runs");
  static_assert(!PAR || (!MASKED && FUSED == 0), "parallel runs: plain selection only");
  constexpr int N = KR * NH;
  constexpr int IT = split_iters(NH);
  constexpr int BW = W < 256 ? W : 256;
  constexpr int NBOX = W / BW;
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  A* buf = reinterpret_cast<A*>(smem_dyn);  // box b at buf + b * N * BW: N rows, the first R filled by TMA
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int tid = threadIdx.x;
  const int t = PAR ? tid % W : tid;   // column of the tile this thread works on
  const int hp = PAR ? tid / W : 0;    // PAR: the run this thread sorts
  const int R = (int)P.R;
  A* mycol = buf + (t / BW) * N * BW + (t % BW);  // row j of this thread's column: mycol[j * BW]
  uint8_t* mbuf = smem_dyn + (size_t)N * W * sizeof(A);  // MASKED: [R][W] bytes behind the data tile
  const int64_t first = blockIdx.x, step = gridDim.x;
  const int64_t n_my = first < T.num_tiles ? (T.num_tiles - first + step - 1) / step : 0;

  auto coords = [&](int64_t k, int64_t& o, int64_t& i0) {
    const int64_t tl = first + k * step;
    if constexpr (MASKED) {
      // pixel-tile-major: the `outer` slices (bands) of one pixel tile are neighbours in tile order, so the CTAs
      // running at the same time share the (t, y, x) mask tile in L2 instead of re-reading it once per band
      // (ncu r1/r2: 11.25 GB read for 9.55 GB algorithmic with the band-major walk)
      const int64_t it = tl / P.outer;
      o = tl - it * P.outer;
      i0 = it * W;
    } else {
      o = tl / T.tiles_per_outer;
      i0 = (tl - o * T.tiles_per_outer) * W;
    }
  };
  auto issue = [&](int64_t k) {  // thread 0 only
    int64_t o, i0;
    coords(k, o, i0);
    mbar_expect_tx(&full, T.tile_bytes);
#pragma unroll
    for (int b = 0; b < NBOX; ++b) {
      if constexpr (KR == 2) {
        tma_load_box(buf + b * N * BW, &tmap, &full, (int32_t)(i0 + b * BW), 0, (int32_t)o);
      } else {
        for (int rb = 0; rb < T.nbox_r; ++rb)  // boxes of box_r <= 256 rows; the last may reach past R (zero fill)
          tma_load_box(buf + b * N * BW + rb * T.box_r * BW, &tmap, &full, (int32_t)(i0 + b * BW), rb * T.box_r,
                       (int32_t)o);
      }
    }
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
  // sample r of this thread's column as the reducers see it (MASKED: a set mask byte reads as NaN)
  auto sample = [&](int r) -> A {
    if constexpr (MASKED)
      return __fadd_rn(mycol[r * BW], __uint_as_float((uint32_t)mbuf[r * W + t] * T.mf_mul + 0x80000000u));
    else
      return mycol[r * BW];
  };

  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (T.stagger_ns) __nanosleep(T.stagger_ns * (uint32_t)(blockIdx.x / T.sms));
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
      for (int b = 0; b < NBOX; ++b)
        for (int rb = 0; rb < (KR == 2 ? 1 : T.nbox_r); ++rb)
          tma_prefetch_box(&tmap, (int32_t)(i2 + b * BW), rb * T.box_r, (int32_t)o2);
    }
    mbar_wait(&full, (uint32_t)(k & 1));
    if constexpr (MASKED) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();  // the mask bytes other threads copied are visible
    }

    if constexpr (FUSED != 0) {
      // streaming reducers of the same column, sequential in r out of the resident tile (numpy's order)
      const ColumnParams& C = P.fused;
      ColStats<A, G_MINMAX | G_RARE> st;
      st.init();
      A cnt = (A)0;
#pragma unroll 4
      for (int r = 0; r < R; ++r) {
        const A x = sample(r);
        sum_step_fcount(x, st.sum, cnt);
        if constexpr (FUSED >= 2) {
          st.mn = min_(st.mn, x);
          st.mx = max_(st.mx, x);
          rare_step(x, st.prod, st.nfin);
        }
      }
      st.n = (int)cnt;
      A sumsq = (A)0;
      if (C.ops & ((1u << EOC_OP_VARIANCE) | (1u << EOC_OP_SD))) {
        const A avg = mean_of(st.sum, st.n);
        if (sqdev_fast_ok(avg, st.n)) {
#pragma unroll 4
          for (int r = 0; r < R; ++r) sumsq = add_rn(sumsq, sqdev_fast(sample(r), avg));
        } else {
          for (int r = 0; r < R; ++r) sumsq = add_rn(sumsq, sqdev(sample(r), avg));
        }
      }
      if (active) finalize_column<A, G_MINMAX | G_RARE>(st, R, sumsq, C, o * P.inner + i, false);
    }

    // the runs: rows [h NH, (h + 1) NH) of the column, sorted in registers by the same code, parked in place
#pragma unroll 1
    for (int h = (PAR ? hp : 0); h < (PAR ? hp + 1 : KR); ++h) {
      A* run = mycol + h * NH * BW;
      const int left = R - h * NH;  // rows of this run that hold samples (FULL: all NH)
      A v[NH];
#pragma unroll
      for (int j = 0; j < NH; ++j) {
        if constexpr (MASKED) {
          const uint8_t* mrun = mbuf + h * NH * W + t;
          // (an unused slot must hold the CANONICAL NaN too -- the hardware's, not the compiler's literal: the
          // mixed-pipe comparators rebuild `hi` from bit patterns)
          if (FULL || j < left)
            v[j] = __fadd_rn(run[j * BW], __uint_as_float((uint32_t)mrun[j * W] * T.mf_mul + 0x80000000u));
          else
            v[j] = __fsub_rn(Lim<A>::nan(), P.zero);
        } else {
          // x - 0.0 (run-time zero): the same number bit for bit, and any NaN becomes the canonical one the
          // mixed-pipe comparators need; one FADD on the FMA pipe
          v[j] = __fsub_rn((FULL || j < left) ? run[j * BW] : Lim<A>::nan(), P.zero);
        }
      }
      sortnet<NH, A, LOCK, true>(v, one);
#pragma unroll
      for (int j = 0; j < NH; ++j) run[j * BW] = v[j];
    }

    // each thread reads back only its own column: no barrier needed before the fetches -- unless other threads
    // sorted the other runs of it
    if constexpr (PAR) __syncthreads();
    if constexpr (KR == 2 && !PAR) {
      ParkedColumn<A> fa, fb;
      fa.col = mycol;
      fa.stride = BW;
      fb.col = mycol + NH * BW;
      fb.stride = BW;
      const int na = first_nan_index(NH, fa);
      const int nb = first_nan_index(NH, fb);
      const int n = na + nb;
      if (active) {
        const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
        if (P.median_out) {
          A m = Lim<A>::nan();
          if (!poisoned) m = median_two_runs<IT, A>(fa, na, fb, nb);
          store_median<A>(P, o * P.inner + i, m);
        }
        for (int q = 0; q < P.nq; ++q) {
          double r = Lim<double>::nan();
          if (!poisoned) r = quantile_two_runs<IT, A>(fa, na, fb, nb, P.probs[q], P.lerp_wrap16 != 0);
          store_quant<A>(P, o, q, i, r);
        }
      }
    } else {
      const A* col = mycol;
      auto f = [col](int r, int idx) { return col[(r * NH + idx) * BW]; };
      int nr[KR], n = 0;
#pragma unroll
      for (int r = 0; r < KR; ++r) {
        nr[r] = first_nan_index(NH, [&](int idx) { return f(r, idx); });
        n += nr[r];
      }
      if (active) {
        const bool poisoned = (!P.ignore_nodata && n < R) || n == 0;
        if (P.median_out && hp == 0) {
          A m = Lim<A>::nan();
          if (!poisoned) m = median_k_runs<KR, A>(f, nr, n);
          store_median<A>(P, o * P.inner + i, m);
        }
        for (int q = 0; q < P.nq; ++q) {
          if (PAR && (q + 1) % KR != hp) continue;  // the threads of a column share its probabilities
          double r = Lim<double>::nan();
          if (!poisoned) r = quantile_k_runs<KR, A>(f, nr, n, P.probs[q], P.lerp_wrap16 != 0);
          store_quant<A>(P, o, q, i, r);
        }
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
