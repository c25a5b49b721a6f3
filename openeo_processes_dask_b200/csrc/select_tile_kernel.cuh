// Kernel C-T: the register selection kernel fed by TMA instead of per-thread global loads.
//
// ncu on k_select_regs (profiles/r1b_quantiles_raw.csv) showed the ALU pipe -- the pipe the sorting
// network saturates -- only 66-70 % busy: a CTA that has just started cannot even issue its 100 loads
// (address arithmetic, predicates) while the other CTAs of the SM are sorting, so its load latency is
// exposed afterwards.  Here a persistent CTA of 128 threads walks tiles of 128 columns; ONE thread
// issues a single cp.async.bulk.tensor for the whole (R x 128) tile and everybody sleeps on the
// mbarrier, then every thread pulls its column into registers with conflict-free LDS (immediate
// offsets, no address arithmetic) and the sort starts.  For the median the next tile's TMA is issued
// right after the registers are filled, so it lands while this tile is being sorted; the quantile
// variant parks the sorted column in the same buffer first (rank fetches are data dependent) and
// issues the next TMA after that, with an L2 prefetch of the next tile up front.
// Arithmetic, networks and outputs are those of k_select_regs (select_kernel.cuh).
#pragma once

#include "select_kernel.cuh"
#include "tile_kernel.cuh"

namespace eoc {

// columns per tile == threads per CTA.  256: the 8 warps of a CTA walk the straight-line network in lockstep
// and share its instruction fetch; measured on B200 (T=100): quantiles 2601 / 2684 / 2473 GB/s and median
// 3352 / 3556 / 3466 GB/s for 128 / 256 / 512 columns.
constexpr int SELT_W = 256;

struct SelectTileParams {
  SelectParams S;
  int64_t tiles_per_outer, num_tiles;
  uint32_t tile_bytes;  // R * SELT_W * sizeof(float)
  const float* data;    // the view's base pointer (rare exact-variance reload of the fused variants)
  uint32_t stagger_ns;  // start delay per resident-CTA slot (de-phases the CTAs of one SM)
  int sms;
  // MASKED variant: canonical uint8 mask (1 = sample reads as NaN) with its own strides (0 = broadcast)
  const uint8_t* mask;
  int64_t mso, msr;
  uint32_t mf_mul;  // 0xFFC00000 at run time: byte * mf_mul + 0x80000000 is -0.0f (keep) or NaN (replace)
  int box_r = 0, nbox_r = 1;  // k_select_split with more than two runs: rows per TMA box and boxes per column box
};

__device__ __forceinline__ void tma_prefetch_box(const CUtensorMap* map, int32_t c0, int32_t c1, int32_t c2) {
  asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global [%0, {%1, %2, %3}];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

// MASKED (mask -> median / quantiles in one pass): the uint8 mask tile (R x W bytes) is
// staged next to the data tile by cp.async (4-byte chunks issued by all threads; a uint8 row stride is rarely
// the multiple of 16 bytes TMA wants), and the register fill turns a set byte into NaN WITHOUT touching the
// ALU pipe the network saturates: byte * 0xFFC00000 + 0x80000000 is the bit pattern of -0.0f (byte 0) or of
// a quiet NaN (byte 1) -- one IMAD -- and x + (-0.0f) == x bit for bit for every x, x + NaN == NaN -- one FADD.
template <int N, bool QUANT, bool FULL, int FUSED, bool LOCK = true, int W = SELT_W, bool MASKED = false>
__global__ void __launch_bounds__(W, MASKED ? 3 : (select_blocks<float, N, QUANT, FUSED>() * 128 / W < 1 ? 1 : select_blocks<float, N, QUANT, FUSED>() * 128 / W))
k_select_tile(const __grid_constant__ CUtensorMap tmap, const SelectTileParams T) {
  using A = float;
  static_assert(!MASKED || (W <= 256 && FUSED == 0), "masked variant: one box per tile, no fused reducers");
  extern __shared__ __align__(1024) unsigned char smem_dyn[];
  // A tile is W columns wide, staged as W / BW boxes of BW <= 256 columns (the TMA box limit); box b lives at
  // buf + b * rows * BW with `rows` = R (median) or N (quantiles: the sorted column is parked in place).
  constexpr int BW = W < 256 ? W : 256;
  constexpr int NBOX = W / BW;
  A* buf = reinterpret_cast<A*>(smem_dyn);
  __shared__ uint64_t full;
  const SelectParams& P = T.S;
  const int t = threadIdx.x;
  const int R = (int)P.R;
  const int rows = QUANT ? N : R;
  A* mycol = buf + (t / BW) * rows * BW + (t % BW);  // row j of this thread's column: mycol[j * BW]
  uint8_t* mbuf = smem_dyn + (size_t)rows * W * sizeof(A);  // MASKED: [R][W] bytes behind the data tile
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
    for (int b = 0; b < NBOX; ++b)
      tma_load_box(buf + b * rows * BW, &tmap, &full, (int32_t)(i0 + b * BW), 0, (int32_t)o);
  };

  // MASKED: every thread copies the 4-byte chunk (t % (W/4)) of the rows t / (W/4) + k * 4 of tile k's mask
  auto issue_mask = [&](int64_t k) {
    int64_t o, i0;
    coords(k, o, i0);
    constexpr int CPR = W / 4;  // chunks per row
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

  if (t == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    mbar_init(&full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (T.stagger_ns) __nanosleep(T.stagger_ns * (uint32_t)(blockIdx.x / T.sms));
  if (t == 0 && n_my > 0) issue(0);
  if (MASKED && n_my > 0) issue_mask(0);

  for (int64_t k = 0; k < n_my; ++k) {
    int64_t o, i0;
    coords(k, o, i0);
    const int64_t i = i0 + t;
    const bool active = i < P.inner;
    if (QUANT && t == 0 && k + 1 < n_my) {  // warm L2 with the next tile; its TMA is issued after the park
      int64_t o2, i2;
      coords(k + 1, o2, i2);
#pragma unroll
      for (int b = 0; b < NBOX; ++b) tma_prefetch_box(&tmap, (int32_t)(i2 + b * BW), 0, (int32_t)o2);
    }
    mbar_wait(&full, (uint32_t)(k & 1));
    if constexpr (MASKED) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();  // the mask bytes other threads copied are visible
    }

    A v[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
      if constexpr (MASKED) {
        if (FULL || j < R) {
          const uint32_t mb = mbuf[j * W + t];
          v[j] = __fadd_rn(mycol[j * BW], __uint_as_float(mb * T.mf_mul + 0x80000000u));
        } else {
          v[j] = Lim<A>::nan();
        }
      } else if constexpr (FULL) {
        v[j] = mycol[j * BW];
      } else {
        v[j] = (j < R) ? mycol[j * BW] : Lim<A>::nan();
      }
    }
    __syncthreads();  // every thread holds its column: the buffers are free
    if (!QUANT && t == 0 && k + 1 < n_my) issue(k + 1);
    if (MASKED && k + 1 < n_my) issue_mask(k + 1);

    if constexpr (FUSED != 0) {
      const float* gcol = T.data + o * P.so + (active ? i : 0);
      const int64_t sr = P.sr;
      fused_column_stats<N, A, FUSED>(v, P.fused, o * P.inner + i, active, [&](int j) { return __ldg(gcol + j * sr); });
    }

    if constexpr (!QUANT) {
      const A m = nanmedian_network<N, A>(v, R, P.ignore_nodata != 0, (A)P.zero);
      if (active) ((A*)P.median_out)[o * P.inner + i] = m;
    } else {
      sort_nan_last<N, A, LOCK>(v, (A)P.zero);  // NaN (nodata, unused slots) last, no canonicalisation pass
#pragma unroll
      for (int j = 0; j < N; ++j) mycol[j * BW] = v[j];
      // each thread reads back only its own column: no barrier needed before the fetches
      const int n = first_nan_index(N, [&](int r) { return mycol[r * BW]; });
      // (explicit conversions: passing the constexpr local BW as it is trips an internal error of the nvcc 12.9 front end)
      if (active) emit_sorted<A>(P, o, i, n, (const A*)mycol, int(BW));
      // the park wrote the buffer through the generic proxy; order those writes before the TMA (async
      // proxy) that refills it
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();  // all rank fetches done: the buffer may be overwritten by the next tile
      if (t == 0 && k + 1 < n_my) issue(k + 1);
      // (Measured: fetching the ranks first and interpolating / storing only after the issue -- to hide the
      // TMA latency behind the float64 work -- was 5 % SLOWER: the values kept live across the barrier pushed
      // the register allocation of the network into spills.)
    }
  }
}

}  // namespace eoc
