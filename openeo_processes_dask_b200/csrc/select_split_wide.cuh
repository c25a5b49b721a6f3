// Long columns (128 < R <= 512, float32) on the split selection kernel: K = 2 .. 4 sorted runs of 100 / 128 samples
// parked in the TMA tile, ranks by K-run elimination.  Replaces k_select_wide (K threads per column, ~150 KB of
// straight-line SASS, 0.11 - 0.22 of the HBM peak: ncu r1i showed it starving on instruction fetch) wherever TMA can
// address the view.  One translation unit per tile width (select_split_wide64.cu / select_split_wide128.cu).
#pragma once

#include "select_split_launch.cuh"

namespace eoc {

template <int NH, int KR, int W>
int launch_select_split_kw(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  constexpr bool PAR = true;              // one thread per (column, run)
  constexpr int MINB = W >= 128 ? 1 : 2;  // CTAs of W * KR threads; shared memory decides the residency
  const int nbox_r = (int)((in.R + 255) / 256);
  const int box_r = (int)((in.R + nbox_r - 1) / nbox_r);
  CUtensorMap map;
  if (int rc = tma_make_map(in, W, box_r, &map)) return rc;
  SelectTileParams T;
  T.S = P;
  T.tiles_per_outer = (in.inner + W - 1) / W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  T.tile_bytes = (uint32_t)((size_t)nbox_r * box_r * W * sizeof(float));
  T.data = (const float*)in.data;
  T.sms = sm_count();
  T.stagger_ns = 0;
  T.mask = nullptr;
  T.mso = T.msr = 0;
  T.mf_mul = 0xFFC00000u;
  T.box_r = box_r;
  T.nbox_r = nbox_r;
  const size_t smem = (size_t)KR * NH * W * sizeof(float);
  if ((size_t)nbox_r * box_r > (size_t)KR * NH) EOC_FAIL(EOC_ERR_UNSUPPORTED, "split selection: boxes exceed the tile");
  int per_sm = 0;
  auto kern = k_select_split<NH, false, 0, false, W, MINB, false, KR, PAR>;
  EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  constexpr int NT = PAR ? W * KR : W;
  EOC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t grid = (int64_t)sm_count() * per_sm;
  if (grid > T.num_tiles) grid = T.num_tiles;
  kern<<<(unsigned)grid, NT, smem, s>>>(map, T);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// Runs for a column of R samples: the cheapest (NH, K) with K * NH >= R.  The parked runs cost R * 4 bytes of shared
// memory per column, so residency falls with R; with one thread per (column, run) the resident warps are K times
// what one thread per column gives.  Measured on B200 (input GB/s, 30 % NaN, median; one thread per column ->
// one thread per run; k_select_wide of round 1 in brackets): t = 200: 2354 -> 2939 (1465); t = 300: 1174 -> 2496;
// t = 365: 1118 -> 2259 (846).
template <int W> int launch_select_split_wide_w(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  if (in.R <= 200) return launch_select_split_kw<100, 2, W>(in, P, s);
  if (in.R <= 256) return launch_select_split_kw<128, 2, W>(in, P, s);
  if (in.R <= 300) return launch_select_split_kw<100, 3, W>(in, P, s);
  if (in.R <= 384) return launch_select_split_kw<128, 3, W>(in, P, s);
  if constexpr (W <= 64) {
    if (in.R > 400) return launch_select_split_kw<128, 4, W>(in, P, s);  // 256 KB per 128-column tile: 64 only
  }
  return launch_select_split_kw<100, 4, W>(in, P, s);
}

int launch_select_split_wide64(const eoc_view& in, const SelectParams& P, cudaStream_t s);
int launch_select_split_wide128(const eoc_view& in, const SelectParams& P, cudaStream_t s);

#ifndef EOC_SPLIT_WIDE_W_DEFAULT
#define EOC_SPLIT_WIDE_W_DEFAULT 64
#endif
inline int launch_select_split_wide(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  // 128-column tiles wherever the tile fits shared memory (K * NH * 512 bytes <= 227 KB), else 64-column tiles
  const int w = split_env("EOC_SPLIT_WIDE_W", 128);
  if (w == 128 && in.R <= 400) return launch_select_split_wide128(in, P, s);
  return launch_select_split_wide64(in, P, s);
}

// select_multi.cu: eight short runs with sentinels, several threads per column (select_multi_kernel.cuh); takes what
// select_split_wide_applies() accepts unless EOC_SELECT_MULTI=0
int launch_select_multi(const eoc_view& in, const SelectParams& P, cudaStream_t s);
int launch_select_multi_int(const eoc_view& in, const SelectParams& P, cudaStream_t s);
inline bool select_multi_enabled() { return split_env("EOC_SELECT_MULTI", 1) == 1; }

// float32, 128 < R <= 512, TMA-addressable, enough columns, no sorted output, few probabilities
inline bool select_split_wide_applies(const eoc_view& in, int nq, bool want_sorted) {
  if (in.dtype != EOC_F32 || in.R <= 128 || in.R > 512) return false;
  if (want_sorted || nq > EOC_SPLIT_MAX_Q) return false;
  if (split_env("EOC_SELECT_SPLIT", -1) == 0 || !tma_view_ok(in)) return false;
  const char* e = getenv("EOC_SELECT_TMA");
  if (e && e[0] == '0') return false;
  if (e && e[0] == '1') return true;
  return in.outer * in.inner >= 65536 && tma_tiles_fill(in, 64);
}

}  // namespace eoc
