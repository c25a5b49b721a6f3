// Order statistics of 8/16-bit integer cubes (raw Sentinel-2 / Landsat digital numbers) on the float32 register
// networks: every such sample is exact in float32, so the order, the two middle samples, their mean
// ((lo + hi) / 2 < 2^17: exact) and the float64 interpolation of the quantiles are numpy's; the results are stored
// as float64, the dtype np.nanmedian / np.nanquantile return for integer input.  Before this the integer cubes
// took the shared-memory sorter: 32 GB/s for a uint16 median at t = 100.  One translation unit per sample type
// (EOC_INT_T) so they compile in parallel; only the predicated-load (FULL = false) variants are instantiated.
#pragma once

#include "select_wide_kernel.cuh"

namespace eoc {

template <typename In, int N, bool QUANT>
int launch_regs_int(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int threads = select_tpb<float, QUANT>();
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const size_t smem = QUANT ? (size_t)N * threads * sizeof(float) : 0;
  auto kern = k_select_regs<In, float, N, QUANT, false, 0>;
  if (smem > 48 * 1024) EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, (uint32_t)(in.stride_R * (int64_t)sizeof(In)));
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In, bool QUANT> int launch_int_n(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int64_t R = in.R;
  if (R <= 16) return launch_regs_int<In, 16, QUANT>(in, P, s);
  if (R <= 32) return launch_regs_int<In, 32, QUANT>(in, P, s);
  if (R <= 64) return launch_regs_int<In, 64, QUANT>(in, P, s);
  if (R <= 100) return launch_regs_int<In, 100, QUANT>(in, P, s);
  if (R <= 128) return launch_regs_int<In, 128, QUANT>(in, P, s);
  return launch_wide_t<In>(in, P, s);
}

template <typename In> int launch_select_int_t(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s) {
  return quant ? launch_int_n<In, true>(in, P, s) : launch_int_n<In, false>(in, P, s);
}

}  // namespace eoc
