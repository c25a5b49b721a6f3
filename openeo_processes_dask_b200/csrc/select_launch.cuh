// Launch templates of the register selection kernel, shared by select.cu and select_fused.cu (the fused
// float32 variants are a translation unit of their own so the two compile in parallel).
#pragma once

#include "select_kernel.cuh"

namespace eoc {

template <typename In, typename A, int N, bool QUANT, int FUSED>
int launch_regs(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int threads = (QUANT && sizeof(A) == 8) ? 64 : select_tpb<A, QUANT>();
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const size_t smem = QUANT ? (size_t)N * threads * sizeof(A) : 0;
  const uint32_t sr_bytes = (uint32_t)(in.stride_R * (int64_t)sizeof(In));
  bool launched = false;
  if constexpr (FUSED < 2) {  // level-2 fused variants always take the predicated loads (fewer kernels)
    if (in.R == N) {
      auto kern = k_select_regs<In, A, N, QUANT, true, FUSED>;
      if (smem > 48 * 1024)
        EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, sr_bytes);
      launched = true;
    }
  }
  if (!launched) {
    auto kern = k_select_regs<In, A, N, QUANT, false, FUSED>;
    if (smem > 48 * 1024)
      EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, sr_bytes);
  }
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In, typename A, bool QUANT, int FUSED>
int launch_regs_n(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int64_t R = in.R;
  if (R <= 8) return launch_regs<In, A, 8, QUANT, FUSED>(in, P, s);
  if (R <= 16) return launch_regs<In, A, 16, QUANT, FUSED>(in, P, s);
  if (R <= 32) return launch_regs<In, A, 32, QUANT, FUSED>(in, P, s);
  if (R <= 64) return launch_regs<In, A, 64, QUANT, FUSED>(in, P, s);
  if constexpr (sizeof(A) == 4) {
    if (R <= 100) return launch_regs<In, A, 100, QUANT, FUSED>(in, P, s);
    if (R <= 128) return launch_regs<In, A, 128, QUANT, FUSED>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_UNSUPPORTED, "register selection kernel: R=%lld too large", (long long)R);
}

// select_wide.cu
bool select_wide_applies(const eoc_view& in);
int launch_select_wide(const eoc_view& in, const SelectParams& P, cudaStream_t s);
bool select_wide64_applies(const eoc_view& in);
int launch_select_wide64(const eoc_view& in, const SelectParams& P, cudaStream_t s);

// select_int_{u8,i16,u16}.cu: 8/16-bit integer cubes on the float32 networks, float64 results (P.median_f64 = 1)
int launch_select_int_u8(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s);
int launch_select_int_i16(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s);
int launch_select_int_u16(const eoc_view& in, const SelectParams& P, bool quant, cudaStream_t s);

// select_fused.cu
int launch_select_fused_f32(const eoc_view& in, const SelectParams& P, bool quant, int level, cudaStream_t s);

}  // namespace eoc
