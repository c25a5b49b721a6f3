// Launcher of eoc_select: picks the register-network kernel for float32/float64 series that fit
// in registers, the shared-memory sorter otherwise.
#include <cstdlib>
#include <type_traits>

#include "select_kernel.cuh"

namespace eoc {
namespace {

template <typename In, typename A, int N, bool QUANT>
int launch_regs(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int threads = (QUANT && sizeof(A) == 8) ? 64 : select_tpb<A, QUANT>();
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  const size_t smem = QUANT ? (size_t)N * threads * sizeof(A) : 0;
  const uint32_t sr_bytes = (uint32_t)(in.stride_R * (int64_t)sizeof(In));
  if (in.R == N) {
    auto kern = k_select_regs<In, A, N, QUANT, true>;
    if (smem > 48 * 1024)
      EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, sr_bytes);
  } else {
    auto kern = k_select_regs<In, A, N, QUANT, false>;
    if (smem > 48 * 1024)
      EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, sr_bytes);
  }
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In, typename A, bool QUANT>
int launch_regs_n(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  const int64_t R = in.R;
  if (R <= 8) return launch_regs<In, A, 8, QUANT>(in, P, s);
  if (R <= 16) return launch_regs<In, A, 16, QUANT>(in, P, s);
  if (R <= 32) return launch_regs<In, A, 32, QUANT>(in, P, s);
  if (R <= 64) return launch_regs<In, A, 64, QUANT>(in, P, s);
  if constexpr (sizeof(A) == 4) {
    if (R <= 100) return launch_regs<In, A, 100, QUANT>(in, P, s);
    if (R <= 128) return launch_regs<In, A, 128, QUANT>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_UNSUPPORTED, "register selection kernel: R=%lld too large", (long long)R);
}

template <typename In>
int launch_smem(const eoc_view& in, const SelectParams& P, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  int npad = 2;
  while (npad < in.R) npad <<= 1;
  int threads = 128;
  while (threads > 32 && (size_t)npad * threads * sizeof(A) > 96 * 1024) threads >>= 1;
  const size_t smem = (size_t)npad * threads * sizeof(A);
  if (smem > 227 * 1024) EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld does not fit in shared memory", (long long)in.R);
  const int64_t items = P.outer * P.inner;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  auto kern = k_select_smem<In, A>;
  if (smem > 48 * 1024)
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)blocks, threads, smem, s>>>((const In*)in.data, P, npad);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace

// set by tests through EOC_SELECT_FORCE_SMEM=1 to cross-check the two implementations on the GPU
static bool force_smem() {
  const char* e = getenv("EOC_SELECT_FORCE_SMEM");
  return e && e[0] == '1';
}

int launch_select(const eoc_view& in, int ignore_nodata, void* median_out, int nq, const double* probs,
                  void* quant_out, int quant_dtype, cudaStream_t s) {
  if (nq < 0 || nq > EOC_SELECT_MAX_Q) EOC_FAIL(EOC_ERR_INVALID, "nq=%d outside [0, %d]", nq, EOC_SELECT_MAX_Q);
  if (nq > 0 && (!probs || !quant_out)) EOC_FAIL(EOC_ERR_INVALID, "nq > 0 needs probs and quant_out");
  if (nq == 0 && !median_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute");
  if (quant_dtype != EOC_F64 && quant_dtype != EOC_F32) EOC_FAIL(EOC_ERR_INVALID, "quant_dtype must be f32 or f64");
  if (in.R < 1) EOC_FAIL(EOC_ERR_INVALID, "R must be >= 1 (empty axes are handled by the host layer)");
  if (in.R > EOC_SELECT_MAX_R)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld exceeds EOC_SELECT_MAX_R=%d", (long long)in.R, EOC_SELECT_MAX_R);
  if (in.outer * in.inner == 0) return EOC_OK;
  SelectParams P;
  P.outer = in.outer;
  P.R = in.R;
  P.inner = in.inner;
  P.so = in.stride_outer;
  P.sr = in.stride_R;
  P.ignore_nodata = ignore_nodata;
  P.median_out = median_out;
  P.nq = nq;
  for (int q = 0; q < EOC_SELECT_MAX_Q; ++q) P.probs[q] = q < nq ? probs[q] : 0.0;
  P.quant_out = quant_out;
  P.quant_is_f64 = quant_dtype == EOC_F64;
  P.zero = 0.0f;

  const bool stride32 = (uint64_t)in.stride_R * dtype_size(in.dtype) < (1ull << 32);
  const bool regs_ok = !force_smem() && stride32 &&
                       ((in.dtype == EOC_F32 && in.R <= 128) || (in.dtype == EOC_F64 && in.R <= 64));
  if (regs_ok) {
    if (in.dtype == EOC_F32)
      return nq == 0 ? launch_regs_n<float, float, false>(in, P, s) : launch_regs_n<float, float, true>(in, P, s);
    return nq == 0 ? launch_regs_n<double, double, false>(in, P, s) : launch_regs_n<double, double, true>(in, P, s);
  }
  switch (in.dtype) {
    case EOC_F32: return launch_smem<float>(in, P, s);
    case EOC_F64: return launch_smem<double>(in, P, s);
    case EOC_U8: return launch_smem<uint8_t>(in, P, s);
    case EOC_I16: return launch_smem<int16_t>(in, P, s);
    case EOC_U16: return launch_smem<uint16_t>(in, P, s);
    case EOC_I32: return launch_smem<int32_t>(in, P, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

}  // namespace eoc
