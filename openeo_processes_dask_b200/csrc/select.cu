// Launcher of eoc_select: picks the register-network kernel for float32/float64 series that fit
// in registers, the shared-memory sorter otherwise.
#include <cstdlib>
#include <type_traits>

#include "select_launch.cuh"
#include "select_tile_launch.cuh"
#include "select_split_launch.cuh"
#include "select_split_wide.cuh"

namespace eoc {
namespace {

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

// every field of SelectParams from the call's arguments (no sorted output)
static SelectParams make_select_params(const eoc_view& in, int ignore_nodata, void* median_out, int nq,
                                       const double* probs, void* quant_out, int quant_dtype, uint32_t ops, int ddof,
                                       const OutPtrs& outs) {
  SelectParams P;
  P.outer = in.outer;
  P.R = in.R;
  P.inner = in.inner;
  P.so = in.stride_outer;
  P.sr = in.stride_R;
  P.ignore_nodata = ignore_nodata;
  P.median_out = median_out;
  P.median_f64 = 0;
  P.lerp_wrap16 = 0;
  P.nq = nq;
  for (int q = 0; q < EOC_SELECT_MAX_Q; ++q) P.probs[q] = q < nq ? probs[q] : 0.0;
  P.quant_out = quant_out;
  P.quant_is_f64 = quant_dtype == EOC_F64;
  P.zero = 0.0f;
  P.sort_out = nullptr;
  P.sort_desc = 0;
  P.sort_nan_first = 0;
  P.fused = ColumnParams{};
  P.fused.outer = in.outer;
  P.fused.R = in.R;
  P.fused.inner = in.inner;
  P.fused.ops = ops;
  P.fused.ignore_nodata = ignore_nodata;
  P.fused.ddof = ddof;
  P.fused.nseg = 1;
  P.fused.seg_offsets = nullptr;
  P.fused.members = nullptr;
  P.fused.outs = outs;
  return P;
}

// mask -> median / quantiles in one pass.  Only what the fused kernel covers; everything else is
// EOC_ERR_UNSUPPORTED and the caller materialises the masked cube first (eoc_masked_reduce with masked_out).
int launch_masked_select(const eoc_view& in, const eoc_view& mk, double replacement, int ignore_nodata,
                         void* median_out, int nq, const double* probs, void* quant_out, int quant_dtype,
                         cudaStream_t s) {
  if (mk.dtype != EOC_U8) EOC_FAIL(EOC_ERR_INVALID, "the mask must be EOC_U8 (see eoc_mask_canonical)");
  if (mk.outer != in.outer || mk.R != in.R || mk.inner != in.inner)
    EOC_FAIL(EOC_ERR_INVALID, "data and mask views differ in shape (broadcast with stride 0, not with extent 1)");
  if (nq < 0 || nq > EOC_SELECT_MAX_Q) EOC_FAIL(EOC_ERR_INVALID, "nq=%d outside [0, %d]", nq, EOC_SELECT_MAX_Q);
  if (nq > 0 && (!probs || !quant_out)) EOC_FAIL(EOC_ERR_INVALID, "nq > 0 needs probs and quant_out");
  if (nq == 0 && !median_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute");
  if (quant_dtype != EOC_F64 && quant_dtype != EOC_F32) EOC_FAIL(EOC_ERR_INVALID, "quant_dtype must be f32 or f64");
  if (in.outer * in.inner == 0) return EOC_OK;
  if (replacement == replacement) EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused masked selection needs replacement = NaN");
  const bool raw_int = in.dtype == EOC_U16 || in.dtype == EOC_I16 || in.dtype == EOC_U8;
  if ((in.dtype != EOC_F32 && !raw_int) || in.R <= 32 || in.R > 128 || !tma_view_ok(in))
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused masked selection: float32 or 8 / 16-bit integers, 32 < R <= 128, TMA-addressable view");
  if ((uintptr_t)mk.data % 4 || mk.stride_R % 4 || mk.stride_outer % 4 || in.inner % 4)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused masked selection: mask base, strides and inner must be multiples of 4");
  {
    const char* e = getenv("EOC_SELECT_TMA");
    if (!(e && e[0] == '1') && !tma_tiles_fill(in, 128))
      EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused masked selection: inner=%lld would leave the 128-column tiles mostly empty",
               (long long)in.inner);
  }
  if (raw_int) {
    // raw digital numbers under a cloud mask: where(mask, NaN) is float64 in numpy, so are the results
    if (nq != 0 && quant_dtype != EOC_F64) EOC_FAIL(EOC_ERR_INVALID, "integer input: quantiles are float64");
    if (!select_quad_applies(in, nq, false, QUAD_MASKED) || split_env("EOC_QUAD_INT", 1) != 1)
      EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused masked selection of integer samples needs the four-run kernel");
    SelectParams PI = make_select_params(in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, 0u, 1, OutPtrs{});
    PI.median_f64 = 1;
    return launch_select_quad_masked_int(in, mk, PI, s);
  }
  const SelectParams P = make_select_params(in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, 0u, 1,
                                            OutPtrs{});
  if (select_quad_applies(in, nq, false, QUAD_MASKED)) return launch_select_quad_masked(in, mk, P, s);
  if (select_split_applies(in, nq, false, false)) return launch_select_split_masked(in, mk, P, s);
  return launch_select_tile_masked(in, mk, P, nq != 0, s);
}

int launch_select(const eoc_view& in, int ignore_nodata, void* median_out, int nq, const double* probs,
                  void* quant_out, int quant_dtype, uint32_t ops, int ddof, const OutPtrs& outs, cudaStream_t s,
                  const SortSpec& sort) {
  if (ops >> EOC_NUM_OPS) EOC_FAIL(EOC_ERR_INVALID, "unknown reducer bit in ops mask 0x%x", ops);
  for (int id = 0; id < EOC_NUM_OPS; ++id)
    if ((ops & (1u << id)) && outs.p[id] == nullptr)
      EOC_FAIL(EOC_ERR_INVALID, "outs[%d] is NULL but selected by ops mask 0x%x", id, ops);
  if (nq < 0 || nq > EOC_SELECT_MAX_Q) EOC_FAIL(EOC_ERR_INVALID, "nq=%d outside [0, %d]", nq, EOC_SELECT_MAX_Q);
  if (nq > 0 && (!probs || !quant_out)) EOC_FAIL(EOC_ERR_INVALID, "nq > 0 needs probs and quant_out");
  if (nq == 0 && !median_out && !sort.out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute");
  if (quant_dtype != EOC_F64 && quant_dtype != EOC_F32) EOC_FAIL(EOC_ERR_INVALID, "quant_dtype must be f32 or f64");
  if (in.R < 1) EOC_FAIL(EOC_ERR_INVALID, "R must be >= 1 (empty axes are handled by the host layer)");
  if (in.outer * in.inner == 0) return EOC_OK;
  SelectParams P = make_select_params(in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, ops, ddof, outs);
  P.sort_out = sort.out;
  P.sort_desc = sort.descending;
  P.sort_nan_first = sort.nan_first;
  const bool want_sorted = sort.out != nullptr;
  if (in.R > EOC_SELECT_MAX_R) {
    // 513 .. 1024 float32 samples on a view the eight-run kernel takes: beyond the documented limit, but served
    // (the host layer tries this before it falls back to eoc_select_general)
    if (in.R <= 1024 && in.dtype == EOC_F32 && !ops && !want_sorted && nq <= EOC_SPLIT_MAX_Q && !force_smem() &&
        select_multi_enabled() && split_env("EOC_SELECT_SPLIT", -1) != 0 && tma_view_ok(in)) {
      const char* e = getenv("EOC_SELECT_TMA");
      const bool want = (e && e[0]) ? e[0] == '1' : (in.outer * in.inner >= 65536 && tma_tiles_fill(in, 32));
      if (want) return launch_select_multi(in, P, s);
    }
    // the same for raw 8 / 16-bit integer cubes (float64 results)
    const bool raw_int = in.dtype == EOC_U8 || in.dtype == EOC_I16 || in.dtype == EOC_U16;
    if (in.R <= 1024 && raw_int && !ops && !want_sorted && nq <= EOC_SPLIT_MAX_Q && !force_smem() && select_multi_enabled() &&
        split_env("EOC_SELECT_SPLIT", -1) != 0 && split_env("EOC_QUAD_INT", 1) == 1 && tma_view_ok(in) &&
        (nq == 0 || P.quant_is_f64)) {
      const char* e = getenv("EOC_SELECT_TMA");
      const bool want = (e && e[0]) ? e[0] == '1' : (in.outer * in.inner >= 65536 && tma_tiles_fill(in, 32));
      if (want) {
        P.median_f64 = 1;
        P.lerp_wrap16 = in.dtype == EOC_I16;
        return launch_select_multi_int(in, P, s);
      }
    }
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "R=%lld exceeds EOC_SELECT_MAX_R=%d", (long long)in.R, EOC_SELECT_MAX_R);
  }

  const bool stride32 = (uint64_t)in.stride_R * dtype_size(in.dtype) < (1ull << 32);
  const bool regs_ok = !force_smem() && stride32 &&
                       ((in.dtype == EOC_F32 && in.R <= 128) || (in.dtype == EOC_F64 && in.R <= 64));
  if (ops) {
    // one pass over HBM when the float32 register kernel applies; otherwise the streaming reducers run as
    // their own launch after the selection (same results, the cube is read twice)
    const bool one_pass = regs_ok && !(ops & kOrderOps) && !want_sorted;
    if (one_pass && select_tile_applies(in, nq != 0, true)) {
      if (select_quad_applies(in, nq, false, QUAD_FUSED)) return launch_select_quad_fused(in, P, fused_level(ops), s);
      if (select_split_applies(in, nq, false, false)) return launch_select_split_fused(in, P, fused_level(ops), s);
      return launch_select_tile_fused(in, P, nq != 0, fused_level(ops), s);
    }
    if (one_pass && in.dtype == EOC_F32) return launch_select_fused_f32(in, P, nq != 0, fused_level(ops), s);
    if (int rc = launch_select(in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, 0u, ddof, outs, s, sort))
      return rc;
    if (reduce_workspace_bytes(in, ops) != 0)
      EOC_FAIL(EOC_ERR_UNSUPPORTED, "fused reducers over a long contiguous axis are not supported");
    return launch_reduce(in, ops, ignore_nodata, ddof, outs, nullptr, 0, s);
  }
  // float64 columns of 33 .. 128 samples on a TMA-addressable view with enough columns: the four-run kernel for doubles
  // (129 .. 512 samples: eight runs, select_multi64.cu)
  if (in.dtype == EOC_F64 && !force_smem() && !want_sorted && in.R > 32 && in.R <= 512 && nq <= EOC_SPLIT_MAX_Q &&
      split_env("EOC_SELECT_QUAD", 1) == 1 && split_env("EOC_QUAD64", 1) == 1 && tma_view_ok(in)) {
    const char* e = getenv("EOC_SELECT_TMA");
    const bool want = (e && e[0]) ? e[0] == '1' : (in.outer * in.inner >= 65536 && tma_tiles_fill(in, 64));
    if (want) return in.R <= 128 ? launch_select_quad64(in, P, s) : launch_select_multi64(in, P, s);
  }
  // the "quantile" variants sort the whole column (data-dependent ranks); a sorted output needs them too
  const bool full_sort = nq != 0 || want_sorted;
  if (regs_ok && select_quad_short_applies(in, nq, want_sorted)) return launch_select_quad(in, P, s);
  if (regs_ok && select_tile_applies(in, full_sort, false)) {
    if (select_quad_applies(in, nq, want_sorted)) return launch_select_quad(in, P, s);
    if (select_split_applies(in, nq, want_sorted, nq == 0)) return launch_select_split(in, P, s);
    return launch_select_tile(in, P, full_sort, s);
  }
  if (regs_ok) {
    if (in.dtype == EOC_F32)
      return !full_sort ? launch_regs_n<float, float, false, 0>(in, P, s) : launch_regs_n<float, float, true, 0>(in, P, s);
    return !full_sort ? launch_regs_n<double, double, false, 0>(in, P, s) : launch_regs_n<double, double, true, 0>(in, P, s);
  }
  // 8/16-bit integer cubes: the float32 networks with a conversion on load, float64 results (select_int.cuh);
  // a sorted output keeps the old path (np.sort of integers is integer: the host layer casts back)
  const bool small_int = in.dtype == EOC_U8 || in.dtype == EOC_I16 || in.dtype == EOC_U16;
  if (!force_smem() && small_int && !want_sorted && (uint64_t)in.stride_R * dtype_size(in.dtype) < (1ull << 32)) {
    P.median_f64 = 1;
    P.lerp_wrap16 = in.dtype == EOC_I16;
    if (nq != 0 && !P.quant_is_f64) EOC_FAIL(EOC_ERR_INVALID, "integer input: quantiles are float64");
    // 64 < R <= 128 on a TMA-addressable view with enough columns: the four-run kernel, converting in its register
    // fill (EOC_QUAD_INT=0: the register kernels below)
    if (split_env("EOC_QUAD_INT", 1) == 1 && tma_view_ok(in)) {
      const char* e = getenv("EOC_SELECT_TMA");
      const bool want = (e && e[0]) ? e[0] == '1' : (in.outer * in.inner >= 65536 && tma_tiles_fill(in, 128));
      if (want && select_quad_applies(in, nq, false)) return launch_select_quad_int(in, P, s);
      // 128 < R <= 512: eight runs, several threads per column (select_multi.cu)
      if (want && in.R > 128 && in.R <= 512 && nq <= EOC_SPLIT_MAX_Q && select_multi_enabled() &&
          split_env("EOC_SELECT_SPLIT", -1) != 0)
        return launch_select_multi_int(in, P, s);
    }
    if (in.dtype == EOC_U8) return launch_select_int_u8(in, P, nq != 0, s);
    if (in.dtype == EOC_I16) return launch_select_int_i16(in, P, nq != 0, s);
    return launch_select_int_u16(in, P, nq != 0, s);
  }
  // long float32 columns (128 < R <= 512): K sorted runs in the TMA tile + rank elimination (select_split_wide.cuh);
  // views TMA cannot address, small cubes, sorted output: K threads per column with a bitonic merge (select_wide.cu)
  if (!force_smem() && select_split_wide_applies(in, nq, want_sorted))
    return select_multi_enabled() ? launch_select_multi(in, P, s) : launch_select_split_wide(in, P, s);
  if (!force_smem() && in.dtype == EOC_F32 && select_wide_applies(in)) return launch_select_wide(in, P, s);
  if (!force_smem() && select_wide64_applies(in)) return launch_select_wide64(in, P, s);
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
