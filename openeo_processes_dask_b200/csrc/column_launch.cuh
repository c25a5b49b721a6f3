// Launch templates of the column kernel, shared by reduce.cu (plain / normalized-difference sources) and
// masked.cu (masked source).
#pragma once

#include "column_kernel.cuh"

namespace eoc {

// elements per thread and row: 16-byte loads for 4- and 8-byte samples, 16-byte loads (8 samples) for the
// 2-byte ones as well -- with 8-byte loads the uint16 ndvi+max path kept too few bytes in flight
template <typename In> struct VecMax { static constexpr int value = 4; };
template <> struct VecMax<double> { static constexpr int value = 2; };
template <> struct VecMax<uint16_t> { static constexpr int value = 8; };
template <> struct VecMax<int16_t> { static constexpr int value = 8; };
template <> struct VecMax<uint8_t> { static constexpr int value = 8; };

inline uint32_t groups_for(uint32_t ops, int ignore_nodata) {
  // min / max alone, NaN skipped: fmin / fmax already ignore NaN and an all-NaN column stays NaN, so neither
  // the running sum nor the not-NaN count is needed (3 of the ~27 instructions per sample of the fused
  // uint16 normalized-difference -> max kernel, which is bound by issue slots)
  if (ignore_nodata && ops && !(ops & ~((1u << EOC_OP_MIN) | (1u << EOC_OP_MAX)))) return G_MINMAX | G_NOSUM;
  uint32_t g = 0;
  if (ops & ((1u << EOC_OP_MIN) | (1u << EOC_OP_MAX))) g |= G_MINMAX;
  if (ops & (1u << EOC_OP_PRODUCT)) g |= G_MINMAX | G_RARE;
  if ((ops & (1u << EOC_OP_COUNT)) && !(g & G_RARE)) g = (g & G_MINMAX) ? (G_MINMAX | G_RARE) : G_COUNT;
  if (ops & kOrderOps) {
    const bool fl = ops & ((1u << EOC_OP_FIRST) | (1u << EOC_OP_LAST));
    const bool lg = ops & ((1u << EOC_OP_ANY) | (1u << EOC_OP_ALL));
    g = (g || (fl && lg)) ? (G_MINMAX | G_RARE | G_ORDER) : (fl ? G_FIRSTLAST : G_LOGIC);
  }
  return g;
}

template <typename Source, typename A, int VEC>
int launch_columns_g(const Source& src, const ColumnParams& P, cudaStream_t s) {
  const int64_t items = P.outer * (P.inner / VEC);
  if (items == 0) return EOC_OK;
  const int threads = 256;
  const int64_t blocks = (items + threads - 1) / threads;
  if (blocks > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "cube too large for one launch");
  constexpr int U = (VEC == 1) ? 8 : 4;
  switch (groups_for(P.ops, P.ignore_nodata)) {
    case G_MINMAX | G_NOSUM:
      k_columns<Source, A, VEC, G_MINMAX | G_NOSUM, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case 0:
      k_columns<Source, A, VEC, 0, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case G_MINMAX:
      k_columns<Source, A, VEC, G_MINMAX, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case G_COUNT:
      k_columns<Source, A, VEC, G_COUNT, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case G_FIRSTLAST:
      k_columns<Source, A, VEC, G_FIRSTLAST, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case G_LOGIC:
      k_columns<Source, A, VEC, G_LOGIC, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    case G_MINMAX | G_RARE | G_ORDER:
      k_columns<Source, A, VEC, G_MINMAX | G_RARE | G_ORDER, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
    default:
      k_columns<Source, A, VEC, G_MINMAX | G_RARE, U><<<(unsigned)blocks, threads, 0, s>>>(src, P);
      break;
  }
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

inline bool aligned_for(const void* p, int64_t so, int64_t sr, int64_t inner, size_t esize, int vec) {
  const size_t bytes = esize * vec;
  return ((uintptr_t)p % bytes == 0) && (so % vec == 0) && (sr % vec == 0) && (inner % vec == 0);
}

inline int check_outs(uint32_t ops, const OutPtrs& outs) {
  if (ops >> EOC_NUM_OPS) EOC_FAIL(EOC_ERR_INVALID, "unknown reducer bit in ops mask 0x%x", ops);
  for (int id = 0; id < EOC_NUM_OPS; ++id)
    if ((ops & (1u << id)) && outs.p[id] == nullptr)
      EOC_FAIL(EOC_ERR_INVALID, "outs[%d] is NULL but selected by ops mask 0x%x", id, ops);
  return EOC_OK;
}

inline ColumnParams make_params(const eoc_view& in, uint32_t ops, int ignore_nodata, int ddof,
                         const OutPtrs& outs) {
  ColumnParams P;
  P.outer = in.outer;
  P.R = in.R;
  P.inner = in.inner;
  P.ops = ops;
  P.ignore_nodata = ignore_nodata;
  P.ddof = ddof;
  P.nseg = 1;
  P.seg_offsets = nullptr;
  P.members = nullptr;
  P.outs = outs;
  return P;
}

}  // namespace eoc
