// Binary elementwise processes with numpy broadcasting: the overlap resolvers of merge_cubes
// (cubes/merge.py:54-295 reduces a 2-long "cubes" dimension with a process of x and y) and the arithmetic /
// comparison / logic processes `apply` chains.
//   arithmetic  math.py:98-115        x / y, x - y, x * y, x + y in the operands' common dtype (numpy: integers
//                                     wrap; true division of integers is float64)
//   comparison  comparison.py:45-127  np.where(notnull(x) & notnull(y), x OP y, nan)  -> float64 1 / 0 / NaN
//   logic       logic.py:11-46        the reference's three-valued and / or / xor       -> float64 1 / 0 / NaN
// One pass, HBM-bound: sizeof(x) + sizeof(y) + sizeof(out) bytes per sample when nothing is broadcast; a
// broadcast operand (stride 0 along a dimension) is re-read from L2.  The output is dense C order; the host
// merges adjacent dimensions, so equal shapes run as ONE dimension with 16-byte loads and stores and no index
// arithmetic; otherwise a thread owns VEC adjacent samples of the innermost dimension and derives the
// operands' offsets from the outer coordinates.
#include <type_traits>

#include "column_kernel.cuh"

namespace eoc {

namespace {

constexpr int MAXD = EOC_BINARY_MAX_DIMS;

struct BinShape {
  int nd;                // collapsed dimensions, innermost last
  int64_t shape[MAXD];
  int64_t sa[MAXD], sb[MAXD];  // element strides (0 = broadcast)
  double scalar;         // operand that is a python scalar (side: 0 none, 1 = x, 2 = y)
  int scalar_side;
};

template <int OP> struct is_arith { static constexpr bool value = OP <= EOC_BIN_DIVIDE; };

// numpy result type of (In OP In): In itself, except true division of integers (float64)
template <typename In, int OP> struct bin_out {
  using type = typename std::conditional<
      is_arith<OP>::value,
      typename std::conditional<OP == EOC_BIN_DIVIDE && !std::is_floating_point<In>::value, double, In>::type,
      double>::type;
};

template <typename T> __device__ __forceinline__ bool nan_of(T x) {
  if constexpr (std::is_floating_point<T>::value)
    return x != x;
  else
    return false;
}

template <typename In, int OP> __device__ __forceinline__ typename bin_out<In, OP>::type bin_apply(In x, In y) {
  using O = typename bin_out<In, OP>::type;
  if constexpr (OP == EOC_BIN_ADD || OP == EOC_BIN_SUBTRACT || OP == EOC_BIN_MULTIPLY) {
    if constexpr (std::is_floating_point<In>::value) {
      return OP == EOC_BIN_ADD ? add_rn(x, y) : (OP == EOC_BIN_SUBTRACT ? sub_rn(x, y) : mul_rn(x, y));
    } else {  // numpy integer arithmetic wraps: do it unsigned (defined behaviour), same bits
      using U = typename std::make_unsigned<In>::type;
      const U a = (U)x, b = (U)y;
      return (In)(U)(OP == EOC_BIN_ADD ? a + b : (OP == EOC_BIN_SUBTRACT ? a - b : a * b));
    }
  } else if constexpr (OP == EOC_BIN_DIVIDE) {
    return div_rn((O)x, (O)y);
  } else if constexpr (OP >= EOC_BIN_EQ && OP <= EOC_BIN_LTE) {
    if (nan_of(x) || nan_of(y)) return Lim<double>::nan();
    bool r;
    if constexpr (OP == EOC_BIN_EQ) r = x == y;
    if constexpr (OP == EOC_BIN_NEQ) r = !(x == y);
    if constexpr (OP == EOC_BIN_GT) r = x > y;
    if constexpr (OP == EOC_BIN_LT) r = x < y;
    if constexpr (OP == EOC_BIN_LTE) r = x <= y;
    if constexpr (OP == EOC_BIN_GTE) {  // comparison.py:117-119: (x - y) >= 0, in the operands' dtype
      if constexpr (std::is_floating_point<In>::value) {
        r = sub_rn(x, y) >= (In)0;
      } else {
        using U = typename std::make_unsigned<In>::type;
        r = (In)(U)((U)x - (U)y) >= (In)0;
      }
    }
    return r ? 1.0 : 0.0;
  } else {
    const bool nx = nan_of(x), ny = nan_of(y);
    const bool tx = x != (In)0, ty = y != (In)0;  // NaN is truthy, like np.logical_and(nan, 1)
    if constexpr (OP == EOC_BIN_AND) {            // logic.py:11-19: NaN only where the other side is true
      const bool r = tx && ty;
      return (r && (nx || ny)) ? Lim<double>::nan() : (r ? 1.0 : 0.0);
    } else if constexpr (OP == EOC_BIN_OR) {      // logic.py:22-32: nan_to_num first, NaN where the rest is false
      const bool r = (tx && !nx) || (ty && !ny);
      return (!r && (nx || ny)) ? Lim<double>::nan() : (r ? 1.0 : 0.0);
    } else {                                      // logic.py:35-40
      return (nx || ny) ? Lim<double>::nan() : ((tx != ty) ? 1.0 : 0.0);
    }
  }
}

// 16-byte (or narrower) load through the read-only path WITHOUT the evict-first hint of ld_vec: for an operand
// that is broadcast along an outer dimension and therefore re-read (from L2) once per output row
template <typename In, int VEC> __device__ __forceinline__ Vec<In, VEC> ld_vec_reuse(const In* p) {
  Vec<In, VEC> r;
  if constexpr (sizeof(In) * VEC == 16) {
    const int4 q = __ldg(reinterpret_cast<const int4*>(p));
    r = *reinterpret_cast<const Vec<In, VEC>*>(&q);
  } else if constexpr (sizeof(In) * VEC == 8) {
    const int2 q = __ldg(reinterpret_cast<const int2*>(p));
    r = *reinterpret_cast<const Vec<In, VEC>*>(&q);
  } else {
#pragma unroll
    for (int k = 0; k < VEC; ++k) r.v[k] = __ldg(p + k);
  }
  return r;
}

// dense fast path: equal shapes (after merging: one dimension), or one scalar operand
template <typename In, int OP, int VEC, int SIDE>
__global__ void __launch_bounds__(256) k_binary_flat(const In* __restrict__ a, const In* __restrict__ b,
                                                     typename bin_out<In, OP>::type* __restrict__ out, int64_t nvec,
                                                     In scalar) {
  using O = typename bin_out<In, OP>::type;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += step) {
    Vec<In, VEC> qa, qb;
    if constexpr (SIDE != 1) qa = ld_vec<In, VEC>(a + v * VEC);
    if constexpr (SIDE != 2) qb = ld_vec<In, VEC>(b + v * VEC);
    O y[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k)
      y[k] = bin_apply<In, OP>(SIDE == 1 ? scalar : qa.v[k], SIDE == 2 ? scalar : qb.v[k]);
    if constexpr (sizeof(O) * VEC <= 16 || (sizeof(O) * VEC) % 16 == 0) {
      st_vec<O, VEC>(out + v * VEC, y);
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) out[v * VEC + k] = y[k];
    }
  }
}

// broadcasting path.  The outer coordinates of an output row (and with them the operands' row offsets) cost a
// chain of 64-bit divisions; ROWWISE (rows of >= 64 vectors): a CTA walks one row at a time, the divisions are
// uniform per CTA and the threads stride over the row; otherwise a thread owns VEC adjacent samples of one row
// and does the divisions itself (short rows: few samples per division, but such cubes are small).
template <typename In, int OP, int VEC, bool ROWWISE>
__global__ void __launch_bounds__(256) k_binary_bcast(const In* __restrict__ a, const In* __restrict__ b,
                                                      typename bin_out<In, OP>::type* __restrict__ out,
                                                      const BinShape S, int64_t vec_per_row, int64_t nunits) {
  using O = typename bin_out<In, OP>::type;
  const int last = S.nd - 1;
  const int64_t inner = S.shape[last];
  const int64_t ia = S.sa[last], ib = S.sb[last];
  bool ra = false, rb = false;  // operand broadcast along an outer dimension: its rows are re-read, keep them in L2
  for (int d = 0; d < last; ++d) {
    ra = ra || S.sa[d] == 0;
    rb = rb || S.sb[d] == 0;
  }
  auto offsets = [&](int64_t row, int64_t& oa, int64_t& ob) {
    oa = ob = 0;
    for (int d = last - 1; d >= 0; --d) {
      const int64_t q = row / S.shape[d];
      const int64_t c = row - q * S.shape[d];
      oa += c * S.sa[d];
      ob += c * S.sb[d];
      row = q;
    }
  };
  auto work = [&](const In* pa, const In* pb, O* po, int64_t i0) {  // VEC samples from column i0 of a row
    if (VEC > 1 && i0 + VEC <= inner) {
      Vec<In, VEC> qa, qb;
      const In* xa = pa + i0 * ia;
      const In* xb = pb + i0 * ib;
      if (ia == 1 && ((uintptr_t)xa % (sizeof(In) * VEC)) == 0) {
        qa = ra ? ld_vec_reuse<In, VEC>(xa) : ld_vec<In, VEC>(xa);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) qa.v[k] = __ldg(xa + k * ia);
      }
      if (ib == 1 && ((uintptr_t)xb % (sizeof(In) * VEC)) == 0) {
        qb = rb ? ld_vec_reuse<In, VEC>(xb) : ld_vec<In, VEC>(xb);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) qb.v[k] = __ldg(xb + k * ib);
      }
      O y[VEC];
#pragma unroll
      for (int k = 0; k < VEC; ++k) y[k] = bin_apply<In, OP>(qa.v[k], qb.v[k]);
      if (((uintptr_t)(po + i0) % (sizeof(O) * VEC)) == 0) {
        st_vec<O, VEC>(po + i0, y);
      } else {  // rows need not be a multiple of VEC long
#pragma unroll
        for (int k = 0; k < VEC; ++k) po[i0 + k] = y[k];
      }
    } else {
      for (int k = 0; k < VEC && i0 + k < inner; ++k)
        po[i0 + k] = bin_apply<In, OP>(__ldg(pa + (i0 + k) * ia), __ldg(pb + (i0 + k) * ib));
    }
  };
  if constexpr (ROWWISE) {
    // unit = (chunk of 1024 vectors, row), rows fastest: enough CTAs even for a few very long rows, and the CTAs
    // running at the same time share the chunk of an operand that is broadcast over the rows (L2 hits)
    const int64_t chunks = (vec_per_row + 1023) / 1024;
    const int64_t rows = nunits / chunks;
    for (int64_t u = blockIdx.x; u < nunits; u += gridDim.x) {
      const int64_t chunk = u / rows;
      const int64_t row = u - chunk * rows;
      const int64_t v0 = chunk * 1024;
      int64_t oa, ob;
      offsets(row, oa, ob);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t v = v0 + k * 256 + threadIdx.x;
        if (v < vec_per_row) work(a + oa, b + ob, out + row * inner, v * VEC);
      }
    }
  } else {
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < nunits; t += step) {
      const int64_t row = t / vec_per_row;
      int64_t oa, ob;
      offsets(row, oa, ob);
      work(a + oa, b + ob, out + row * inner, (t - row * vec_per_row) * VEC);
    }
  }
}

template <typename In, int OP>
int launch_bin_op(const void* a, const void* b, const BinShape& S, void* out, cudaStream_t s) {
  using O = typename bin_out<In, OP>::type;
  constexpr int VEC = 16 / (sizeof(In) > sizeof(O) ? sizeof(In) : sizeof(O)) < 1
                          ? 1
                          : 16 / (sizeof(In) > sizeof(O) ? sizeof(In) : sizeof(O));
  int64_t n = 1;
  for (int d = 0; d < S.nd; ++d) n *= S.shape[d];
  if (n == 0) return EOC_OK;
  const int64_t cap = (int64_t)sm_count() * 32;
  const bool flat = S.scalar_side != 0 || (S.nd == 1 && S.sa[0] == 1 && S.sb[0] == 1);
  if (flat) {
    const bool al = (S.scalar_side == 1 || (uintptr_t)a % 16 == 0) && (S.scalar_side == 2 || (uintptr_t)b % 16 == 0) &&
                    (uintptr_t)out % 16 == 0;
    const int64_t nvec = al ? n / VEC : 0;
    const In sc = (In)S.scalar;
    auto go = [&](auto vec_tag, const In* pa, const In* pb, O* po, int64_t count) -> int {
      constexpr int V = decltype(vec_tag)::value;
      int64_t blocks = (count + 255) / 256;
      if (blocks > cap) blocks = cap;
      if (S.scalar_side == 1)
        k_binary_flat<In, OP, V, 1><<<(unsigned)blocks, 256, 0, s>>>(pa, pb, po, count, sc);
      else if (S.scalar_side == 2)
        k_binary_flat<In, OP, V, 2><<<(unsigned)blocks, 256, 0, s>>>(pa, pb, po, count, sc);
      else
        k_binary_flat<In, OP, V, 0><<<(unsigned)blocks, 256, 0, s>>>(pa, pb, po, count, sc);
      EOC_LAUNCH_CHECK();
      return EOC_OK;
    };
    if (nvec > 0) {
      const int rc = go(std::integral_constant<int, VEC>{}, (const In*)a, (const In*)b, (O*)out, nvec);
      if (rc != EOC_OK) return rc;
    }
    const int64_t done = nvec * VEC;
    if (done < n)
      return go(std::integral_constant<int, 1>{}, a ? (const In*)a + done : nullptr, b ? (const In*)b + done : nullptr,
                (O*)out + done, n - done);
    return EOC_OK;
  }
  const int64_t inner = S.shape[S.nd - 1];
  const int64_t vec_per_row = (inner + VEC - 1) / VEC;
  const int64_t rows = n / inner;
  if (vec_per_row >= 64) {
    const int64_t units = rows * ((vec_per_row + 1023) / 1024);
    const int64_t blocks = units < cap ? units : cap;
    k_binary_bcast<In, OP, VEC, true><<<(unsigned)blocks, 256, 0, s>>>((const In*)a, (const In*)b, (O*)out, S,
                                                                       vec_per_row, units);
  } else {
    const int64_t nthreads = rows * vec_per_row;
    int64_t blocks = (nthreads + 255) / 256;
    if (blocks > cap) blocks = cap;
    k_binary_bcast<In, OP, VEC, false><<<(unsigned)blocks, 256, 0, s>>>((const In*)a, (const In*)b, (O*)out, S,
                                                                        vec_per_row, nthreads);
  }
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

template <typename In> int launch_bin_t(const void* a, const void* b, const BinShape& S, int op, void* out, cudaStream_t s) {
  switch (op) {
    case EOC_BIN_ADD: return launch_bin_op<In, EOC_BIN_ADD>(a, b, S, out, s);
    case EOC_BIN_SUBTRACT: return launch_bin_op<In, EOC_BIN_SUBTRACT>(a, b, S, out, s);
    case EOC_BIN_MULTIPLY: return launch_bin_op<In, EOC_BIN_MULTIPLY>(a, b, S, out, s);
    case EOC_BIN_DIVIDE: return launch_bin_op<In, EOC_BIN_DIVIDE>(a, b, S, out, s);
    case EOC_BIN_EQ: return launch_bin_op<In, EOC_BIN_EQ>(a, b, S, out, s);
    case EOC_BIN_NEQ: return launch_bin_op<In, EOC_BIN_NEQ>(a, b, S, out, s);
    case EOC_BIN_GT: return launch_bin_op<In, EOC_BIN_GT>(a, b, S, out, s);
    case EOC_BIN_GTE: return launch_bin_op<In, EOC_BIN_GTE>(a, b, S, out, s);
    case EOC_BIN_LT: return launch_bin_op<In, EOC_BIN_LT>(a, b, S, out, s);
    case EOC_BIN_LTE: return launch_bin_op<In, EOC_BIN_LTE>(a, b, S, out, s);
    case EOC_BIN_AND: return launch_bin_op<In, EOC_BIN_AND>(a, b, S, out, s);
    case EOC_BIN_OR: return launch_bin_op<In, EOC_BIN_OR>(a, b, S, out, s);
    case EOC_BIN_XOR: return launch_bin_op<In, EOC_BIN_XOR>(a, b, S, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: unknown op %d", op);
}

}  // namespace

int launch_binary(const void* a, const void* b, int dtype, int ndim, const int64_t* shape, const int64_t* sa,
                  const int64_t* sb, int op, double scalar, int scalar_side, void* out, cudaStream_t s) {
  if (ndim < 0 || ndim > MAXD) EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: ndim %d outside 0..%d", ndim, MAXD);
  if (scalar_side < 0 || scalar_side > 2) EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: scalar_side %d", scalar_side);
  if ((scalar_side != 1 && !a) || (scalar_side != 2 && !b) || !out || (ndim > 0 && !shape))
    EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: NULL pointer");
  if (scalar_side == 0 && ndim > 0 && (!sa || !sb)) EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: NULL strides");
  // merge adjacent dimensions that both operands walk contiguously (or both broadcast), drop 1-long ones
  BinShape S;
  S.nd = 0;
  S.scalar = scalar;
  S.scalar_side = scalar_side;
  for (int d = 0; d < ndim; ++d) {
    if (shape[d] < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: negative extent");
    if (shape[d] == 1) continue;
    const int64_t a_s = scalar_side ? 0 : sa[d], b_s = scalar_side ? 0 : sb[d];
    if (a_s < 0 || b_s < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: negative strides are not supported");
    if (S.nd > 0) {
      const int p = S.nd - 1;
      if (scalar_side || (S.sa[p] == a_s * shape[d] && S.sb[p] == b_s * shape[d])) {
        S.shape[p] *= shape[d];
        S.sa[p] = a_s;
        S.sb[p] = b_s;
        continue;
      }
    }
    S.shape[S.nd] = shape[d];
    S.sa[S.nd] = a_s;
    S.sb[S.nd] = b_s;
    ++S.nd;
  }
  if (S.nd == 0) {  // a single sample
    S.nd = 1;
    S.shape[0] = 1;
    S.sa[0] = S.sb[0] = 1;
  }
  if (scalar_side) S.sa[0] = S.sb[0] = 1;
  switch (dtype) {
    case EOC_F32: return launch_bin_t<float>(a, b, S, op, out, s);
    case EOC_F64: return launch_bin_t<double>(a, b, S, op, out, s);
    case EOC_U8: return launch_bin_t<uint8_t>(a, b, S, op, out, s);
    case EOC_I16: return launch_bin_t<int16_t>(a, b, S, op, out, s);
    case EOC_U16: return launch_bin_t<uint16_t>(a, b, S, op, out, s);
    case EOC_I32: return launch_bin_t<int32_t>(a, b, S, op, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "eoc_binary: unknown dtype %d", dtype);
}

}  // namespace eoc
