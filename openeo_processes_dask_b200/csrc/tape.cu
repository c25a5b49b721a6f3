// eoc_apply_tape -- a chain of elementwise processes in ONE pass (`apply`, cubes/apply.py:24-39).
//
// The reference runs the child process graph of `apply` node by node: every node is a numpy expression that reads
// and writes the whole cube (linear_scale_range(clip((a - b) / (a + b))) = 4+ reads and writes).  Here the chain is
// recorded by the host layer (LazyElementwise) and evaluated per sample by a tiny accumulator machine:
//   acc <- operand | acc <- acc OP operand | acc <- f(acc) | t0 / t1 <- acc
// with operand one of {immediate, x, y, t0, t1} -- named values, no indexed register file (that would live in local
// memory).  The tape is uniform over the grid, so the switch is branch-uniform; a 10-step chain is ~60 instructions
// per sample: still HBM-bound at 8-12 bytes per sample.  Every step rounds to the compute dtype (float32 for
// float32 input, else float64) with the *_rn intrinsics -- no FMA contraction -- i.e. exactly numpy's operation by
// operation arithmetic for + - * / sqrt abs sign floor ceil trunc min max; exp, ln, the trigonometric and
// hyperbolic functions, power and arctan2 use CUDA's double-precision routines (<= 2 ulp in float64; float32 results
// are the correctly rounded double results), which agree with numpy's within the 1e-6 relative tolerance of the
// path but are not bit-identical (libm implementations differ by an ulp).
#include <type_traits>

#include "column_kernel.cuh"

namespace eoc {
namespace {

struct TapeOp {
  int code, operand;  // operand: 0 immediate, 1 x, 2 y, 3 t0, 4 t1
  double imm;
};
struct Tape {
  int n;
  TapeOp op[EOC_TAPE_MAX_OPS];
};

template <typename A> __device__ __forceinline__ A tape_unary(A v, int code, A imm) {
  const double d = (double)v;
  switch (code) {
    case EOC_TAPE_ABS: return abs_(v);
    case EOC_TAPE_SGN: return is_nan(v) ? v : (v > (A)0 ? (A)1 : (v < (A)0 ? (A)-1 : (A)0));
    case EOC_TAPE_SQRT: return sqrt_rn(v);
    case EOC_TAPE_FLOOR: return floor(v);
    case EOC_TAPE_CEIL: return ceil(v);
    case EOC_TAPE_TRUNC: return trunc(v);
    case EOC_TAPE_ROUND: {  // np.around(x, p): rint(x * 10^p) / 10^p (p >= 0), rint(x / 10^-p) * 10^-p (p < 0)
      const A s = imm;  // 10^|p|, sign of p carried by the code's operand field being unused: imm > 0 always
      return div_rn(rint(mul_rn(v, s)), s);
    }
    case EOC_TAPE_ROUND_NEG: return mul_rn(rint(div_rn(v, imm)), imm);
    case EOC_TAPE_EXP: return (A)exp(d);
    case EOC_TAPE_LN: return (A)log(d);
    case EOC_TAPE_SIN: return (A)sin(d);
    case EOC_TAPE_COS: return (A)cos(d);
    case EOC_TAPE_TAN: return (A)tan(d);
    case EOC_TAPE_ASIN: return (A)asin(d);
    case EOC_TAPE_ACOS: return (A)acos(d);
    case EOC_TAPE_ATAN: return (A)atan(d);
    case EOC_TAPE_SINH: return (A)sinh(d);
    case EOC_TAPE_COSH: return (A)cosh(d);
    case EOC_TAPE_TANH: return (A)tanh(d);
    case EOC_TAPE_ASINH: return (A)asinh(d);
    case EOC_TAPE_ACOSH: return (A)acosh(d);
    default: return (A)atanh(d);
  }
}

// np.mod: the remainder carries the sign of the divisor (python's %), NaN for a zero divisor
__device__ __forceinline__ double np_mod(double a, double b) {
  if (b == 0.0) return Lim<double>::nan();
  double m = fmod(a, b);
  if (m != 0.0 && ((m < 0.0) != (b < 0.0))) m += b;
  if (m == 0.0) m = copysign(0.0, b);
  return m;
}

template <typename A> __device__ __forceinline__ A tape_binary(A acc, A o, int code) {
  switch (code) {
    case EOC_TAPE_ADD: return add_rn(acc, o);
    case EOC_TAPE_SUB: return sub_rn(acc, o);
    case EOC_TAPE_RSUB: return sub_rn(o, acc);
    case EOC_TAPE_MUL: return mul_rn(acc, o);
    case EOC_TAPE_DIV: return div_rn(acc, o);
    case EOC_TAPE_RDIV: return div_rn(o, acc);
    case EOC_TAPE_MAXNUM: return (is_nan(acc) || is_nan(o)) ? (is_nan(acc) ? acc : o) : (acc > o ? acc : o);  // np.maximum
    case EOC_TAPE_MINNUM: return (is_nan(acc) || is_nan(o)) ? (is_nan(acc) ? acc : o) : (acc < o ? acc : o);  // np.minimum
    case EOC_TAPE_POW: return (A)pow((double)acc, (double)o);
    case EOC_TAPE_RPOW: return (A)pow((double)o, (double)acc);
    case EOC_TAPE_MOD: return (A)np_mod((double)acc, (double)o);
    case EOC_TAPE_RMOD: return (A)np_mod((double)o, (double)acc);
    case EOC_TAPE_ATAN2: return (A)atan2((double)acc, (double)o);
    default: return (A)atan2((double)o, (double)acc);
  }
}

template <typename A> __device__ __forceinline__ A tape_eval(const A vx, const A vy, const Tape& T) {
  A acc = vx, t0 = (A)0, t1 = (A)0;
  for (int s = 0; s < T.n; ++s) {
    const int code = T.op[s].code, which = T.op[s].operand;
    const A imm = (A)T.op[s].imm;
    const A o = which == 0 ? imm : (which == 1 ? vx : (which == 2 ? vy : (which == 3 ? t0 : t1)));
    if (code == EOC_TAPE_LOAD)
      acc = o;
    else if (code == EOC_TAPE_STORE_T0)
      t0 = acc;
    else if (code == EOC_TAPE_STORE_T1)
      t1 = acc;
    else if (code < EOC_TAPE_ABS)
      acc = tape_binary<A>(acc, o, code);
    else
      acc = tape_unary<A>(acc, code, imm);
  }
  return acc;
}

// n samples: nvec = n / VEC vectors grid-strided over all CTAs, the n % VEC samples after them by CTA 0
template <typename In, typename A, int VEC, bool TWO>
__global__ void __launch_bounds__(256) k_tape(const In* __restrict__ x, const In* __restrict__ y, A* __restrict__ out,
                                              int64_t n, const Tape T) {
  const int64_t nvec = n / VEC;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += step) {
    const Vec<In, VEC> qx = ld_vec<In, VEC>(x + v * VEC);
    Vec<In, VEC> qy = qx;
    if constexpr (TWO) qy = ld_vec<In, VEC>(y + v * VEC);
    A res[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) res[k] = tape_eval<A>((A)qx.v[k], (A)qy.v[k], T);
    st_vec<A, VEC>(out + v * VEC, res);
  }
  if (VEC > 1 && blockIdx.x == 0) {
    for (int64_t i = nvec * VEC + threadIdx.x; i < n; i += blockDim.x)
      out[i] = tape_eval<A>((A)x[i], TWO ? (A)y[i] : (A)x[i], T);
  }
}

template <typename In, typename A>
int launch_tape_t(const void* x, const void* y, int64_t n, const Tape& T, void* out, cudaStream_t s) {
  constexpr int VM = (sizeof(A) == 4) ? 4 : 2;
  const bool al = ((uintptr_t)x % (sizeof(In) * VM) == 0) && ((uintptr_t)out % (sizeof(A) * VM) == 0) &&
                  (!y || (uintptr_t)y % (sizeof(In) * VM) == 0);
  const int64_t cap = (int64_t)sm_count() * 32;
  auto go = [&](auto vec_tag, const In* px, const In* py, A* po, int64_t nv, int64_t blocks, int threads) {
    constexpr int V = decltype(vec_tag)::value;
    if (py)
      k_tape<In, A, V, true><<<(unsigned)blocks, threads, 0, s>>>(px, py, po, nv, T);
    else
      k_tape<In, A, V, false><<<(unsigned)blocks, threads, 0, s>>>(px, px, po, nv, T);
  };
  const In* px = (const In*)x;
  const In* py = (const In*)y;
  A* po = (A*)out;
  if (al) {
    int64_t blocks = (n / VM + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > cap) blocks = cap;
    go(std::integral_constant<int, VM>(), px, py, po, n, blocks, 256);
    EOC_LAUNCH_CHECK();
    return EOC_OK;
  }
  int64_t blocks = (n + 255) / 256;
  if (blocks > cap) blocks = cap;
  go(std::integral_constant<int, 1>(), px, py, po, n, blocks, 256);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" int eoc_apply_tape(const void* x, const void* y, int dtype, int64_t n, const eoc_tape_op* ops, int nops,
                              void* out, int out_dtype, void* stream) {
  if (n < 0 || nops < 0 || nops > EOC_TAPE_MAX_OPS)
    EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape: n=%lld, nops=%d (at most %d steps)", (long long)n, nops, EOC_TAPE_MAX_OPS);
  if (n == 0) return EOC_OK;
  if (!x || !out || (nops > 0 && !ops)) EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape: NULL pointer");
  Tape T;
  T.n = nops;
  for (int s = 0; s < nops; ++s) {
    if (ops[s].code < 0 || ops[s].code > EOC_TAPE_ATANH || ops[s].operand < 0 || ops[s].operand > 4)
      EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape: step %d has code %d / operand %d", s, ops[s].code, ops[s].operand);
    if (ops[s].operand == 2 && !y) EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape: step %d reads y but y is NULL", s);
    T.op[s].code = ops[s].code;
    T.op[s].operand = ops[s].operand;
    T.op[s].imm = ops[s].imm;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (out_dtype == EOC_F32) {  // float32 arithmetic: float32 samples and the integers float32 holds exactly
    switch (dtype) {
      case EOC_F32: return launch_tape_t<float, float>(x, y, n, T, out, s);
      case EOC_U8: return launch_tape_t<uint8_t, float>(x, y, n, T, out, s);
      case EOC_I16: return launch_tape_t<int16_t, float>(x, y, n, T, out, s);
      case EOC_U16: return launch_tape_t<uint16_t, float>(x, y, n, T, out, s);
    }
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_apply_tape: float32 arithmetic on dtype %d would round the samples", dtype);
  }
  if (out_dtype != EOC_F64) EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape: out_dtype %d is neither float32 nor float64", out_dtype);
  switch (dtype) {
    case EOC_F32: return launch_tape_t<float, double>(x, y, n, T, out, s);
    case EOC_F64: return launch_tape_t<double, double>(x, y, n, T, out, s);
    case EOC_U8: return launch_tape_t<uint8_t, double>(x, y, n, T, out, s);
    case EOC_I16: return launch_tape_t<int16_t, double>(x, y, n, T, out, s);
    case EOC_U16: return launch_tape_t<uint16_t, double>(x, y, n, T, out, s);
    case EOC_I32: return launch_tape_t<int32_t, double>(x, y, n, T, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", dtype);
}
