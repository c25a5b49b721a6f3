// Elementwise processes that `apply` runs over a whole cube: clip (math.py:332-338) and
// linear_scale_range (math.py:292-298).  One pass, 16-byte loads and stores, grid-stride over a dense array:
// sizeof(in) + sizeof(out) bytes per sample, HBM-bound.  The arithmetic is numpy's, operation by operation in
// the compute dtype (float32 for float32 input -- python scalars are weak under NEP 50 -- else float64),
// without contraction: clip = minimum(maximum(x, lo), hi) with NaN passing through, then
// ((x - inMin) / (inMax - inMin)) * (outMax - outMin) + outMin with the two differences formed on the host in
// double and rounded to the compute dtype once, exactly like the python expressions of the reference.
#include <type_traits>

#include "column_kernel.cuh"

namespace eoc {

namespace {

template <typename A> struct ElemParams {
  int op;
  A lo, hi;         // clip bounds
  A in_min, in_span, out_span, out_min;
};

// the unary math processes (math.py:212-222, 305-315): np.abs, np.sign, np.sqrt, np.floor, np.ceil, np.trunc --
// all exactly rounded / exact operations, so the results are numpy's bit for bit (signed zeros included:
// np.sign(-0.0) is +0.0, ceil(-0.5) is -0.0, sqrt(-0.0) is -0.0)
template <typename A> __device__ __forceinline__ A apply_unary(A x, int op) {
  switch (op) {
    case EOC_ELEM_ABSOLUTE: return fabs(x);
    case EOC_ELEM_SGN: return is_nan(x) ? x : (x > (A)0 ? (A)1 : (x < (A)0 ? (A)-1 : (A)0));
    case EOC_ELEM_SQRT: return sqrt_rn(x);
    case EOC_ELEM_FLOOR: return floor(x);
    case EOC_ELEM_CEIL: return ceil(x);
    default: return trunc(x);
  }
}

template <typename A> __device__ __forceinline__ A apply_elem(A x, const ElemParams<A>& E) {
  if (E.op >= EOC_ELEM_ABSOLUTE) return apply_unary<A>(x, E.op);
  A y = is_nan(x) ? x : min_(max_(x, E.lo), E.hi);
  if (E.op == EOC_ELEM_LINEAR_SCALE_RANGE)
    y = add_rn(mul_rn(div_rn(sub_rn(y, E.in_min), E.in_span), E.out_span), E.out_min);
  return y;
}

template <typename In, typename A, int VEC>
__global__ void __launch_bounds__(256) k_elementwise(const In* __restrict__ in, A* __restrict__ out, int64_t nvec,
                                                     const ElemParams<A> E) {
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += step) {
    const Vec<In, VEC> q = ld_vec<In, VEC>(in + v * VEC);
    A y[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) y[k] = apply_elem<A>((A)q.v[k], E);
    st_vec<A, VEC>(out + v * VEC, y);
  }
}

template <typename In>
int launch_elem_t(const void* in, int64_t n, const ElemParams<double>& D, void* out, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  ElemParams<A> E;
  E.op = D.op;
  E.lo = (A)D.lo;
  E.hi = (A)D.hi;
  E.in_min = (A)D.in_min;
  E.in_span = (A)D.in_span;
  E.out_span = (A)D.out_span;
  E.out_min = (A)D.out_min;
  constexpr int VM = (sizeof(A) == 4) ? 4 : 2;
  const bool al = ((uintptr_t)in % (sizeof(In) * VM) == 0) && ((uintptr_t)out % (sizeof(A) * VM) == 0);
  const int64_t cap = (int64_t)sm_count() * 32;
  if (al) {
    const int64_t nvec = n / VM;
    if (nvec > 0) {
      int64_t blocks = (nvec + 255) / 256;
      if (blocks > cap) blocks = cap;
      k_elementwise<In, A, VM><<<(unsigned)blocks, 256, 0, s>>>((const In*)in, (A*)out, nvec, E);
      EOC_LAUNCH_CHECK();
    }
    const int64_t done = nvec * VM;
    if (done < n) {
      k_elementwise<In, A, 1><<<1, 32, 0, s>>>((const In*)in + done, (A*)out + done, n - done, E);
      EOC_LAUNCH_CHECK();
    }
    return EOC_OK;
  }
  int64_t blocks = (n + 255) / 256;
  if (blocks > cap) blocks = cap;
  k_elementwise<In, A, 1><<<(unsigned)blocks, 256, 0, s>>>((const In*)in, (A*)out, n, E);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace

int launch_elementwise(const void* in, int dtype, int64_t n, int op, const double* params, void* out,
                       cudaStream_t s) {
  if (n == 0) return EOC_OK;
  if (!in || !out || !params) EOC_FAIL(EOC_ERR_INVALID, "eoc_elementwise: NULL pointer");
  ElemParams<double> D;
  D.op = op;
  const double inf = __builtin_inf();
  if (op == EOC_ELEM_CLIP) {
    D.lo = params[0];
    D.hi = params[1];
    D.in_min = 0.0;
    D.in_span = D.out_span = 1.0;
    D.out_min = 0.0;
  } else if (op == EOC_ELEM_LINEAR_SCALE_RANGE) {
    const double in_min = params[0], in_max = params[1], out_min = params[2], out_max = params[3];
    D.lo = in_max < in_min ? in_max : in_min;  // math.py:293-296: clip(x, inputMax, inputMin) when swapped
    D.hi = in_max < in_min ? in_min : in_max;
    D.in_min = in_min;
    D.in_span = in_max - in_min;
    D.out_span = out_max - out_min;
    D.out_min = out_min;
  } else if (op >= EOC_ELEM_ABSOLUTE && op <= EOC_ELEM_TRUNC) {
    D.lo = D.hi = D.in_min = D.out_min = 0.0;
    D.in_span = D.out_span = 1.0;
  } else {
    EOC_FAIL(EOC_ERR_INVALID, "unknown elementwise op %d", op);
  }
  (void)inf;
  switch (dtype) {
    case EOC_F32: return launch_elem_t<float>(in, n, D, out, s);
    case EOC_F64: return launch_elem_t<double>(in, n, D, out, s);
    case EOC_U8: return launch_elem_t<uint8_t>(in, n, D, out, s);
    case EOC_I16: return launch_elem_t<int16_t>(in, n, D, out, s);
    case EOC_U16: return launch_elem_t<uint16_t>(in, n, D, out, s);
    case EOC_I32: return launch_elem_t<int32_t>(in, n, D, out, s);
  }
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", dtype);
}

}  // namespace eoc
