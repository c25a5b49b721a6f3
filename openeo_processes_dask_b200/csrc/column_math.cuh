// Per-column arithmetic of the reducers, shared by every kernel.
//
// Everything here is __host__ __device__ so tests/hostmodel can compile the very same code with
// g++ and compare it bit-for-bit with numpy on the CPU box before any GPU time is spent.  The
// arithmetic mirrors what numpy does for a reduction over a NON-contiguous axis of a C-ordered
// array (the (t,b,y,x) cube reduced over t):
//   np.nansum / np.nanmean   sequential adds in r order, NaN replaced by 0
//                            (numpy/lib/_nanfunctions_impl.py:70 _replace_nan, :1031-1054 nanmean)
//   np.nanvar(ddof)          two passes: avg = sum/n; sum of (x-avg)^2 with NaN slots zeroed; / (n-ddof)
//                            (numpy/lib/_nanfunctions_impl.py:1850-1892)
//   np.nanmedian             sort with NaN masked to +inf, (lo+hi)/2   (numpy/ma/extras.py:814-890)
//   np.nanquantile(linear)   virtual index (n-1)q, two-sided lerp in float64
//                            (numpy/lib/_function_base_impl.py:4657-4678, 4753-4785)
// No operation here may be contracted into an FMA: explicit *_rn intrinsics on the device,
// -ffp-contract=off on the host.
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define EOC_HD __host__ __device__ __forceinline__
#else
#define EOC_HD inline
#endif

namespace eoc {

// ---- correctly rounded, never-contracted primitives -------------------------------------------
EOC_HD float add_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fadd_rn(a, b);
#else
  return a + b;
#endif
}
EOC_HD double add_rn(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dadd_rn(a, b);
#else
  return a + b;
#endif
}
EOC_HD float sub_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fsub_rn(a, b);
#else
  return a - b;
#endif
}
EOC_HD double sub_rn(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dsub_rn(a, b);
#else
  return a - b;
#endif
}
EOC_HD float mul_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fmul_rn(a, b);
#else
  return a * b;
#endif
}
EOC_HD double mul_rn(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dmul_rn(a, b);
#else
  return a * b;
#endif
}
EOC_HD float div_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fdiv_rn(a, b);
#else
  return a / b;
#endif
}
EOC_HD double div_rn(double a, double b) {
#ifdef __CUDA_ARCH__
  return __ddiv_rn(a, b);
#else
  return a / b;
#endif
}
EOC_HD float sqrt_rn(float a) {
#ifdef __CUDA_ARCH__
  return __fsqrt_rn(a);
#else
  return sqrtf(a);
#endif
}
EOC_HD double sqrt_rn(double a) {
#ifdef __CUDA_ARCH__
  return __dsqrt_rn(a);
#else
  return sqrt(a);
#endif
}

template <typename A> struct Lim;
template <> struct Lim<float> {
  static EOC_HD float inf() { return __builtin_huge_valf(); }
  static EOC_HD float nan() { return __builtin_nanf(""); }
  static EOC_HD float maxv() { return 3.402823466e+38f; }
};
template <> struct Lim<double> {
  static EOC_HD double inf() { return __builtin_huge_val(); }
  static EOC_HD double nan() { return __builtin_nan(""); }
  static EOC_HD double maxv() { return 1.7976931348623157e+308; }
};

template <typename A> EOC_HD bool is_nan(A x) { return x != x; }
EOC_HD float abs_(float x) { return fabsf(x); }
EOC_HD double abs_(double x) { return fabs(x); }
// fmin/fmax with "NaN loses" semantics (np.fmin / np.fmax; CUDA fminf/fmaxf behave the same)
EOC_HD float min_(float a, float b) { return fminf(a, b); }
EOC_HD float max_(float a, float b) { return fmaxf(a, b); }
EOC_HD double min_(double a, double b) { return fmin(a, b); }
EOC_HD double max_(double a, double b) { return fmax(a, b); }

// ---- pass-1 statistics of one column ---------------------------------------------------------
// Accumulator groups (compile-time, so an HBM-bound kernel does not pay for unused reducers)
constexpr uint32_t G_MINMAX = 1u;  // min, max
constexpr uint32_t G_RARE = 2u;    // product + finite count
constexpr uint32_t G_COUNT = 4u;   // finite count only (the `count` process, e.g. aggregate_temporal(count))
constexpr uint32_t G_FIRSTLAST = 8u;  // first, last
constexpr uint32_t G_LOGIC = 16u;     // any, all
constexpr uint32_t G_ORDER = G_FIRSTLAST | G_LOGIC;
constexpr uint32_t G_NOSUM = 32u;     // min / max alone with ignore_nodata: no running sum, no not-NaN count

// NOTE on the NaN tests in sum_step / rare_step / pad_step: each spells "x is NaN" differently (against 0,
// against 1, against itself).  Written identically, ptxas merges the tests of one sample across the
// stages of a register-resident column (fused selection kernel) and keeps ~100 predicates alive as
// registers in between: that spilled the whole column (392 bytes/thread) until the spellings differed.
// if x is not NaN: sum += x, n += 1.  numpy adds 0 for a NaN slot instead of skipping it, which is the
// same thing: np.add.reduce starts from the identity +0.0 (np.sum of an all -0.0 column is +0.0, probed
// with numpy 2.3.5), so the running total is never -0.0 and adding +0 cannot change it.  Skipping makes
// the step 3 instructions on the device (FSETP, predicated FADD, predicated IADD) with a single one on the
// ALU pipe besides the count -- the pipe the selection networks saturate.
EOC_HD void sum_step(float x, float& sum, int& n) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred p;\n\t"
      "setp.num.f32 p, %2, 0f00000000;\n\t"
      "@p add.rn.f32 %0, %0, %2;\n\t"
      "@p add.s32 %1, %1, 1;\n\t}"
      : "+f"(sum), "+r"(n)
      : "f"(x));
#else
  if (!is_nan(x)) {
    sum = add_rn(sum, x);
    n += 1;
  }
#endif
}
EOC_HD void sum_step(double x, double& sum, int& n) {
  if (!is_nan(x)) {
    sum = add_rn(sum, x);
    n += 1;
  }
}
// sum_step with the count kept as a float (exact below 2^24; used for register-resident columns of at
// most 128 samples): both updates then issue on the FMA pipe and only the FSETP is left on the ALU pipe
EOC_HD void sum_step_fcount(float x, float& sum, float& cnt) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred p;\n\t"
      "setp.num.f32 p, %2, 0f00000000;\n\t"
      "@p add.rn.f32 %0, %0, %2;\n\t"
      "@p add.rn.f32 %1, %1, 0f3F800000;\n\t}"
      : "+f"(sum), "+f"(cnt)
      : "f"(x));
#else
  if (!is_nan(x)) {
    sum = add_rn(sum, x);
    cnt += 1.0f;
  }
#endif
}
EOC_HD void sum_step_fcount(double x, double& sum, double& cnt) {
  if (!is_nan(x)) {
    sum = add_rn(sum, x);
    cnt += 1.0;
  }
}
// prod *= (x is NaN ? 1 : x); nfin += (|x| <= FLT_MAX ? 1 : 0)  (false for NaN and +-inf).  Spelled out on
// the device for the same reason as sum_step: written in C++ over a register-resident column the compiler
// turns the 100 counts into a tree of temporaries and spills the column.
EOC_HD void rare_step(float x, float& prod, int& nfin) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred p, q;\n\t.reg .f32 a;\n\t"
      "setp.num.f32 p, %2, 0f3F800000;\n\t"
      "@p mul.rn.f32 %0, %0, %2;\n\t"
      "abs.f32 a, %2;\n\t"
      "setp.le.f32 q, a, 0f7F7FFFFF;\n\t"
      "@q add.s32 %1, %1, 1;\n\t}"
      : "+f"(prod), "+r"(nfin)
      : "f"(x));
#else
  if (!is_nan(x)) prod = mul_rn(prod, x);
  nfin += (fabsf(x) <= 3.402823466e+38f) ? 1 : 0;
#endif
}
// nfin += (|x| <= FLT_MAX ? 1 : 0) alone
EOC_HD void fin_step(float x, int& nfin) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred q;\n\t.reg .f32 a;\n\t"
      "abs.f32 a, %1;\n\t"
      "setp.le.f32 q, a, 0f7F7FFFFF;\n\t"
      "@q add.s32 %0, %0, 1;\n\t}"
      : "+r"(nfin)
      : "f"(x));
#else
  nfin += (fabsf(x) <= 3.402823466e+38f) ? 1 : 0;
#endif
}
EOC_HD void fin_step(double x, int& nfin) { nfin += (fabs(x) <= 1.7976931348623157e+308) ? 1 : 0; }
EOC_HD void rare_step(double x, double& prod, int& nfin) {
  if (!is_nan(x)) prod = mul_rn(prod, x);
  nfin += (fabs(x) <= 1.7976931348623157e+308) ? 1 : 0;
}

template <typename A, uint32_t G>
struct ColStats {
  A sum;     // nansum, starts at +0.0 like np.add.reduce (identity), so an all -0.0 column sums to +0.0
  A prod;
  A mn, mx;  // start at NaN: fmin(NaN, x) == x
  int n;     // not-NaN samples
  int nfin;  // not-NaN and not-inf samples (the `count` process)
  // G_FIRSTLAST: first / last not-NaN sample, sample 0 / sample R-1 as they are; G_LOGIC: any non-zero, no zero
  A first, last, raw0, rawl;
  bool seen, any_, all_;
  EOC_HD void init() {
    sum = (A)0;
    prod = (A)1;
    mn = Lim<A>::nan();
    mx = Lim<A>::nan();
    n = 0;
    nfin = 0;
    first = last = raw0 = rawl = Lim<A>::nan();
    seen = false;
    any_ = false;
    all_ = true;
  }
  EOC_HD void add(A x) {
    if (!(G & G_NOSUM)) sum_step(x, sum, n);
    if (G & G_MINMAX) {
      mn = min_(mn, x);
      mx = max_(mx, x);
    }
    if (G & G_RARE)
      rare_step(x, prod, nfin);
    else if (G & G_COUNT)
      fin_step(x, nfin);
    if (G & G_FIRSTLAST) {
      first = is_nan(first) ? x : first;  // stays NaN until the first valid sample arrives
      last = is_nan(x) ? last : x;
      raw0 = seen ? raw0 : x;
      rawl = x;
      seen = true;
    }
    if (G & G_LOGIC) {
      any_ = any_ || (x < (A)0 || x > (A)0);  // np.any(np.nan_to_num(x)): NaN -> 0, +-inf stay non-zero
      all_ = all_ && !(x == (A)0);            // np.all(x): NaN is truthy
    }
  }
};

// avg exactly as numpy's _divide_by_count: float64 division of (sum, count) rounded to A.  For
// A=float that double rounding is innocuous (53 >= 2*24+2), so a correctly rounded float division
// by the (exactly representable) count gives the same bits.
template <typename A> EOC_HD A mean_of(A sum, int n) { return div_rn(sum, (A)n); }

// pass-2 term of the variance: (x - avg)^2 with NaN slots contributing 0
template <typename A> EOC_HD A sqdev(A x, A avg) {
  const A d = is_nan(x) ? (A)0 : sub_rn(x, avg);
  return mul_rn(d, d);
}
// The same term in one ALU-pipe instruction less, valid when `sqdev_fast_ok(avg, n)`: with a finite
// avg, x - avg is NaN exactly when x is NaN, the square of anything else is >= +0, and
// fmax(NaN, 0) == 0 puts the NaN slots at 0 (an all-NaN column, avg == NaN and n == 0, likewise sums
// to 0).  A non-finite avg with valid samples (a +-inf sample, or an overflowed sum) must take
// sqdev(): there numpy propagates inf - inf = NaN from the valid slots.
template <typename A> EOC_HD A sqdev_fast(A x, A avg) {
  const A d = sub_rn(x, avg);
  return max_(mul_rn(d, d), (A)0);
}
template <typename A> EOC_HD bool sqdev_fast_ok(A avg, int n) { return n == 0 || abs_(avg) <= Lim<A>::maxv(); }

// variance from the pass-2 sum; dof <= 0 -> NaN (numpy/lib/_nanfunctions_impl.py:1882-1891)
template <typename A> EOC_HD A variance_of(A sumsq, int n, int ddof) {
  const int dof = n - ddof;
  return dof <= 0 ? Lim<A>::nan() : div_rn(sumsq, (A)dof);
}

// every value reducer of one column, from the pass-1 statistics and the pass-2 sum
template <typename A> struct ColumnResult {
  A sum, mean, mn, mx, prod, var, sd, count_all;
  long long count, not_nan;
  A first, last, any_, all_;
};

// ignore_nodata == 0 and a NaN in the column -> NaN for every value reducer (np.mean/np.min/...
// propagate NaN); `empty_nan` marks a missing group of a segmented reduce (xarray re-indexes it to NaN)
template <typename A, uint32_t G>
EOC_HD ColumnResult<A> finish_column(const ColStats<A, G>& st, int len, A sumsq, bool ignore_nodata,
                                     int ddof, bool empty_nan) {
  // (G_NOSUM is only selected with ignore_nodata: the not-NaN count is not kept there)
  const bool poisoned = (!(G & G_NOSUM) && !ignore_nodata && st.n < len) || empty_nan;
  const A nanv = Lim<A>::nan();
  const A total = st.sum;
  ColumnResult<A> r;
  r.sum = poisoned ? nanv : total;
  r.mean = poisoned ? nanv : mean_of(total, st.n);
  r.mn = poisoned ? nanv : st.mn;
  r.mx = poisoned ? nanv : st.mx;
  r.prod = poisoned ? nanv : st.prod;
  r.var = poisoned ? nanv : variance_of(sumsq, st.n, ddof);
  r.sd = sqrt_rn(r.var);
  r.count = st.nfin;
  r.not_nan = st.n;
  r.count_all = (A)len;
  if (G & G_ORDER) {
    // first / last (arrays.py:461-495): ignore_nodata = False takes sample 0 / R-1 whatever it holds
    r.first = empty_nan ? nanv : (ignore_nodata ? st.first : st.raw0);
    r.last = empty_nan ? nanv : (ignore_nodata ? st.last : st.rawl);
    // _any / _all (logic.py:58-87): with ignore_nodata = False a NaN in the column makes the result NaN
    // unless the valid samples already decide it (any: a non-zero sample; all: a zero sample)
    const bool has_nan = st.n < len;
    r.any_ = (empty_nan || (!ignore_nodata && has_nan && !st.any_)) ? nanv : (st.any_ ? (A)1 : (A)0);
    r.all_ = (empty_nan || (!ignore_nodata && has_nan && st.all_)) ? nanv : (st.all_ ? (A)1 : (A)0);
  } else {
    r.first = r.last = r.any_ = r.all_ = nanv;
  }
  return r;
}

// ---- order statistics --------------------------------------------------------------------------
// max that lets NaN win (PTX max.NaN -> FMNMX.NAN): with min_ (NaN loses) it makes a comparator that
// treats NaN as larger than everything, +inf included
EOC_HD float max_nan_(float a, float b) {
#ifdef __CUDA_ARCH__
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
#else
  return (a != a || b != b) ? __builtin_nanf("") : fmaxf(a, b);
#endif
}
EOC_HD double max_nan_(double a, double b) { return (a != a || b != b) ? __builtin_nan("") : fmax(a, b); }

// compare-exchange of the networks.  NANHI = false: plain min/max (inputs hold no NaN any more).
// NANHI = true: NaN sorts above +inf natively, so a NaN-padded column needs NO canonicalisation pass
// before the network (saves one FMNMX and one NaN test per sample on the ALU pipe, the bottleneck of the
// quantile kernel) and real +inf samples stay distinguishable from the pads.
template <bool NANHI, typename A> EOC_HD void cmpx(A& a, A& b) {
  const A lo = min_(a, b);
  if constexpr (NANHI)
    b = max_nan_(a, b);
  else
    b = max_(a, b);
  a = lo;
}
// Mixed-pipe form of the same comparator for float32 on the device (see cmpx_mix below for the idea):
// lo = min(a, b) on the ALU pipe, hi = bits(a) + bits(b) - bits(lo) by two IMADs on the FMA pipe.  With
// NANHI the columns must hold CANONICAL NaNs (0x7FFFFFFF, what FMNMX / FADD produce): min(NaN, NaN) returns
// the canonical NaN, so the integer identity only yields a NaN again if both operands were canonical --
// canonical_nans() below does that with one FADD per sample.
#ifndef EOC_MIXQ_NUM
#define EOC_MIXQ_NUM 3
#endif
#ifndef EOC_MIXQ_DEN
#define EOC_MIXQ_DEN 5
#endif
template <bool NANHI, bool MIX, typename A> EOC_HD void cmpx_k(A& a, A& b, int one) {
#ifdef __CUDA_ARCH__
  if constexpr (MIX && sizeof(A) == 4) {
    const float lo = fminf(a, b);
    int s, h;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(s) : "r"(__float_as_int(a)), "r"(one), "r"(__float_as_int(b)));
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(h) : "r"(__float_as_int(lo)), "r"(-one), "r"(s));
    a = lo;
    b = __int_as_float(h);
    return;
  }
#endif
  (void)one;
  cmpx<NANHI, A>(a, b);
}
#define EOC_CS(a, b, k) cmpx_k<NANHI, (((k) % EOC_MIXQ_DEN) < EOC_MIXQ_NUM), A>((a), (b), one)
// "raw" networks (sortraw_N, k_select_quad): the first comparator a wire meets is pure -- min (NaN loses) and
// max.NaN hand back a sample or the CANONICAL NaN whatever payload came in -- so the later, possibly mixed-pipe
// comparators see canonical NaNs without a canonicalisation pass.  Their own mixing share (the quad kernel has
// fewer other ALU-pipe instructions per comparator than the two-run kernel, so its optimum differs).
#ifndef EOC_MIXR_NUM
#define EOC_MIXR_NUM 3  // measured on B200 (median t = 100, GB/s): 3/5 4865, 3/4 5064, 4/5 4982, 5/6 4897, 1/1 4531
#endif
#ifndef EOC_MIXR_DEN
#define EOC_MIXR_DEN 4
#endif
#define EOC_CSF(a, b) cmpx<true, A>((a), (b))
#define EOC_CSQ(a, b, k) cmpx_k<true, (((k) % EOC_MIXR_DEN) < EOC_MIXR_NUM), A>((a), (b), one)

// Comparator of the median cones (inputs hold no NaN: the pads are +-inf).  The selection kernels are bound
// by the ALU pipe (FMNMX: one warp instruction per 2 cycles per SM sub-partition) while the FMA pipe idles,
// so a share of the comparators whose BOTH outputs are live is spelled with one ALU-pipe and two FMA-pipe
// instructions instead of two ALU-pipe ones: lo = min(a, b) is bit for bit one of the operands (-0 < +0 in
// FMNMX, no flush to zero), hence the other operand is bits(a) + bits(b) - bits(lo) in wrapping integer
// arithmetic -- two IMADs whose multipliers (+1, -1) are RUN-TIME values so that the assembler cannot turn
// them back into an IADD3 on the ALU pipe.  EOC_MIX_NUM of every EOC_MIX_DEN full comparators take this form
// (tuned on B200; the instruction issue rate is the next limit).
#ifndef EOC_MIX_NUM
#define EOC_MIX_NUM 1
#endif
#ifndef EOC_MIX_DEN
#define EOC_MIX_DEN 2
#endif
template <bool MIX> EOC_HD void cmpx_mix(float& a, float& b, int one) {
#ifdef __CUDA_ARCH__
  if constexpr (MIX) {
    const float lo = fminf(a, b);
    int s, h;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(s) : "r"(__float_as_int(a)), "r"(one), "r"(__float_as_int(b)));
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(h) : "r"(__float_as_int(lo)), "r"(-one), "r"(s));
    a = lo;
    b = __int_as_float(h);
    return;
  }
#endif
  (void)one;
  const float lo = min_(a, b);
  b = max_(a, b);
  a = lo;
}
template <bool MIX> EOC_HD void cmpx_mix(double& a, double& b, int) {
  const double lo = min_(a, b);
  b = max_(a, b);
  a = lo;
}
#define EOC_CSX(a, b, k, one) cmpx_mix<(((k) % EOC_MIX_DEN) < EOC_MIX_NUM)>((a), (b), (one))
#define EOC_MIN(a, b) (a) = min_((a), (b))
#define EOC_MAX(a, b) (b) = max_((a), (b))

// lockstep point inside a network (see select_kernel.cuh): a CTA barrier on the device when S
#if defined(__CUDA_ARCH__)
#define EOC_NET_SYNC(S)               \
  do {                                \
    if constexpr (S) __syncthreads(); \
  } while (0)
#else
#define EOC_NET_SYNC(S) \
  do {                  \
  } while (0)
#endif

// candidate k of the generated networks (one per >= 64 comparators; measured on B200: 1, 2 or 3 times denser or sparser changes nothing) is a barrier when k % EOC_SYNC_MOD == 0
#ifndef EOC_SYNC_MOD
#define EOC_SYNC_MOD 1
#endif
#define EOC_NET_SYNC_K(S, k) EOC_NET_SYNC((S) && ((k) % EOC_SYNC_MOD) == (EOC_SYNC_MOD - 1))

#include "generated/sort_networks.inc"

// median from the two middle order statistics, as numpy/ma/extras.py:884-890: s = lo + hi; s / 2
// (for an odd count numpy duplicates the middle element, so lo == hi and (h+h)/2 == h unless h+h
// overflows -- reproduced on purpose).
template <typename A> EOC_HD A median_of_pair(A lo, A hi) { return mul_rn(add_rn(lo, hi), (A)0.5); }

// nan-median of N register-resident samples (NaN = nodata or unused slot; R = slots in use).
// NaNs are replaced ALTERNATELY by -inf / +inf: with P pads in total ceil(P/2) sort below and
// floor(P/2) above the n valid samples, so their median sits at the fixed wires N/2-1, N/2 (n even)
// or N/2 (n odd), and the compiler drops every comparator outside the cone of those two wires.
// One step of the pad substitution: v <- pad if v is NaN, then flip the pad's sign and count it.
// The selection kernel is bound by the ALU pipe (FMNMX), so on the device the flip and the count are
// issued as predicated FADDs on the FMA pipe, and so is the substitution itself (`v <- pad + zero` under
// the predicate instead of a select): ONE ALU-pipe instruction per sample (FSETP).  `zero` is a RUNTIME
// 0.0f (a kernel parameter): with a literal the assembler folds either form back into an FSEL / MOV on
// the ALU pipe.
EOC_HD void pad_step(float& v, float& pad, float& nnan, float zero) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred p;\n\t"
      "setp.nan.f32 p, %0, %0;\n\t"
      "@p add.f32 %0, %1, %3;\n\t"
      "@p sub.f32 %1, %3, %1;\n\t"
      "@p add.f32 %2, %2, 0f3F800000;\n\t}"
      : "+f"(v), "+f"(pad), "+f"(nnan)
      : "f"(zero));
#else
  (void)zero;
  const bool nan = is_nan(v);
  v = nan ? pad : v;
  pad = nan ? -pad : pad;
  nnan += nan ? 1.0f : 0.0f;
#endif
}
EOC_HD void pad_step(double& v, double& pad, double& nnan, double) {
  const bool nan = is_nan(v);
  v = nan ? pad : v;
  pad = nan ? -pad : pad;
  nnan += nan ? 1.0 : 0.0;
}

// NaN -> +inf, full sort; returns n = number of valid samples, which then occupy v[0..n)
// (fmin(NaN, +inf) == +inf does the substitution in one FMNMX; the count is a predicated FADD on
// the FMA pipe on the device, for the same reason as in pad_step)
EOC_HD void count_nan(float v, float& nnan) {
#ifdef __CUDA_ARCH__
  asm("{\n\t.reg .pred p;\n\t"
      "setp.nan.f32 p, %1, %1;\n\t"
      "@p add.f32 %0, %0, 0f3F800000;\n\t}"
      : "+f"(nnan)
      : "f"(v));
#else
  nnan += is_nan(v) ? 1.0f : 0.0f;
#endif
}
EOC_HD void count_nan(double v, double& nnan) { nnan += is_nan(v) ? 1.0 : 0.0; }

// +1 the assembler cannot see through (zero is the kernel's run-time 0.0)
EOC_HD int one_from_zero(float zero) {
#ifdef __CUDA_ARCH__
  return __float_as_int(zero) + 1;
#else
  (void)zero;
  return 1;
#endif
}
EOC_HD int one_from_zero(double) { return 1; }

template <int N, typename A> EOC_HD A nanmedian_network(A* v, int R, bool ignore_nodata, A zero = (A)0) {
  static_assert(N % 2 == 0, "network size must be even");
  // (Measured on B200: dropping the count from pad_step -- n's parity is the pad's final sign, n == 0 leaves
  // -inf / +inf on the median wires -- made the kernel 10 % SLOWER, 4382 -> 3964 GB/s; the predicated count
  // rides on the FMA pipe for free and the shorter dependency chain did not pay.)
  A pad = -Lim<A>::inf();
  A nnan = (A)0;  // exact: at most N <= 128
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int j = 0; j < N; ++j) pad_step(v[j], pad, nnan, zero);
  const int n = N - (int)nnan;
  mediannet<N, A>(v, one_from_zero(zero));
  const A lo = v[N / 2 - 1], hi = v[N / 2];
  A m = (n & 1) ? median_of_pair<A>(hi, hi) : median_of_pair<A>(lo, hi);
  if (n == 0 || (!ignore_nodata && n < R)) m = Lim<A>::nan();
  return m;
}

// full ascending sort with NaN (nodata and unused slots) last; the valid samples, real infinities included,
// end up in v[0..n) with n = index of the first NaN (the caller finds it by bisection over the parked column)
// `zero` = the kernel's run-time 0.0: x - 0.0 == x bit for bit for every number (-0.0 included) and turns any
// NaN into the canonical one the mixed-pipe comparators need; one FADD on the FMA pipe per sample.
template <int N, typename A, bool SYNC = false> EOC_HD void sort_nan_last(A* v, A zero = (A)0) {
#if defined(__CUDA_ARCH__)
  if constexpr (sizeof(A) == 4 && EOC_MIXQ_NUM > 0) {
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = sub_rn(v[j], zero);
  }
#endif
  sortnet<N, A, SYNC, true>(v, one_from_zero(zero));
}
template <typename Fetch> EOC_HD int first_nan_index(int N, Fetch fetch) {
  int lo = 0, hi = N;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const auto x = fetch(mid);
    if (x != x)
      hi = mid;
    else
      lo = mid + 1;
  }
  return lo;
}

template <int N, typename A, bool SYNC = false> EOC_HD int sort_valid_first(A* v, int one = 1) {
  A nnan = (A)0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int j = 0; j < N; ++j) {
    count_nan(v[j], nnan);
    v[j] = min_(v[j], Lim<A>::inf());
  }
  sortnet<N, A, SYNC>(v, one);
  return N - (int)nnan;
}

// Linear-method quantile from the neighbours of the virtual index.  n = valid samples (> 0).
// `fetch(k)` returns the k-th smallest valid sample (0-based).
//   virtual = (n-1)*q; prev = floor(virtual), next = prev+1; virtual >= n-1 -> both are the last
//   element and gamma = virtual - (-1)  (numpy indexes with -1 there; harmless because the
//   neighbours are equal unless they are infinite, which numpy turns into NaN too); virtual < 0 -> 0.
// the two ranks numpy's "linear" method interpolates between, and the weight gamma
EOC_HD void quantile_ranks(int n, double q, int& ip, int& in_, double& gamma) {
  const double virt = mul_rn((double)(n - 1), q);
  double prevd = floor(virt);
  if (virt >= (double)(n - 1)) {
    ip = in_ = n - 1;
    prevd = -1.0;
  } else if (virt < 0.0) {
    ip = in_ = 0;
    prevd = 0.0;
  } else {
    ip = (int)prevd;
    in_ = ip + 1;
  }
  if (is_nan(virt)) {  // q is NaN: numpy takes the last element for both
    ip = in_ = n - 1;
    prevd = -1.0;
  }
  gamma = sub_rn(virt, prevd);
}
// numpy's two-sided _lerp (numpy/lib/_function_base_impl.py:4657-4678) between samples a <= b
// wrap16: the samples are int16 digital numbers carried in float32; numpy forms b - a in int16, which WRAPS
// when the neighbours are more than 32767 apart -- reproduced, since parity is with what the reference returns.
template <typename A> EOC_HD double quantile_lerp(A a, A b, double gamma, bool wrap16 = false) {
  double diff = (double)sub_rn(b, a);  // subtract(b, a) happens in the sample dtype
  if (wrap16) diff = (double)(short)(int)diff;
  double r = add_rn((double)a, mul_rn(diff, gamma));
  if (gamma >= 0.5) r = sub_rn((double)b, mul_rn(diff, sub_rn(1.0, gamma)));
  return r;
}
template <typename A, typename Fetch> EOC_HD double quantile_linear(int n, double q, Fetch fetch, bool wrap16 = false) {
  int ip, in_;
  double gamma;
  quantile_ranks(n, q, ip, in_, gamma);
  const A a = fetch(ip);
  const A b = fetch(in_);
  return quantile_lerp<A>(a, b, gamma, wrap16);
}

// ---- split selection: two sorted runs and a rank search (round 2) --------------------------------------------
// A column of R <= 2*NH samples is sorted as TWO runs of NH (NaN last in each, cmpx<NANHI>): 2 * S(NH)
// comparators instead of S(2*NH) -- 806 instead of 1104 for R = 100 -- and no pad substitution at all.  The
// order statistics numpy needs are then ranks of the UNION of the two runs, found by the classic partition
// search: the k+1 smallest elements of the union are the first i of run a and the first k+1-i of run b, where i
// is the smallest index in [max(0, k+1-nb), min(k+1, na)] with  i == upper bound  or  !(a[i] < b[k-i]).
// Fixed trip count (ceil(log2(NH+1)) halvings: every thread of a warp runs the same instruction stream), data
// dependent addresses only (the runs are parked in shared memory on the device).
//   select_two_runs(...)  ->  x_k and x_{k+1} of the union (0-based; x_{k+1} = x_k when k is the last rank)
constexpr int split_iters(int nh) {
  int it = 0;
  while ((1 << it) < nh + 1) ++it;
  return it;
}
template <typename A> EOC_HD bool less_(A x, A y) { return x < y; }
template <int ITERS, typename A, typename FA, typename FB>
EOC_HD void select_two_runs(FA a, int na, FB b, int nb, int k, A& xk, A& xk1) {
  int lo = (k + 1 - nb) > 0 ? (k + 1 - nb) : 0;
  int hi = (k + 1) < na ? (k + 1) : na;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int it = 0; it < ITERS; ++it) {
    const int mid = (lo + hi) >> 1;
    // mid < hi keeps both fetches inside the runs: mid < na, and 0 <= k - mid < nb by the bounds on lo / hi
    const int ma = mid < hi ? mid : (hi > 0 ? hi - 1 : 0);
    const int mb = (k - ma) < 0 ? 0 : ((k - ma) < nb ? (k - ma) : (nb > 0 ? nb - 1 : 0));
    const bool more = (mid < hi) && (nb > 0) && less_(a(ma), b(mb));
    lo = more ? mid + 1 : lo;
    hi = more ? hi : mid;
  }
  const int i = lo, j = k + 1 - lo;
  const A ninf = -Lim<A>::inf(), pinf = Lim<A>::inf();
  const A ap = i > 0 ? a(i - 1) : ninf;
  const A bp = j > 0 ? b(j - 1) : ninf;
  xk = bp > ap ? bp : ap;
  const A an = i < na ? a(i) : pinf;
  const A bn = j < nb ? b(j) : pinf;
  const A nx = bn < an ? bn : an;
  xk1 = (k + 1 < na + nb) ? nx : xk;
}

// median / linear quantile of the union of two sorted runs holding na + nb valid samples (> 0)
template <int ITERS, typename A, typename FA, typename FB> EOC_HD A median_two_runs(FA a, int na, FB b, int nb) {
  const int n = na + nb;
  A lo, hi;
  select_two_runs<ITERS, A>(a, na, b, nb, (n - 1) >> 1, lo, hi);
  return (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
}
template <int ITERS, typename A, typename FA, typename FB>
EOC_HD double quantile_two_runs(FA a, int na, FB b, int nb, double q, bool wrap16 = false) {
  int ip, in_;
  double gamma;
  quantile_ranks(na + nb, q, ip, in_, gamma);
  A x0, x1;
  select_two_runs<ITERS, A>(a, na, b, nb, ip, x0, x1);
  return quantile_lerp<A>(x0, in_ == ip ? x0 : x1, gamma, wrap16);
}

// K sorted runs (K = 3, 4: columns of 129 .. 512 samples as runs of 100 or 128).  `f(r, idx)` returns element idx
// of run r, n[r] its number of valid samples.  Rank k of the union by elimination: with `rem` elements still to
// skip and d = (rem - 1) / K + 1, look at element off[r] + d - 1 of every run that has one; the run holding the
// SMALLEST of those candidates can drop its next d elements -- at most K (d - 1) < rem remaining elements are
// smaller than that candidate, so all d are among the `rem` smallest.  rem shrinks by a factor (1 - 1/K) per
// step (17 steps for k = 182, K = 4); the step sequence depends on k only, not on the data.  No dynamic
// indexing of off[] (it would go to local memory): the update is a predicated add per run.
template <int K, typename A, typename F>
EOC_HD void select_k_runs(F f, const int* n, int k, A& xk, A& xk1) {
  int off[K];
  int total = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int r = 0; r < K; ++r) {
    off[r] = 0;
    total += n[r];
  }
  int rem = k;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  while (rem > 0) {
    const int d = (rem - 1) / K + 1;
    int best = -1;
    A bestv = (A)0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < K; ++r) {
      const int idx = off[r] + d - 1;
      const bool has = idx < n[r];
      const A v = f(r, has ? idx : 0);
      const bool take = has && (best < 0 || v < bestv);
      best = take ? r : best;
      bestv = take ? v : bestv;
    }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < K; ++r) off[r] += (r == best) ? d : 0;
    rem -= d;
  }
  // heads: x_k is the smallest, x_{k+1} the smallest once that one is removed
  A h0 = (A)0, h1 = (A)0;
  int r0 = -1;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int r = 0; r < K; ++r) {
    const bool has = off[r] < n[r];
    const A v = f(r, has ? off[r] : 0);
    const bool take = has && (r0 < 0 || v < h0);
    r0 = take ? r : r0;
    h0 = take ? v : h0;
  }
  bool any1 = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int r = 0; r < K; ++r) {
    const int idx = off[r] + ((r == r0) ? 1 : 0);
    const bool has = idx < n[r];
    const A v = f(r, has ? idx : 0);
    const bool take = has && (!any1 || v < h1);
    any1 = any1 || has;
    h1 = take ? v : h1;
  }
  xk = h0;
  xk1 = (k + 1 < total && any1) ? h1 : h0;
}
// FOUR short runs with NaN sentinels (k_select_quad: plain median of 64 < R <= 128 samples as 4 runs of NH <= 32).
// Layout seen through fetch(pos): run r is rows [r * (NH + 1), r * (NH + 1) + NH) of the column, sorted ascending
// with NaN (nodata, unused slots) last, and row r * (NH + 1) + NH holds a NaN that is never overwritten -- so a
// candidate position clamped to the sentinel reads NaN, NaN never is the smallest candidate (min_: NaN loses), and
// the elimination of select_k_runs needs neither per-run counts nor bounds predicates: per step and run one add,
// one clamp, one address, one load.
template <int NH, typename F> EOC_HD int quad_valid_prefix(F fetch, int r) {
  // number of leading non-NaN elements of run r: branch-free bisection, steps 32 / 16 / 8 .. 1 (NH <= 64)
  const int base = r * (NH + 1);
  int pos = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int s = (NH > 64 ? 64 : (NH > 32 ? 32 : (NH > 16 ? 16 : 8))); s >= 1; s >>= 1) {
    const int t = pos + s;
    const int idx = (t - 1) < NH ? (t - 1) : NH;  // the sentinel reads NaN
    const auto x = fetch(base + idx);
    pos = (x == x) ? t : pos;
  }
  // NH = 32 (or 16): a full run of 2^m valid samples is one more than the steps can reach
  if ((NH & (NH - 1)) == 0) {
    const auto x = fetch(base + (pos < NH ? pos : NH));
    pos = (x == x) ? pos + 1 : pos;
  }
  return pos;
}
// min in the TOTAL order the device's FMNMX implements (NaN loses, -0.0 < +0.0, the result is one of the
// operands bit for bit); fminf on the host may hand back either zero
EOC_HD float min_ord_(float a, float b) {
#ifdef __CUDA_ARCH__
  return fminf(a, b);
#else
  if (a != a) return b;
  if (b != b) return a;
  if (a < b) return a;
  if (b < a) return b;
  return signbit(a) ? a : b;
#endif
}
EOC_HD double min_ord_(double a, double b) {
  if (a != a) return b;
  if (b != b) return a;
  if (a < b) return a;
  if (b < a) return b;
  return signbit(a) ? a : b;
}
EOC_HD bool same_bits_(float a, float b) {
#ifdef __CUDA_ARCH__
  return __float_as_int(a) == __float_as_int(b);
#else
  uint32_t x, y;
  __builtin_memcpy(&x, &a, 4);
  __builtin_memcpy(&y, &b, 4);
  return x == y;
#endif
}
EOC_HD bool same_bits_(double a, double b) {
  uint64_t x, y;
  __builtin_memcpy(&x, &a, 8);
  __builtin_memcpy(&y, &b, 8);
  return x == y;
}
// (the kernel runs the same steps on shared-memory addresses in hand-written PTX: select_quad_runs_dev)
template <int NH, typename A, typename F> EOC_HD void select_quad_runs(F fetch, int k, A& xk, A& xk1) {
  constexpr int S = NH + 1;
  int h0 = 0, h1 = S, h2 = 2 * S, h3 = 3 * S;  // heads (row positions) of the four runs
  int rem = k;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  while (rem > 0) {
    const int d = ((rem - 1) >> 2) + 1;
    const int c0 = (h0 + d - 1) < (NH) ? (h0 + d - 1) : (NH);
    const int c1 = (h1 + d - 1) < (S + NH) ? (h1 + d - 1) : (S + NH);
    const int c2 = (h2 + d - 1) < (2 * S + NH) ? (h2 + d - 1) : (2 * S + NH);
    const int c3 = (h3 + d - 1) < (3 * S + NH) ? (h3 + d - 1) : (3 * S + NH);
    const A v0 = fetch(c0), v1 = fetch(c1), v2 = fetch(c2), v3 = fetch(c3);
    const A m = min_ord_(min_ord_(v0, v1), min_ord_(v2, v3));
    // the winner: the first run whose candidate IS the minimum bit for bit (NaN is never the minimum)
    const bool p0 = same_bits_(v0, m);
    const bool p1 = !p0 && same_bits_(v1, m);
    const bool p2 = !p0 && !p1 && same_bits_(v2, m);
    const bool p3 = !p0 && !p1 && !p2;
    h0 += p0 ? d : 0;
    h1 += p1 ? d : 0;
    h2 += p2 ? d : 0;
    h3 += p3 ? d : 0;
    rem -= d;
  }
  // x_k = the smallest head; x_{k+1} = the smallest once that one is removed (NaN when nothing is left)
  const A v0 = fetch(h0), v1 = fetch(h1), v2 = fetch(h2), v3 = fetch(h3);
  const A m = min_ord_(min_ord_(v0, v1), min_ord_(v2, v3));
  const bool p0 = same_bits_(v0, m);
  const bool p1 = !p0 && same_bits_(v1, m);
  const bool p2 = !p0 && !p1 && same_bits_(v2, m);
  const bool p3 = !p0 && !p1 && !p2;
  const int hn = (p0 ? h0 : (p1 ? h1 : (p2 ? h2 : h3))) + 1;
  const A nx = fetch(hn);
  const A w0 = p0 ? nx : v0, w1 = p1 ? nx : v1, w2 = p2 ? nx : v2, w3 = p3 ? nx : v3;
  xk = m;
  xk1 = min_ord_(min_ord_(w0, w1), min_ord_(w2, w3));
}
// The same elimination over K runs with sentinels (k_select_multi: K = 8; the kernel runs it in PTX, MultiCursor8)
template <int NH, int K, typename A, typename F> EOC_HD void select_multi_runs(F fetch, int k, A& xk, A& xk1) {
  constexpr int S = NH + 1;
  int h[K];
  for (int r = 0; r < K; ++r) h[r] = r * S;
  int rem = k;
  while (rem > 0) {
    const int d = (rem - 1) / K + 1;
    A v[K];
    A m = Lim<A>::nan();
    for (int r = 0; r < K; ++r) {
      const int c = (h[r] + d - 1) < (r * S + NH) ? (h[r] + d - 1) : (r * S + NH);
      v[r] = fetch(c);
      m = min_ord_(m, v[r]);
    }
    bool taken = false;
    for (int r = 0; r < K; ++r) {
      const bool win = !taken && (r == K - 1 || same_bits_(v[r], m));
      h[r] += win ? d : 0;
      taken = taken || win;
    }
    rem -= d;
  }
  A v[K];
  A m = Lim<A>::nan();
  for (int r = 0; r < K; ++r) {
    v[r] = fetch(h[r]);
    m = min_ord_(m, v[r]);
  }
  bool taken = false;
  A nx = Lim<A>::nan();
  for (int r = 0; r < K; ++r) {
    const bool win = !taken && (r == K - 1 || same_bits_(v[r], m));
    if (win) {
      nx = fetch(h[r] + 1);
      v[r] = Lim<A>::nan();
    }
    taken = taken || win;
  }
  A m1 = nx;
  for (int r = 0; r < K; ++r) m1 = min_ord_(m1, v[r]);
  xk = m;
  xk1 = m1;
}
template <int NH, int K, typename A, typename F> EOC_HD A median_multi_runs(F fetch, int n) {
  A lo, hi;
  select_multi_runs<NH, K, A>(fetch, (n - 1) >> 1, lo, hi);
  return (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
}

// median of the union of the four runs, n = number of valid samples (> 0)
template <int NH, typename A, typename F> EOC_HD A median_quad_runs(F fetch, int n) {
  A lo, hi;
  select_quad_runs<NH, A>(fetch, (n - 1) >> 1, lo, hi);
  return (n & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
}

template <int K, typename A, typename F> EOC_HD A median_k_runs(F f, const int* n, int total) {
  A lo, hi;
  select_k_runs<K, A>(f, n, (total - 1) >> 1, lo, hi);
  return (total & 1) ? median_of_pair<A>(lo, lo) : median_of_pair<A>(lo, hi);
}
template <int K, typename A, typename F>
EOC_HD double quantile_k_runs(F f, const int* n, int total, double q, bool wrap16 = false) {
  int ip, in_;
  double gamma;
  quantile_ranks(total, q, ip, in_, gamma);
  A x0, x1;
  select_k_runs<K, A>(f, n, ip, x0, x1);
  return quantile_lerp<A>(x0, in_ == ip ? x0 : x1, gamma, wrap16);
}

}  // namespace eoc
