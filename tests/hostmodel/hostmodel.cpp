// Host build of the kernels' per-column arithmetic (TEST INFRASTRUCTURE).
//
// Compiles openeo_processes_dask_b200/csrc/column_math.cuh -- the very header the CUDA kernels
// include -- with g++ (-ffp-contract=off) and exposes the per-column routines through a C ABI so the
// CPU test-suite can compare them bit-for-bit with numpy before any GPU time is spent.  This is not
// a product path: nothing under openeo_processes_dask_b200/ loads it.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../openeo_processes_dask_b200/csrc/column_math.cuh"

using namespace eoc;

namespace {

template <typename A>
void reduce_columns(const A* data, int R, int inner, int ignore, int ddof, A* sum, A* mean, A* mn, A* mx,
                    A* prod, A* var, A* sd, long long* count) {
  for (int i = 0; i < inner; ++i) {
    ColStats<A, G_MINMAX | G_RARE> st;
    st.init();
    for (int r = 0; r < R; ++r) st.add(data[(size_t)r * inner + i]);
    const A avg = mean_of(st.sum, st.n);
    A sumsq = (A)0;
    // same path choice as the kernels: one-instruction-shorter term whenever it is exact
    if (sqdev_fast_ok(avg, st.n))
      for (int r = 0; r < R; ++r) sumsq = add_rn(sumsq, sqdev_fast(data[(size_t)r * inner + i], avg));
    else
      for (int r = 0; r < R; ++r) sumsq = add_rn(sumsq, sqdev(data[(size_t)r * inner + i], avg));
    const ColumnResult<A> c = finish_column<A, G_MINMAX | G_RARE>(st, R, sumsq, ignore != 0, ddof, false);
    sum[i] = c.sum;
    mean[i] = c.mean;
    mn[i] = c.mn;
    mx[i] = c.mx;
    prod[i] = c.prod;
    var[i] = c.var;
    sd[i] = c.sd;
    count[i] = c.count;
  }
}

template <typename A>
void order_columns(const A* data, int R, int inner, int ignore, A* first, A* last, A* any_, A* all_) {
  for (int i = 0; i < inner; ++i) {
    ColStats<A, G_ORDER> st;
    st.init();
    for (int r = 0; r < R; ++r) st.add(data[(size_t)r * inner + i]);
    const ColumnResult<A> c = finish_column<A, G_ORDER>(st, R, (A)0, ignore != 0, 1, false);
    first[i] = c.first;
    last[i] = c.last;
    any_[i] = c.any_;
    all_[i] = c.all_;
  }
}

template <int N, typename A>
void median_columns_n(const A* data, int R, int inner, int ignore, A* out) {
  for (int i = 0; i < inner; ++i) {
    A v[N];
    for (int j = 0; j < N; ++j) v[j] = j < R ? data[(size_t)j * inner + i] : Lim<A>::nan();
    out[i] = nanmedian_network<N, A>(v, R, ignore != 0);
  }
}

template <int N, typename A>
void quantile_columns_n(const A* data, int R, int inner, int ignore, int nq, const double* probs, double* qout,
                        A* med) {
  for (int i = 0; i < inner; ++i) {
    A v[N];
    for (int j = 0; j < N; ++j) v[j] = j < R ? data[(size_t)j * inner + i] : Lim<A>::nan();
    int n;
    if (sizeof(A) == 4) {  // the float32 kernels sort NaN last natively and find n by bisection
      sort_nan_last<N, A>(v);
      n = first_nan_index(N, [&](int k) { return v[k]; });
    } else {
      n = sort_valid_first<N, A>(v);
    }
    const bool poisoned = (!ignore && n < R) || n == 0;
    auto fetch = [&](int k) { return v[k]; };
    med[i] = poisoned ? Lim<A>::nan() : median_of_pair<A>(fetch((n - 1) >> 1), fetch(n >> 1));
    for (int q = 0; q < nq; ++q)
      qout[(size_t)q * inner + i] = poisoned ? Lim<double>::nan() : quantile_linear<A>(n, probs[q], fetch);
  }
}

// the split selection of the round-2 kernels: two runs of NH sorted separately (NaN last), ranks by partition search
template <int NH, typename A>
void split_columns_n(const A* data, int R, int inner, int ignore, int nq, const double* probs, double* qout, A* med) {
  constexpr int IT = split_iters(NH);
  for (int i = 0; i < inner; ++i) {
    A v[2 * NH];
    for (int j = 0; j < 2 * NH; ++j) v[j] = j < R ? data[(size_t)j * inner + i] : Lim<A>::nan();
    sortnet<NH, A, false, true>(v);
    sortnet<NH, A, false, true>(v + NH);
    const int na = first_nan_index(NH, [&](int k) { return v[k]; });
    const int nb = first_nan_index(NH, [&](int k) { return v[NH + k]; });
    const int n = na + nb;
    const bool poisoned = (!ignore && n < R) || n == 0;
    auto fa = [&](int k) { return v[k]; };
    auto fb = [&](int k) { return v[NH + k]; };
    med[i] = poisoned ? Lim<A>::nan() : median_two_runs<IT, A>(fa, na, fb, nb);
    for (int q = 0; q < nq; ++q)
      qout[(size_t)q * inner + i] = poisoned ? Lim<double>::nan() : quantile_two_runs<IT, A>(fa, na, fb, nb, probs[q]);
  }
}

// K runs of NH (long columns, 129 .. 512 samples): the same per-run sort, ranks by elimination over K runs
template <int NH, int K, typename A>
void kruns_columns_n(const A* data, int R, int inner, int ignore, int nq, const double* probs, double* qout, A* med) {
  std::vector<A> v(K * NH);
  for (int i = 0; i < inner; ++i) {
    for (int j = 0; j < K * NH; ++j) v[j] = j < R ? data[(size_t)j * inner + i] : Lim<A>::nan();
    int n[K], total = 0;
    for (int r = 0; r < K; ++r) {
      sortnet<NH, A, false, true>(v.data() + r * NH);
      n[r] = first_nan_index(NH, [&](int k) { return v[r * NH + k]; });
      total += n[r];
    }
    const bool poisoned = (!ignore && total < R) || total == 0;
    auto f = [&](int r, int k) { return v[r * NH + k]; };
    med[i] = poisoned ? Lim<A>::nan() : median_k_runs<K, A>(f, n, total);
    for (int q = 0; q < nq; ++q)
      qout[(size_t)q * inner + i] = poisoned ? Lim<double>::nan() : quantile_k_runs<K, A>(f, n, total, probs[q]);
  }
}

// k_select_quad: four runs of NH sorted by the raw network (no canonicalisation pass), parked NH + 1 rows apart with
// a NaN sentinel behind each run, valid counts by branch-free bisection, the middle pair by select_quad_runs
template <int NH, typename A>
void quad_columns_n(const A* data, int R, int inner, int ignore, A* med) {
  constexpr int S = NH + 1;
  for (int i = 0; i < inner; ++i) {
    A col[4 * S];
    for (int r = 0; r < 4; ++r) {
      A v[NH];
      for (int j = 0; j < NH; ++j) v[j] = (r * NH + j) < R ? data[(size_t)(r * NH + j) * inner + i] : Lim<A>::nan();
      sortraw<NH, A>(v);
      for (int j = 0; j < NH; ++j) col[r * S + j] = v[j];
      col[r * S + NH] = Lim<A>::nan();
    }
    auto fetch = [&](int pos) { return col[pos]; };
    int n = 0;
    for (int r = 0; r < 4; ++r) n += quad_valid_prefix<NH>(fetch, r);
    const bool poisoned = (!ignore && n < R) || n == 0;
    med[i] = poisoned ? Lim<A>::nan() : median_quad_runs<NH, A>(fetch, n);
  }
}

// k_select_multi: eight raw-sorted runs of NH with sentinels, ranks by select_multi_runs
template <int NH, typename A>
void multi_columns_n(const A* data, int R, int inner, int ignore, A* med) {
  constexpr int K = 8, S = NH + 1;
  for (int i = 0; i < inner; ++i) {
    A col[K * S];
    for (int r = 0; r < K; ++r) {
      A v[NH];
      for (int j = 0; j < NH; ++j) v[j] = (r * NH + j) < R ? data[(size_t)(r * NH + j) * inner + i] : Lim<A>::nan();
      sortraw<NH, A>(v);
      for (int j = 0; j < NH; ++j) col[r * S + j] = v[j];
      col[r * S + NH] = Lim<A>::nan();
    }
    auto fetch = [&](int pos) { return col[pos]; };
    int n = 0;
    for (int r = 0; r < K; ++r) n += quad_valid_prefix<NH>(fetch, r);
    const bool poisoned = (!ignore && n < R) || n == 0;
    med[i] = poisoned ? Lim<A>::nan() : median_multi_runs<NH, K, A>(fetch, n);
  }
}

template <typename A>
int median_columns(const A* d, int R, int inner, int N, int ignore, A* out) {
  switch (N) {
    case 8: median_columns_n<8, A>(d, R, inner, ignore, out); return 0;
    case 16: median_columns_n<16, A>(d, R, inner, ignore, out); return 0;
    case 32: median_columns_n<32, A>(d, R, inner, ignore, out); return 0;
    case 64: median_columns_n<64, A>(d, R, inner, ignore, out); return 0;
    case 100: median_columns_n<100, A>(d, R, inner, ignore, out); return 0;
    case 128: median_columns_n<128, A>(d, R, inner, ignore, out); return 0;
  }
  return -1;
}

template <typename A>
int quantile_columns(const A* d, int R, int inner, int N, int ignore, int nq, const double* p, double* q, A* med) {
  switch (N) {
    case 8: quantile_columns_n<8, A>(d, R, inner, ignore, nq, p, q, med); return 0;
    case 16: quantile_columns_n<16, A>(d, R, inner, ignore, nq, p, q, med); return 0;
    case 32: quantile_columns_n<32, A>(d, R, inner, ignore, nq, p, q, med); return 0;
    case 64: quantile_columns_n<64, A>(d, R, inner, ignore, nq, p, q, med); return 0;
    case 100: quantile_columns_n<100, A>(d, R, inner, ignore, nq, p, q, med); return 0;
    case 128: quantile_columns_n<128, A>(d, R, inner, ignore, nq, p, q, med); return 0;
  }
  return -1;
}

}  // namespace

extern "C" {

void hm_reduce_f32(const float* d, int R, int inner, int ignore, int ddof, float* sum, float* mean, float* mn,
                   float* mx, float* prod, float* var, float* sd, long long* count) {
  reduce_columns<float>(d, R, inner, ignore, ddof, sum, mean, mn, mx, prod, var, sd, count);
}
void hm_reduce_f64(const double* d, int R, int inner, int ignore, int ddof, double* sum, double* mean,
                   double* mn, double* mx, double* prod, double* var, double* sd, long long* count) {
  reduce_columns<double>(d, R, inner, ignore, ddof, sum, mean, mn, mx, prod, var, sd, count);
}
void hm_order_f32(const float* d, int R, int inner, int ignore, float* first, float* last, float* any_, float* all_) {
  order_columns<float>(d, R, inner, ignore, first, last, any_, all_);
}
void hm_order_f64(const double* d, int R, int inner, int ignore, double* first, double* last, double* any_,
                  double* all_) {
  order_columns<double>(d, R, inner, ignore, first, last, any_, all_);
}
int hm_median_f32(const float* d, int R, int inner, int N, int ignore, float* out) {
  return median_columns<float>(d, R, inner, N, ignore, out);
}
int hm_median_f64(const double* d, int R, int inner, int N, int ignore, double* out) {
  return median_columns<double>(d, R, inner, N, ignore, out);
}
int hm_quantiles_f32(const float* d, int R, int inner, int N, int ignore, int nq, const double* probs,
                     double* qout, float* med) {
  return quantile_columns<float>(d, R, inner, N, ignore, nq, probs, qout, med);
}
int hm_quantiles_f64(const double* d, int R, int inner, int N, int ignore, int nq, const double* probs,
                     double* qout, double* med) {
  return quantile_columns<double>(d, R, inner, N, ignore, nq, probs, qout, med);
}
int hm_split_f32(const float* d, int R, int inner, int NH, int ignore, int nq, const double* probs, double* qout,
                 float* med) {
  switch (NH) {
    case 16: split_columns_n<16, float>(d, R, inner, ignore, nq, probs, qout, med); return 0;
    case 25: split_columns_n<25, float>(d, R, inner, ignore, nq, probs, qout, med); return 0;
    case 32: split_columns_n<32, float>(d, R, inner, ignore, nq, probs, qout, med); return 0;
    case 50: split_columns_n<50, float>(d, R, inner, ignore, nq, probs, qout, med); return 0;
    case 64: split_columns_n<64, float>(d, R, inner, ignore, nq, probs, qout, med); return 0;
  }
  return -1;
}
int hm_quad_f32(const float* d, int R, int inner, int NH, int ignore, float* med) {
  switch (NH) {
    case 6: quad_columns_n<6, float>(d, R, inner, ignore, med); return 0;
    case 8: quad_columns_n<8, float>(d, R, inner, ignore, med); return 0;
    case 12: quad_columns_n<12, float>(d, R, inner, ignore, med); return 0;
    case 13: quad_columns_n<13, float>(d, R, inner, ignore, med); return 0;
    case 14: quad_columns_n<14, float>(d, R, inner, ignore, med); return 0;
    case 16: quad_columns_n<16, float>(d, R, inner, ignore, med); return 0;
    case 20: quad_columns_n<20, float>(d, R, inner, ignore, med); return 0;
    case 25: quad_columns_n<25, float>(d, R, inner, ignore, med); return 0;
    case 32: quad_columns_n<32, float>(d, R, inner, ignore, med); return 0;
  }
  return -1;
}
int hm_multi_f32(const float* d, int R, int inner, int NH, int ignore, float* med) {
  switch (NH) {
    case 20: multi_columns_n<20, float>(d, R, inner, ignore, med); return 0;
    case 25: multi_columns_n<25, float>(d, R, inner, ignore, med); return 0;
    case 32: multi_columns_n<32, float>(d, R, inner, ignore, med); return 0;
    case 40: multi_columns_n<40, float>(d, R, inner, ignore, med); return 0;
    case 50: multi_columns_n<50, float>(d, R, inner, ignore, med); return 0;
    case 64: multi_columns_n<64, float>(d, R, inner, ignore, med); return 0;
    case 128: multi_columns_n<128, float>(d, R, inner, ignore, med); return 0;
  }
  return -1;
}
int hm_kruns_f32(const float* d, int R, int inner, int NH, int K, int ignore, int nq, const double* probs,
                 double* qout, float* med) {
#define EOC_KR(nh, k) \
  if (NH == nh && K == k) { kruns_columns_n<nh, k, float>(d, R, inner, ignore, nq, probs, qout, med); return 0; }
  EOC_KR(16, 3) EOC_KR(16, 4) EOC_KR(25, 3) EOC_KR(100, 3) EOC_KR(100, 4) EOC_KR(128, 3) EOC_KR(128, 4) EOC_KR(100, 2) EOC_KR(128, 2)
#undef EOC_KR
  return -1;
}
void hm_nd_f32(const float* a, const float* b, int n, float* out) {
  for (int i = 0; i < n; ++i) out[i] = div_rn(sub_rn(a[i], b[i]), add_rn(a[i], b[i]));
}

}  // extern "C"
