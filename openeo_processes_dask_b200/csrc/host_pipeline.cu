// Host-buffer entry points: strip pipeline host -> device -> kernel -> host (eoc_reduce_host,
// eoc_select_host).  This is the device-array bridge a numpy / dask chunk goes through.
//
// The cube view [outer, R, inner] is cut along `inner` into strips.  Strip k uses stream k%2 and
// its own staging buffers, so the H2D copy of strip k+1 overlaps the kernel and the D2H copy of
// strip k.  Copies are cudaMemcpy2DAsync straight from/to the caller's buffers (rows of one strip
// are `stride_R` elements apart in the cube and dense on the device); with pinned host buffers
// they run at PCIe speed, with pageable buffers the driver stages them.
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "eoc_common.cuh"

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace eoc {
namespace {

// Copy into the pinned staging strip with NON-TEMPORAL stores: the strip is written once and then read by the DMA
// engine only, so pulling its lines into the cache first (read-for-ownership) wastes a third of the host memory
// traffic of the gather (read source + RFO + write-back -> read source + write).  Rows of a strip are ~100 KB, below
// the size at which memcpy itself switches to streaming stores.  EOC_HOST_NT=0 restores plain memcpy.
#if defined(__x86_64__)
__attribute__((target("avx2"))) void copy_stream_avx2(char* dst, const char* src, size_t n) {
  size_t head = (32 - ((uintptr_t)dst & 31)) & 31;
  if (head > n) head = n;
  if (head) {
    std::memcpy(dst, src, head);
    dst += head;
    src += head;
    n -= head;
  }
  size_t i = 0;
  for (; i + 128 <= n; i += 128) {
    const __m256i a = _mm256_loadu_si256((const __m256i*)(src + i));
    const __m256i b = _mm256_loadu_si256((const __m256i*)(src + i + 32));
    const __m256i c = _mm256_loadu_si256((const __m256i*)(src + i + 64));
    const __m256i d = _mm256_loadu_si256((const __m256i*)(src + i + 96));
    _mm256_stream_si256((__m256i*)(dst + i), a);
    _mm256_stream_si256((__m256i*)(dst + i + 32), b);
    _mm256_stream_si256((__m256i*)(dst + i + 64), c);
    _mm256_stream_si256((__m256i*)(dst + i + 96), d);
  }
  if (i < n) std::memcpy(dst + i, src + i, n - i);
}
#endif
bool use_stream_copy() {
  static const bool on = [] {
    const char* e = getenv("EOC_HOST_NT");
    if (e && e[0] == '0') return false;
#if defined(__x86_64__)
    return (bool)__builtin_cpu_supports("avx2");
#else
    return false;
#endif
  }();
  return on;
}
inline void copy_row(char* dst, const char* src, size_t n, bool stream) {
#if defined(__x86_64__)
  if (stream && n >= 4096) {
    copy_stream_avx2(dst, src, n);
    return;
  }
#endif
  (void)stream;
  std::memcpy(dst, src, n);
}

struct Lane {
  cudaStream_t stream = nullptr;
  void* in = nullptr;
  size_t in_bytes = 0;
  void* out = nullptr;
  size_t out_bytes = 0;
  // pageable input only: pinned staging strip + "its H2D has drained" event
  void* stage = nullptr;
  size_t stage_bytes = 0;
  cudaEvent_t staged = nullptr;
  bool staged_pending = false;
};

struct HostCtx {
  int device = -1;
  Lane lane[2];
  void* aux = nullptr;  // small per-call device arrays (segment tables)
  size_t aux_bytes = 0;
  ~HostCtx() { release(); }
  void release() {
    if (aux) cudaFree(aux);
    aux = nullptr;
    aux_bytes = 0;
    for (Lane& l : lane) {
      if (l.in) cudaFree(l.in);
      if (l.out) cudaFree(l.out);
      if (l.stage) cudaFreeHost(l.stage);
      if (l.staged) cudaEventDestroy(l.staged);
      if (l.stream) cudaStreamDestroy(l.stream);
      l = Lane();
    }
    device = -1;
  }
};

thread_local HostCtx tl_ctx;

int ensure(Lane& l, size_t in_bytes, size_t out_bytes) {
  if (!l.stream) EOC_CUDA(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
  if (l.in_bytes < in_bytes) {
    if (l.in) EOC_CUDA(cudaFree(l.in));
    l.in = nullptr;
    l.in_bytes = 0;
    EOC_CUDA(cudaMalloc(&l.in, in_bytes));
    l.in_bytes = in_bytes;
  }
  if (l.out_bytes < out_bytes) {
    if (l.out) EOC_CUDA(cudaFree(l.out));
    l.out = nullptr;
    l.out_bytes = 0;
    EOC_CUDA(cudaMalloc(&l.out, out_bytes));
    l.out_bytes = out_bytes;
  }
  return EOC_OK;
}

int ctx_for_current_device() {
  int dev = 0;
  EOC_CUDA(cudaGetDevice(&dev));
  if (tl_ctx.device != dev) {
    tl_ctx.release();
    tl_ctx.device = dev;
  }
  return EOC_OK;
}

// cudaMemcpy2D's documented limit on a pitch (cudaDevAttrMaxPitch, 2 GiB - 1): a t-reduction of a pinned
// (t, b, y, x) float32 cube whose planes exceed 2 GiB has a larger source pitch.  Such strips go row by
// row -- one cudaMemcpyAsync per row of the strip -- instead of failing.
static size_t max_pitch() {
  if (const char* e = getenv("EOC_MAX_PITCH")) {  // tests: force the row-by-row path on a small cube
    if (e[0]) return (size_t)atoll(e);
  }
  static size_t p = [] {
    int v = 0, dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrMaxPitch, dev) != cudaSuccess || v <= 0) {
      cudaGetLastError();
      return (size_t)0x7fffffff;
    }
    return (size_t)v;
  }();
  return p;
}

// copy strip [i0, i0+w) of the host view into a dense device strip [outer, R, w]
int h2d_strip(const eoc_view& v, int64_t i0, int64_t w, void* dst, cudaStream_t s) {
  const size_t es = dtype_size(v.dtype);
  const char* base = (const char*)v.data + (size_t)i0 * es;
  if ((size_t)v.stride_R * es > max_pitch()) {
    for (int64_t o = 0; o < v.outer; ++o)
      for (int64_t r = 0; r < v.R; ++r)
        EOC_CUDA(cudaMemcpyAsync((char*)dst + (size_t)((o * v.R + r) * w) * es,
                                 base + (size_t)(o * v.stride_outer + r * v.stride_R) * es, (size_t)w * es,
                                 cudaMemcpyHostToDevice, s));
    return EOC_OK;
  }
  if (v.outer == 1 || v.stride_outer == v.R * v.stride_R) {
    EOC_CUDA(cudaMemcpy2DAsync(dst, (size_t)w * es, base, (size_t)v.stride_R * es, (size_t)w * es,
                               (size_t)(v.outer * v.R), cudaMemcpyHostToDevice, s));
  } else {
    for (int64_t o = 0; o < v.outer; ++o)
      EOC_CUDA(cudaMemcpy2DAsync((char*)dst + (size_t)(o * v.R * w) * es, (size_t)w * es,
                                 base + (size_t)(o * v.stride_outer) * es, (size_t)v.stride_R * es,
                                 (size_t)w * es, (size_t)v.R, cudaMemcpyHostToDevice, s));
  }
  return EOC_OK;
}

// Is this host buffer page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory)?
bool host_is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();  // plain malloc memory on old drivers: an error that must not stick
    return false;
  }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// PAGEABLE input (an ordinary numpy array, a dask chunk): cudaMemcpy2DAsync would go through the driver's own
// staging copy, one host thread, ~10 GB/s end to end against 55 GB/s for a pinned cube (tools/pageable_e2e.py).
// Here a few host threads gather the strip's rows into the lane's pinned staging strip and ONE dense async copy
// moves it; the gather of strip k+1 overlaps the DMA of strip k (two lanes), and the staging strip is only
// refilled once its previous DMA has drained (event).
int h2d_strip_staged(const eoc_view& v, int64_t i0, int64_t w, Lane& L, size_t cap_bytes, void* dst = nullptr) {
  const size_t es = dtype_size(v.dtype);
  const int64_t rows = v.outer * v.R;
  const size_t row_bytes = (size_t)w * es;
  if (L.stage_bytes < cap_bytes) {
    if (L.staged_pending) EOC_CUDA(cudaEventSynchronize(L.staged));
    L.staged_pending = false;
    if (L.stage) EOC_CUDA(cudaFreeHost(L.stage));
    L.stage = nullptr;
    L.stage_bytes = 0;
    EOC_CUDA(cudaHostAlloc(&L.stage, cap_bytes, cudaHostAllocDefault));
    L.stage_bytes = cap_bytes;
  }
  if (!L.staged) EOC_CUDA(cudaEventCreateWithFlags(&L.staged, cudaEventDisableTiming));
  if (L.staged_pending) EOC_CUDA(cudaEventSynchronize(L.staged));
  const char* base = (const char*)v.data + (size_t)i0 * es;
  char* stage = (char*)L.stage;
  const int64_t R = v.R, so = v.stride_outer, sr = v.stride_R;
  const bool nt = use_stream_copy();
  auto gather = [=](int64_t q0, int64_t q1) {
    for (int64_t q = q0; q < q1; ++q) {
      const int64_t o = q / R, r = q - o * R;
      copy_row(stage + (size_t)q * row_bytes, base + (size_t)(o * so + r * sr) * es, row_bytes, nt);
    }
#if defined(__x86_64__)
    if (nt) _mm_sfence();  // the streamed lines are globally visible before the DMA is queued
#endif
  };
  // gather threads: EOC_HOST_THREADS, else the host CPUs divided by the visible GPUs (one process per GPU), at
  // most 8.  Measured on the box (2.25 GB cube, one GPU): 1 / 2 / 4 / 8 / 16 threads -> 9 / 16 / 22 / 34 / 36 GB/s.
  unsigned hw = std::thread::hardware_concurrency();
  int ngpu = 1;
  if (cudaGetDeviceCount(&ngpu) != cudaSuccess || ngpu < 1) ngpu = 1;
  int nthr = (int)hw / ngpu;
  nthr = nthr > 16 ? 16 : (nthr < 1 ? 1 : nthr);
  if (const char* e = getenv("EOC_HOST_THREADS")) {
    const int n = atoi(e);
    if (n >= 1 && n <= 64) nthr = n;
  }
  if ((size_t)rows * row_bytes < ((size_t)4 << 20) || rows < nthr) nthr = 1;
  if (nthr == 1) {
    gather(0, rows);
  } else {
    std::vector<std::thread> pool;
    const int64_t per = (rows + nthr - 1) / nthr;
    for (int t = 1; t < nthr; ++t)
      pool.emplace_back(gather, t * per, (t + 1) * per < rows ? (t + 1) * per : rows);
    gather(0, per < rows ? per : rows);
    for (std::thread& th : pool) th.join();
  }
  EOC_CUDA(cudaMemcpyAsync(dst ? dst : L.in, L.stage, (size_t)rows * row_bytes, cudaMemcpyHostToDevice, L.stream));
  EOC_CUDA(cudaEventRecord(L.staged, L.stream));
  L.staged_pending = true;
  return EOC_OK;
}

// copy a dense device result [rows, w] into columns [i0, i0+w) of a dense host result [rows, inner]
int d2h_strip(void* host_out, const void* dev, int64_t rows, int64_t inner, int64_t i0, int64_t w,
              size_t es, cudaStream_t s) {
  if ((size_t)inner * es > max_pitch()) {
    for (int64_t r = 0; r < rows; ++r)
      EOC_CUDA(cudaMemcpyAsync((char*)host_out + (size_t)(r * inner + i0) * es, (const char*)dev + (size_t)(r * w) * es,
                               (size_t)w * es, cudaMemcpyDeviceToHost, s));
    return EOC_OK;
  }
  EOC_CUDA(cudaMemcpy2DAsync((char*)host_out + (size_t)i0 * es, (size_t)inner * es, dev, (size_t)w * es,
                             (size_t)w * es, (size_t)rows, cudaMemcpyDeviceToHost, s));
  return EOC_OK;
}

// strip plan: width along inner (multiple of 1024 unless it is all of inner) and, when even a
// 1024-wide strip of every outer slice is over budget, a chunk of outer slices per strip
void strip_plan(const eoc_view& v, size_t strip_bytes, int64_t* oc, int64_t* w_out) {
  if (strip_bytes == 0) strip_bytes = (size_t)256 << 20;
  const size_t es = dtype_size(v.dtype);
  int64_t w = (int64_t)(strip_bytes / ((size_t)v.outer * (size_t)v.R * es + 1));
  w = w / 1024 * 1024;  // keeps every strip start 16-byte aligned for the vector kernels
  if (w < 1024) w = 1024;
  if (w > v.inner) w = v.inner;
  int64_t o = (int64_t)(strip_bytes / ((size_t)v.R * (size_t)w * es + 1));
  if (o < 1) o = 1;
  if (o > v.outer) o = v.outer;
  *oc = o;
  *w_out = w;
}

// A device -> host copy into a PAGEABLE result (a fresh numpy array) blocks the issuing thread until the
// strip's kernel has run.  Issued right after the kernel, that stall would keep the NEXT strip's H2D from being
// queued and leave the copy engine idle once per strip (measured: 51.5 GB/s end to end against 55.5 GB/s of raw
// pinned H2D).  The result copies of strip k are therefore recorded and only issued after strip k+1's H2D and
// kernel are in their stream.
struct PendingD2H {
  void* host;
  const void* dev;
  int64_t rows, inner, i0, w;
  size_t es;
};
int flush_d2h(std::vector<PendingD2H>& q, cudaStream_t s) {
  for (const PendingD2H& c : q)
    if (int rc = d2h_strip(c.host, c.dev, c.rows, c.inner, c.i0, c.w, c.es, s)) return rc;
  q.clear();
  return EOC_OK;
}

size_t out_elem_size(int op, int accdt) {
  return (op == EOC_OP_COUNT || op == EOC_OP_NOT_NAN) ? 8 : dtype_size(accdt);
}

// ---- generic two-lane strip driver (round 2): every host entry point besides reduce / select --------------------
// `nin` host views of identical [outer, R, inner] geometry are cut along `inner`; strip k is copied to lane k % 2
// (pinned: cudaMemcpy2DAsync from the caller's buffer; pageable: threaded gather into the lane's pinned staging
// strip + one dense copy), `launch` runs the kernels on the lane's stream with dense device views, and every
// result -- host array dense [outer * rows_per_outer, inner] -- is copied back behind the NEXT strip's H2D.
struct HostOut {
  void* host;
  size_t es;
  int64_t rows_per_outer;  // 1 for reducers, nseg for a segmented reduce, R for axis-kept processes
};

template <typename Launch>
int strip_pipeline(const eoc_view* const* ins, int nin, const std::vector<HostOut>& outs, size_t strip_bytes,
                   Launch launch) {
  if (int rc = ctx_for_current_device()) return rc;
  const eoc_view& v = *ins[0];
  const size_t es = dtype_size(v.dtype);
  bool pinned = true;
  for (int a = 0; a < nin; ++a) pinned = pinned && host_is_pinned(ins[a]->data);
  if (!pinned && strip_bytes == 0) strip_bytes = (size_t)64 << 20;
  int64_t w, oc;
  // the budget covers the inputs AND the axis-kept outputs of a strip
  int64_t kept = 0;
  for (const HostOut& o : outs) kept += o.rows_per_outer >= v.R ? 1 : 0;
  eoc_view plan = v;
  size_t budget = strip_bytes ? strip_bytes : ((size_t)256 << 20);
  budget /= (size_t)(nin + 2 * kept > 0 ? nin + 2 * kept : 1);
  strip_plan(plan, budget, &oc, &w);
  const size_t in_one = ((size_t)oc * v.R * w * es + 255) / 256 * 256;
  std::vector<size_t> out_off(outs.size() + 1, 0);
  for (size_t q = 0; q < outs.size(); ++q)
    out_off[q + 1] = out_off[q] + ((size_t)oc * outs[q].rows_per_outer * w * outs[q].es + 255) / 256 * 256;
  int k = 0;
  std::vector<PendingD2H> pending[2];
  auto fail_sync = [&](int rc) {  // never return with copies into the caller's buffers still in flight
    for (int l = 0; l < 2; ++l)
      if (tl_ctx.lane[l].stream) cudaStreamSynchronize(tl_ctx.lane[l].stream);
    return rc;
  };
  for (int64_t o0 = 0; o0 < v.outer; o0 += oc) {
    const int64_t ock = (o0 + oc <= v.outer) ? oc : v.outer - o0;
    for (int64_t i0 = 0; i0 < v.inner; i0 += w, ++k) {
      const int64_t wk = (i0 + w <= v.inner) ? w : v.inner - i0;
      Lane& L = tl_ctx.lane[k & 1];
      if (int rc = flush_d2h(pending[k & 1], L.stream)) return fail_sync(rc);
      if (int rc = ensure(L, in_one * nin, out_off.back() ? out_off.back() : 256)) return fail_sync(rc);
      std::vector<eoc_view> dv(nin);
      for (int a = 0; a < nin; ++a) {
        eoc_view hv = *ins[a];
        hv.data = (const char*)hv.data + (size_t)(o0 * hv.stride_outer) * es;
        hv.outer = ock;
        void* dst = (char*)L.in + in_one * a;
        if (pinned) {
          if (int rc = h2d_strip(hv, i0, wk, dst, L.stream)) return fail_sync(rc);
        } else {
          if (int rc = h2d_strip_staged(hv, i0, wk, L, (size_t)oc * v.R * w * es, dst)) return fail_sync(rc);
        }
        dv[a] = hv;
        dv[a].data = dst;
        dv[a].inner = wk;
        dv[a].stride_R = wk;
        dv[a].stride_outer = v.R * wk;
      }
      std::vector<void*> dout(outs.size());
      for (size_t q = 0; q < outs.size(); ++q) dout[q] = (char*)L.out + out_off[q];
      if (int rc = launch(dv.data(), dout.data(), L.stream)) return fail_sync(rc);
      for (size_t q = 0; q < outs.size(); ++q)
        if (outs[q].host)
          pending[k & 1].push_back({(char*)outs[q].host + (size_t)(o0 * outs[q].rows_per_outer * v.inner) * outs[q].es,
                                    dout[q], ock * outs[q].rows_per_outer, v.inner, i0, wk, outs[q].es});
      if (tl_ctx.lane[(k + 1) & 1].stream)
        if (int rc = flush_d2h(pending[(k + 1) & 1], tl_ctx.lane[(k + 1) & 1].stream)) return fail_sync(rc);
    }
  }
  for (int l = 0; l < 2; ++l)
    if (tl_ctx.lane[l].stream)
      if (int rc = flush_d2h(pending[l], tl_ctx.lane[l].stream)) return fail_sync(rc);
  EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[0].stream));
  if (tl_ctx.lane[1].stream) EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[1].stream));
  return EOC_OK;
}

int ensure_aux(size_t bytes) {
  if (tl_ctx.aux_bytes < bytes) {
    if (tl_ctx.aux) EOC_CUDA(cudaFree(tl_ctx.aux));
    tl_ctx.aux = nullptr;
    tl_ctx.aux_bytes = 0;
    EOC_CUDA(cudaMalloc(&tl_ctx.aux, bytes));
    tl_ctx.aux_bytes = bytes;
  }
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" {

int eoc_reduce_host(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                    size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_reduce_host")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_reduce_host: outs is NULL");
  if (in->outer * in->inner == 0) return EOC_OK;
  if (int rc = ctx_for_current_device()) return rc;
  const eoc_view& v = *in;
  const int accdt = acc_dtype(v.dtype);
  const size_t es = dtype_size(v.dtype);
  int64_t w, oc;
  const bool pinned = host_is_pinned(v.data);
  if (!pinned && strip_bytes == 0) strip_bytes = (size_t)64 << 20;  // staging strips are pinned host memory: keep them small
  strip_plan(v, strip_bytes, &oc, &w);
  // device result block: one dense [oc, w] plane per selected reducer, 8-byte slots
  int nsel = 0;
  for (int id = 0; id < EOC_NUM_OPS; ++id) nsel += (ops >> id) & 1;
  const size_t plane = (size_t)oc * (size_t)w * 8;
  {  // reject what the strips cannot do BEFORE the first copy is queued (no return with copies in flight)
    eoc_view probe = v;
    probe.inner = w;
    probe.stride_R = w;
    probe.stride_outer = v.R * w;
    probe.outer = oc;
    if (reduce_workspace_bytes(probe, ops) != 0)
      EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_host: long-row (spatial) reductions need the device entry point");
  }
  int k = 0;
  std::vector<PendingD2H> pending[2];
  for (int64_t o0 = 0; o0 < v.outer; o0 += oc) {
    const int64_t ock = (o0 + oc <= v.outer) ? oc : v.outer - o0;
    for (int64_t i0 = 0; i0 < v.inner; i0 += w, ++k) {
      const int64_t wk = (i0 + w <= v.inner) ? w : v.inner - i0;
      Lane& L = tl_ctx.lane[k & 1];
      if (int rc = flush_d2h(pending[k & 1], L.stream)) return rc;  // (only non-empty if a lane is reused early)
      if (int rc = ensure(L, (size_t)oc * v.R * w * es, plane * (nsel ? nsel : 1))) return rc;
      eoc_view hv = v;
      hv.data = (const char*)v.data + (size_t)(o0 * v.stride_outer) * es;
      hv.outer = ock;
      if (pinned) {
        if (int rc = h2d_strip(hv, i0, wk, L.in, L.stream)) return rc;
      } else {
        if (int rc = h2d_strip_staged(hv, i0, wk, L, (size_t)oc * v.R * w * es)) return rc;
      }
      eoc_view dv = hv;
      dv.data = L.in;
      dv.inner = wk;
      dv.stride_R = wk;
      dv.stride_outer = v.R * wk;
      if (reduce_workspace_bytes(dv, ops) != 0)
        EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_host: long-row (spatial) reductions need the device entry point");
      OutPtrs o;
      int slot = 0;
      for (int id = 0; id < EOC_NUM_OPS; ++id)
        o.p[id] = (ops & (1u << id)) ? (char*)L.out + plane * (slot++) : nullptr;
      if (int rc = launch_reduce(dv, ops, ignore_nodata, ddof, o, nullptr, 0, L.stream)) return rc;
      for (int id = 0; id < EOC_NUM_OPS; ++id)
        if (ops & (1u << id)) {
          const size_t oes = out_elem_size(id, accdt);
          pending[k & 1].push_back({(char*)outs[id] + (size_t)(o0 * v.inner) * oes, o.p[id], ock, v.inner, i0, wk, oes});
        }
      // the previous strip's results go out now, behind this strip's H2D in the copy queue
      if (tl_ctx.lane[(k + 1) & 1].stream)
        if (int rc = flush_d2h(pending[(k + 1) & 1], tl_ctx.lane[(k + 1) & 1].stream)) return rc;
    }
  }
  for (int l = 0; l < 2; ++l)
    if (tl_ctx.lane[l].stream)
      if (int rc = flush_d2h(pending[l], tl_ctx.lane[l].stream)) return rc;
  EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[0].stream));
  if (tl_ctx.lane[1].stream) EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[1].stream));
  return EOC_OK;
}

int eoc_select_fused_host(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                          void* quant_out, int quant_dtype, uint32_t ops, int ddof, void* const* outs,
                          size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_select_host")) return rc;
  if (in->outer * in->inner == 0) return EOC_OK;
  if (nq < 0 || nq > EOC_SELECT_MAX_Q) EOC_FAIL(EOC_ERR_INVALID, "nq=%d outside [0, %d]", nq, EOC_SELECT_MAX_Q);
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_select_fused_host: outs is NULL");
  if (int rc = ctx_for_current_device()) return rc;
  const eoc_view& v = *in;
  const size_t es = dtype_size(v.dtype);
  const int accdt = acc_dtype(v.dtype);
  const size_t aes = dtype_size(accdt);
  const size_t qes = quant_dtype == EOC_F64 ? 8 : 4;
  int64_t w, oc;
  const bool pinned = host_is_pinned(v.data);
  if (!pinned && strip_bytes == 0) strip_bytes = (size_t)64 << 20;
  strip_plan(v, strip_bytes, &oc, &w);
  int nsel = 0;
  for (int id = 0; id < EOC_NUM_OPS; ++id) nsel += (ops >> id) & 1;
  const size_t plane = (size_t)oc * w * 8;  // one dense [oc, w] result plane, 8-byte slots
  int k = 0;
  std::vector<PendingD2H> pending[2];
  for (int64_t o0 = 0; o0 < v.outer; o0 += oc) {
    const int64_t ock = (o0 + oc <= v.outer) ? oc : v.outer - o0;
    for (int64_t i0 = 0; i0 < v.inner; i0 += w, ++k) {
      const int64_t wk = (i0 + w <= v.inner) ? w : v.inner - i0;
      Lane& L = tl_ctx.lane[k & 1];
      if (int rc = flush_d2h(pending[k & 1], L.stream)) return rc;
      if (int rc = ensure(L, (size_t)oc * v.R * w * es, plane * (size_t)(1 + nq + nsel))) return rc;
      eoc_view hv = v;
      hv.data = (const char*)v.data + (size_t)(o0 * v.stride_outer) * es;
      hv.outer = ock;
      if (pinned) {
        if (int rc = h2d_strip(hv, i0, wk, L.in, L.stream)) return rc;
      } else {
        if (int rc = h2d_strip_staged(hv, i0, wk, L, (size_t)oc * v.R * w * es)) return rc;
      }
      eoc_view dv = hv;
      dv.data = L.in;
      dv.inner = wk;
      dv.stride_R = wk;
      dv.stride_outer = v.R * wk;
      void* dmed = median_out ? L.out : nullptr;
      void* dq = nq ? (char*)L.out + plane : nullptr;
      OutPtrs o;
      int slot = 0;
      for (int id = 0; id < EOC_NUM_OPS; ++id)
        o.p[id] = (ops & (1u << id)) ? (char*)L.out + plane * (size_t)(1 + nq + slot++) : nullptr;
      if (int rc = launch_select(dv, ignore_nodata, dmed, nq, probs, dq, quant_dtype, ops, ddof, o, L.stream))
        return rc;
      std::vector<PendingD2H>& mine = pending[k & 1];
      if (median_out) mine.push_back({(char*)median_out + (size_t)(o0 * v.inner) * aes, dmed, ock, v.inner, i0, wk, aes});
      if (nq) mine.push_back({(char*)quant_out + (size_t)(o0 * nq * v.inner) * qes, dq, ock * nq, v.inner, i0, wk, qes});
      for (int id = 0; id < EOC_NUM_OPS; ++id)
        if (ops & (1u << id)) {
          const size_t oes = out_elem_size(id, accdt);
          mine.push_back({(char*)outs[id] + (size_t)(o0 * v.inner) * oes, o.p[id], ock, v.inner, i0, wk, oes});
        }
      if (tl_ctx.lane[(k + 1) & 1].stream)
        if (int rc = flush_d2h(pending[(k + 1) & 1], tl_ctx.lane[(k + 1) & 1].stream)) return rc;
    }
  }
  for (int l = 0; l < 2; ++l)
    if (tl_ctx.lane[l].stream)
      if (int rc = flush_d2h(pending[l], tl_ctx.lane[l].stream)) return rc;
  EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[0].stream));
  if (tl_ctx.lane[1].stream) EOC_CUDA(cudaStreamSynchronize(tl_ctx.lane[1].stream));
  return EOC_OK;
}

int eoc_select_host(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                    void* quant_out, int quant_dtype, size_t strip_bytes) {
  return eoc_select_fused_host(in, ignore_nodata, median_out, nq, probs, quant_out, quant_dtype, 0u, 1, nullptr,
                               strip_bytes);
}


// ---- round 2: the remaining host entry points on the generic strip driver ---------------------------------------
int eoc_segmented_reduce_host(const eoc_view* in, int nseg, const int32_t* seg_offsets, const int32_t* members,
                              int max_seg_len, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                              size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_segmented_reduce_host")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_segmented_reduce_host: outs is NULL");
  if (nseg < 1 || !seg_offsets) EOC_FAIL(EOC_ERR_INVALID, "segmented reduce needs nseg >= 1 and offsets");
  if (in->outer * in->inner == 0) return EOC_OK;
  if (int rc = ctx_for_current_device()) return rc;
  const int nmem = seg_offsets[nseg];
  const size_t off_bytes = ((size_t)(nseg + 1) * 4 + 255) / 256 * 256;
  if (int rc = ensure_aux(off_bytes + (size_t)(nmem > 0 ? nmem : 1) * 4)) return rc;
  EOC_CUDA(cudaMemcpy(tl_ctx.aux, seg_offsets, (size_t)(nseg + 1) * 4, cudaMemcpyHostToDevice));
  if (nmem > 0) EOC_CUDA(cudaMemcpy((char*)tl_ctx.aux + off_bytes, members, (size_t)nmem * 4, cudaMemcpyHostToDevice));
  const int32_t* d_off = (const int32_t*)tl_ctx.aux;
  const int32_t* d_mem = (const int32_t*)((char*)tl_ctx.aux + off_bytes);
  const int accdt = acc_dtype(in->dtype);
  std::vector<HostOut> ho;
  std::vector<int> ids;
  for (int id = 0; id < EOC_NUM_OPS; ++id)
    if (ops & (1u << id)) {
      if (!outs[id]) EOC_FAIL(EOC_ERR_INVALID, "outs[%d] is NULL but selected by ops mask 0x%x", id, ops);
      ho.push_back({outs[id], out_elem_size(id, accdt), nseg});
      ids.push_back(id);
    }
  const eoc_view* ins[1] = {in};
  return strip_pipeline(ins, 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    OutPtrs o{};
    for (size_t q = 0; q < ids.size(); ++q) o.p[ids[q]] = dout[q];
    return launch_segmented(dv[0], nseg, d_off, d_mem, max_seg_len, ops, ignore_nodata, ddof, o, s);
  });
}

int eoc_nd_reduce_host(const eoc_view* a, const eoc_view* b, uint32_t ops, int ignore_nodata, int ddof,
                       void* const* outs, void* nd_out, int integer_wrap, size_t strip_bytes) {
  if (int rc = check_view(a, "eoc_nd_reduce_host(a)")) return rc;
  if (int rc = check_view(b, "eoc_nd_reduce_host(b)")) return rc;
  if (!outs && ops) EOC_FAIL(EOC_ERR_INVALID, "eoc_nd_reduce_host: outs is NULL");
  if (a->dtype != b->dtype || a->outer != b->outer || a->R != b->R || a->inner != b->inner)
    EOC_FAIL(EOC_ERR_INVALID, "normalized difference operands differ in dtype or shape");
  if (ops == 0 && !nd_out) EOC_FAIL(EOC_ERR_INVALID, "nothing to compute: ops == 0 and nd_out == NULL");
  if (a->outer * a->R * a->inner == 0) return EOC_OK;
  const bool isint = a->dtype != EOC_F32 && a->dtype != EOC_F64;
  const int accdt = (integer_wrap && isint) ? EOC_F64 : nd_acc_dtype(a->dtype);
  std::vector<HostOut> ho;
  std::vector<int> ids;
  for (int id = 0; id < EOC_NUM_OPS; ++id)
    if (ops & (1u << id)) {
      if (!outs[id]) EOC_FAIL(EOC_ERR_INVALID, "outs[%d] is NULL but selected by ops mask 0x%x", id, ops);
      ho.push_back({outs[id], out_elem_size(id, accdt), 1});
      ids.push_back(id);
    }
  const bool want_nd = nd_out != nullptr;
  if (want_nd) ho.push_back({nd_out, dtype_size(accdt), a->R});
  const eoc_view* ins[2] = {a, b};
  return strip_pipeline(ins, 2, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    OutPtrs o{};
    for (size_t q = 0; q < ids.size(); ++q) o.p[ids[q]] = dout[q];
    return launch_nd_reduce(dv[0], dv[1], ops, ignore_nodata, ddof, o, want_nd ? dout[ids.size()] : nullptr,
                            integer_wrap != 0, s);
  });
}

int eoc_scan_host(const eoc_view* in, int op, int ignore_nodata, void* out, size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_scan_host")) return rc;
  if (!out) EOC_FAIL(EOC_ERR_INVALID, "eoc_scan_host: out is NULL");
  if (in->outer * in->R * in->inner == 0) return EOC_OK;
  std::vector<HostOut> ho{{out, dtype_size(acc_dtype(in->dtype)), in->R}};
  const eoc_view* ins[1] = {in};
  return strip_pipeline(ins, 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    return launch_scan(dv[0], op, ignore_nodata, dout[0], s);
  });
}

int eoc_sort_host(const eoc_view* in, int descending, int nan_first, void* out, size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_sort_host")) return rc;
  if (!out) EOC_FAIL(EOC_ERR_INVALID, "eoc_sort_host: out is NULL");
  if (in->R > EOC_SELECT_MAX_R)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_sort_host: R=%lld exceeds EOC_SELECT_MAX_R=%d (stage the array and use "
             "eoc_select_general)", (long long)in->R, EOC_SELECT_MAX_R);
  if (in->outer * in->R * in->inner == 0) return EOC_OK;
  std::vector<HostOut> ho{{out, dtype_size(acc_dtype(in->dtype)), in->R}};
  const eoc_view* ins[1] = {in};
  SortSpec sort;
  sort.descending = descending;
  sort.nan_first = nan_first;
  return strip_pipeline(ins, 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    SortSpec sp = sort;
    sp.out = dout[0];
    return launch_select(dv[0], 1, nullptr, 0, nullptr, nullptr, EOC_F64, 0u, 1, OutPtrs{}, s, sp);
  });
}

// array_interpolate_linear of a numpy chunk: strips along `inner`, the (host) positions copied to the thread's table
// buffer once.  out: dense [outer, R, inner] in the input dtype (float32 / float64).
int eoc_interpolate_linear_host(const eoc_view* in, const double* x, void* out, size_t strip_bytes) {
  if (int rc = check_view(in, "eoc_interpolate_linear_host")) return rc;
  if (!out) EOC_FAIL(EOC_ERR_INVALID, "eoc_interpolate_linear_host: out is NULL");
  if (in->dtype != EOC_F32 && in->dtype != EOC_F64)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "interpolation is defined for floating-point samples (integers have no gaps)");
  if (in->outer * in->R * in->inner == 0) return EOC_OK;
  if (int rc = ctx_for_current_device()) return rc;
  const double* xd = nullptr;
  if (x) {
    if (int rc = ensure_aux((size_t)in->R * sizeof(double))) return rc;
    EOC_CUDA(cudaMemcpy(tl_ctx.aux, x, (size_t)in->R * sizeof(double), cudaMemcpyHostToDevice));
    xd = (const double*)tl_ctx.aux;
  }
  std::vector<HostOut> ho{{out, dtype_size(in->dtype), in->R}};
  const eoc_view* ins[1] = {in};
  return strip_pipeline(ins, 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    return launch_interpolate_linear(dv[0], xd, dout[0], s);
  });
}

// elementwise processes / recorded chains over a dense numpy array of n samples: the flat array is the view
// [1, 1, n], cut into strips; out in the compute dtype (eoc_elementwise) / out_dtype (eoc_apply_tape).
static eoc_view flat_view(const void* p, int dtype, int64_t n) {
  eoc_view v;
  v.data = p;
  v.dtype = dtype;
  v.outer = 1;
  v.R = 1;
  v.inner = n;
  v.stride_outer = n;
  v.stride_R = n;
  return v;
}
int eoc_elementwise_host(const void* in, int dtype, int64_t n, int op, const double* params, void* out,
                         size_t strip_bytes) {
  if (n < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_elementwise_host: negative length");
  if (n == 0) return EOC_OK;
  if (!in || !out) EOC_FAIL(EOC_ERR_INVALID, "eoc_elementwise_host: NULL pointer");
  const eoc_view v = flat_view(in, dtype, n);
  if (int rc = check_view(&v, "eoc_elementwise_host")) return rc;
  std::vector<HostOut> ho{{out, dtype_size(acc_dtype(dtype)), 1}};
  const eoc_view* ins[1] = {&v};
  return strip_pipeline(ins, 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    return launch_elementwise(dv[0].data, dtype, dv[0].inner, op, params, dout[0], s);
  });
}
int eoc_apply_tape_host(const void* x, const void* y, int dtype, int64_t n, const eoc_tape_op* ops, int nops, void* out,
                        int out_dtype, size_t strip_bytes) {
  if (n < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape_host: negative length");
  if (n == 0) return EOC_OK;
  if (!x || !out) EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape_host: NULL pointer");
  if (out_dtype != EOC_F32 && out_dtype != EOC_F64)
    EOC_FAIL(EOC_ERR_INVALID, "eoc_apply_tape_host: out_dtype %d is neither float32 nor float64", out_dtype);
  const eoc_view vx = flat_view(x, dtype, n), vy = flat_view(y ? y : x, dtype, n);
  if (int rc = check_view(&vx, "eoc_apply_tape_host")) return rc;
  std::vector<HostOut> ho{{out, dtype_size(out_dtype), 1}};
  const eoc_view* ins[2] = {&vx, &vy};
  return strip_pipeline(ins, y ? 2 : 1, ho, strip_bytes, [&](const eoc_view* dv, void* const* dout, cudaStream_t s) {
    return eoc_apply_tape(dv[0].data, y ? dv[1].data : nullptr, dtype, dv[0].inner, ops, nops, dout[0], out_dtype, (void*)s);
  });
}

// Frees the calling thread's staging: two lanes of device strips (in + out), the pinned staging strips of the
// pageable path, the segment tables, two streams.  The *_host entry points keep them between calls (a dask worker
// thread calls them chunk after chunk) until the thread exits or calls this.
int eoc_host_release(void) {
  tl_ctx.release();
  return EOC_OK;
}

}  // extern "C"
