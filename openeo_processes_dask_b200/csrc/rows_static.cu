// eoc_reduce_rows_static -- the long-row (spatial) reduction of eoc_reduce for callers that cannot allocate.
//
// No counterpart in the reference.  Code under CUDA graph capture,
// tight serving loops and harnesses that forbid cudaMalloc cannot provide the workspace eoc_reduce asks for when
// it takes the row-cooperative kernels (k_rows_partial / k_rows_final, rows_kernel.cuh: one RowPartial of 64
// bytes per (row, chunk)).  This entry point runs the SAME two kernels with the partials in a file-scope
// __device__ array of EOC_ROWS_STATIC_BYTES (512 KiB, held by every GPU that loads the library until exit):
//   * it calls no cudaMalloc / cudaFree / allocating library routine -- inputs and outputs are the caller's;
//   * in host code the array's name is only a host-side stand-in for the symbol: the device address is obtained
//     ONCE PER DEVICE, with that device current, by cudaGetSymbolAddress (return code checked) and that pointer
//     is what the kernels get;
//   * requests whose partials do not fit are rejected with EOC_ERR_UNSUPPORTED and a message naming the limit;
//   * nothing has to start at zero: every partial slot k_rows_final reads is written by k_rows_partial of the
//     same call first (slot (o, c) for all o < outer, c < nchunks), so the previous call's contents are dead.
//     Every call on a device shares the array: two calls that OVERLAP on the GPU (different streams, or
//     different host threads without ordering) race on it -- serialise them on one stream.
#include <mutex>
#include <type_traits>

#include "rows_kernel.cuh"

namespace eoc {
namespace {

static __device__ __align__(256) unsigned char g_rows_static_scratch[EOC_ROWS_STATIC_BYTES];

// device address of the scratch array on the CURRENT device, looked up once per device
int rows_static_ptr(void** out) {
  static std::mutex mu;
  static void* cached[64];
  int dev = 0;
  EOC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_rows_static: device ordinal %d outside [0, 64)", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    void* p = nullptr;
    EOC_CUDA(cudaGetSymbolAddress(&p, g_rows_static_scratch));
    if (!p) EOC_FAIL(EOC_ERR_CUDA, "eoc_reduce_rows_static: cudaGetSymbolAddress returned a null address");
    cached[dev] = p;
  }
  *out = cached[dev];
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" int eoc_reduce_rows_static(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof,
                                      void* const* outs, void* stream) {
  if (int rc = check_view(in, "eoc_reduce_rows_static")) return rc;
  if (!outs) EOC_FAIL(EOC_ERR_INVALID, "eoc_reduce_rows_static: outs is NULL");
  if (ops == 0 || (ops >> EOC_NUM_OPS)) EOC_FAIL(EOC_ERR_INVALID, "eoc_reduce_rows_static: bad ops mask 0x%x", ops);
  if (ops & kOrderOps)
    EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_rows_static: first / last / any / all are column-kernel reducers");
  if (in->inner != 1 || in->stride_R != 1)
    EOC_FAIL(EOC_ERR_UNSUPPORTED,
             "eoc_reduce_rows_static: rows of contiguous samples only (inner == 1, stride_R == 1); other views "
             "need no scratch memory: use eoc_reduce");
  if (in->R > 0x7fffffffLL) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_rows_static: R=%lld too large", (long long)in->R);
  OutPtrs o;
  for (int id = 0; id < EOC_NUM_OPS; ++id) {
    o.p[id] = (ops & (1u << id)) ? outs[id] : nullptr;
    if ((ops & (1u << id)) && !o.p[id])
      EOC_FAIL(EOC_ERR_INVALID, "eoc_reduce_rows_static: outs[%d] is NULL but selected by ops mask 0x%x", id, ops);
  }
  if (in->outer == 0) return EOC_OK;
  if (in->R == 0) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_reduce_rows_static: empty rows (R == 0) are the host layer's case");
  const size_t need = rows_workspace_bytes(*in);
  if (need > (size_t)EOC_ROWS_STATIC_BYTES)
    EOC_FAIL(EOC_ERR_UNSUPPORTED,
             "eoc_reduce_rows_static: %lld rows x %d chunks need %zu bytes of partials, the static scratch holds "
             "EOC_ROWS_STATIC_BYTES = %d (use eoc_reduce with a caller-provided workspace)",
             (long long)in->outer, rows_chunks(*in), need, (int)EOC_ROWS_STATIC_BYTES);
  void* scratch = nullptr;
  if (int rc = rows_static_ptr(&scratch)) return rc;
  return launch_rows(*in, ops, ignore_nodata, ddof, o, scratch, (size_t)EOC_ROWS_STATIC_BYTES, (cudaStream_t)stream);
}
