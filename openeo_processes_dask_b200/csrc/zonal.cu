// eoc_rasterize_polygon / eoc_zonal_reduce -- aggregate_spatial (reference cubes/aggregate.py:256-323).
//
// The reference hands the geometries to xvec.zonal_stats(method="iterate"): per geometry a rasterio geometry_mask
// (all_touched=False: a pixel belongs to the polygon when its CENTRE is inside) and a reduce of the where()-masked
// cube over (y, x).  Here:
//   k_rasterize     even-odd crossing test of every pixel centre of the geometry's bounding window against all rings
//                   (holes and multi-polygons are just more rings), float64, the operation order of the numpy oracle
//                   -- no FMA contraction -- so both agree on every pixel, edge cases included;
//   k_zonal_partial one CTA per (plane, chunk of window rows): coalesced mask bytes and samples, RowPartial
//                   accumulators of the rows kernel (rows_kernel.cuh), warp-shuffle + shared-memory combine;
//   k_rows_final    (rows_kernel.cuh, unchanged) folds the chunks of a plane into the selected reducers.
// Partials are float64 and exchangeable: y-strip sharded cubes combine them with one small all_reduce / all_gather
// (sharding.zonal_reduce) -- the one place of this path where a collective sits on the data path.
// Algorithmic bytes per geometry: window pixels * (planes * sizeof(sample) + 1 mask byte).
#include <type_traits>

#include "rows_kernel.cuh"

namespace eoc {
namespace {

__global__ void __launch_bounds__(256) k_rasterize(const double* __restrict__ xs, const double* __restrict__ ys, int64_t W,
                                                   int64_t H, const double* __restrict__ verts,
                                                   const int32_t* __restrict__ ring_offsets, int nrings,
                                                   uint8_t* __restrict__ mask) {
  const int64_t total = W * H;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += step) {
    const double px = xs[e % W], py = ys[e / W];
    bool inside = false;
    for (int r = 0; r < nrings; ++r) {
      const int beg = ring_offsets[r], end = ring_offsets[r + 1];
      for (int k = beg; k < end; ++k) {
        const int k2 = (k + 1 < end) ? k + 1 : beg;
        const double x1 = verts[2 * k], y1 = verts[2 * k + 1], x2 = verts[2 * k2], y2 = verts[2 * k2 + 1];
        if ((y1 > py) != (y2 > py)) {
          const double xi = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn(x2, x1), __dsub_rn(py, y1)), __dsub_rn(y2, y1)), x1);
          if (px < xi) inside = !inside;
        }
      }
    }
    mask[e] = inside ? 1 : 0;
  }
}

template <typename In, uint32_t RG>
__global__ void __launch_bounds__(ROWS_THREADS) k_zonal_partial(const In* __restrict__ data, int64_t plane_stride,
                                                                int64_t row_stride, int64_t H, int64_t W,
                                                                const uint8_t* __restrict__ mask, int64_t mask_row_stride,
                                                                int64_t rows_per_chunk, int nchunks,
                                                                RowPartial* __restrict__ partials) {
  const int64_t p = blockIdx.y;
  const int c = blockIdx.x;
  const In* plane = data + p * plane_stride;
  // shift of the variance sums: the first finite sample of the window's first row (shared by all chunks)
  const double K = (RG & RG_VAR) ? row_shift_cta(plane, W) : 0.0;
  const int64_t y0 = (int64_t)c * rows_per_chunk;
  const int64_t y1 = (y0 + rows_per_chunk < H) ? y0 + rows_per_chunk : H;
  RowPartial acc;
  rp_init(acc);
  for (int64_t y = y0; y < y1; ++y) {
    const In* row = plane + y * row_stride;
    const uint8_t* mrow = mask + y * mask_row_stride;
    // V pixels per thread and step: one 128-bit (or narrower) sample load behind V mask bytes.  The window starts at
    // an arbitrary column of the cube, so the first `lead` pixels up to the next 16-byte boundary and the tail go one
    // by one; with one pixel per thread and step the loop had 4 bytes per thread in flight behind a dependent mask
    // load and ran at 1.0 TB/s (ncu launch list r2d: 75 ms for a 78 GB window).
    constexpr int V = sizeof(In) == 8 ? 2 : 4;
    const int64_t mis = (int64_t)(((uintptr_t)row / sizeof(In)) % V);
    const int64_t lead = mis ? (V - mis < W ? V - mis : W) : 0;
    for (int64_t x = threadIdx.x; x < lead; x += ROWS_THREADS)
      if (mrow[x]) rp_add<RG>(acc, (double)row[x], K);
    const int64_t wv = lead + (W - lead) / V * V;
    // (the samples are loaded whether or not a mask byte is set: a load that waits for the mask halves the bytes in
    // flight, and a polygon rarely covers less than a third of its bounding window)
#pragma unroll 4
    for (int64_t x = lead + (int64_t)threadIdx.x * V; x < wv; x += (int64_t)ROWS_THREADS * V) {
      uint8_t mb[V];
#pragma unroll
      for (int k = 0; k < V; ++k) mb[k] = mrow[x + k];
      const Vec<In, V> q = ld_vec<In, V>(row + x);
#pragma unroll
      for (int k = 0; k < V; ++k)
        if (mb[k]) rp_add<RG>(acc, (double)q.v[k], K);
    }
    for (int64_t x = wv + threadIdx.x; x < W; x += ROWS_THREADS)
      if (mrow[x]) rp_add<RG>(acc, (double)row[x], K);
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) {
    const RowPartial q = rp_shfl_down(acc, d);
    rp_merge(acc, q);
  }
  __shared__ RowPartial warp_part[ROWS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_part[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    RowPartial t = warp_part[0];
    for (int w = 1; w < ROWS_THREADS / 32; ++w) rp_merge(t, warp_part[w]);
    partials[p * nchunks + c] = t;
  }
}

int zonal_chunks(int64_t planes, int64_t H) {
  int64_t want = (8LL * sm_count() + planes - 1) / planes;
  if (want > H) want = H;
  if (want < 1) want = 1;
  if (want > 4096) want = 4096;
  return (int)want;
}

template <typename In>
int launch_zonal_t(const void* data, int64_t planes, int64_t plane_stride, int64_t row_stride, int64_t H, int64_t W,
                   const uint8_t* mask, int64_t mask_row_stride, int64_t full_len, uint32_t ops, int ignore_nodata,
                   int ddof, const OutPtrs& outs, RowPartial* part, int nchunks, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  const int64_t rpc = (H + nchunks - 1) / nchunks;
  const dim3 grid((unsigned)nchunks, (unsigned)planes);
#define EOC_ZONAL_GO(RGV)                                                                                          \
  k_zonal_partial<In, RGV><<<grid, ROWS_THREADS, 0, s>>>((const In*)data, plane_stride, row_stride, H, W, mask,       \
                                                         mask_row_stride, rpc, nchunks, part)
  switch (row_groups_for(ops)) {
    case RG_SUM: EOC_ZONAL_GO(RG_SUM); break;
    case RG_SUM | RG_VAR: EOC_ZONAL_GO(RG_SUM | RG_VAR); break;
    case RG_MM: EOC_ZONAL_GO(RG_MM); break;
    default: EOC_ZONAL_GO(RG_ALL); break;
  }
#undef EOC_ZONAL_GO
  EOC_LAUNCH_CHECK();
  // the finishing kernel of the rows path: `R` is the length the reference reduces over -- the whole where()-masked
  // raster -- which decides ignore_nodata = 0 (any NaN, i.e. any pixel outside the polygon, poisons) and count(True)
  k_rows_final<In, A><<<(unsigned)((planes + 127) / 128), 128, 0, s>>>((const In*)data, plane_stride, planes, full_len,
                                                                       nchunks, part, ops, ignore_nodata, ddof, outs);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

}  // namespace
}  // namespace eoc

using namespace eoc;

extern "C" int eoc_rasterize_polygon(const double* xs, const double* ys, int64_t W, int64_t H, const double* verts,
                                     const int32_t* ring_offsets, int nrings, uint8_t* mask, void* stream) {
  if (W < 0 || H < 0 || nrings < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_rasterize_polygon: negative size");
  if (W * H == 0) return EOC_OK;
  if (!xs || !ys || !mask || (nrings > 0 && (!verts || !ring_offsets)))
    EOC_FAIL(EOC_ERR_INVALID, "eoc_rasterize_polygon: NULL pointer");
  int64_t blocks = (W * H + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 32;
  if (blocks > cap) blocks = cap;
  k_rasterize<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(xs, ys, W, H, verts, ring_offsets, nrings, mask);
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

extern "C" size_t eoc_zonal_reduce_workspace_bytes(int64_t planes, int64_t H) {
  if (planes <= 0 || H <= 0) return 0;
  return sizeof(RowPartial) * (size_t)planes * (size_t)zonal_chunks(planes, H);
}

extern "C" int eoc_zonal_reduce(const void* data, int dtype, int64_t planes, int64_t plane_stride, int64_t row_stride,
                                int64_t H, int64_t W, const uint8_t* mask, int64_t mask_row_stride, int64_t full_len,
                                uint32_t ops, int ignore_nodata, int ddof, void* const* outs, void* workspace,
                                size_t workspace_bytes, void* stream) {
  if (dtype_size(dtype) == 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_zonal_reduce: unknown dtype %d", dtype);
  if (planes < 0 || H < 0 || W < 0) EOC_FAIL(EOC_ERR_INVALID, "eoc_zonal_reduce: negative extent");
  if (!outs || ops == 0 || (ops >> EOC_NUM_OPS) || (ops & kOrderOps))
    EOC_FAIL(EOC_ERR_INVALID, "eoc_zonal_reduce: bad ops mask 0x%x (streaming reducers only)", ops);
  OutPtrs o;
  for (int id = 0; id < EOC_NUM_OPS; ++id) {
    o.p[id] = (ops & (1u << id)) ? outs[id] : nullptr;
    if ((ops & (1u << id)) && !o.p[id]) EOC_FAIL(EOC_ERR_INVALID, "eoc_zonal_reduce: outs[%d] is NULL", id);
  }
  if (planes == 0) return EOC_OK;
  if (planes > 65535) EOC_FAIL(EOC_ERR_UNSUPPORTED, "eoc_zonal_reduce: planes=%lld > 65535", (long long)planes);
  if (H * W == 0 || !data || !mask) EOC_FAIL(EOC_ERR_INVALID, "eoc_zonal_reduce: empty window or NULL data / mask");
  const int nchunks = zonal_chunks(planes, H);
  const size_t need = sizeof(RowPartial) * (size_t)planes * (size_t)nchunks;
  if (!workspace || workspace_bytes < need)
    EOC_FAIL(EOC_ERR_WORKSPACE, "eoc_zonal_reduce needs %zu workspace bytes, got %zu", need, workspace_bytes);
  RowPartial* part = (RowPartial*)workspace;
  cudaStream_t s = (cudaStream_t)stream;
#define EOC_Z(T) launch_zonal_t<T>(data, planes, plane_stride, row_stride, H, W, mask, mask_row_stride, full_len, ops, \
                                   ignore_nodata, ddof, o, part, nchunks, s)
  switch (dtype) {
    case EOC_F32: return EOC_Z(float);
    case EOC_F64: return EOC_Z(double);
    case EOC_U8: return EOC_Z(uint8_t);
    case EOC_I16: return EOC_Z(int16_t);
    case EOC_U16: return EOC_Z(uint16_t);
    case EOC_I32: return EOC_Z(int32_t);
  }
#undef EOC_Z
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", dtype);
}
