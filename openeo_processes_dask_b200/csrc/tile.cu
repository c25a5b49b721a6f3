// Launcher of the TMA tile kernel (tile_kernel.cuh): builds the CUtensorMap of the [outer, R, inner]
// view, picks tile width / stage count from the shared-memory budget and launches a persistent grid.
#include <cstdlib>
#include <mutex>
#include <type_traits>

#include "tile_kernel.cuh"
#include "tma_host.cuh"

namespace eoc {
namespace {

constexpr size_t kSmemPerSm = 228 * 1024;   // shared memory of one SM
constexpr size_t kSmemPerCtaMax = 227 * 1024;
constexpr size_t kCtaReserve = 1024;        // driver-reserved shared memory per resident CTA
constexpr size_t kSlack = 128 + 8 * TILE_MAX_STAGES;  // alignment of the first buffer + mbarriers

struct TilePlan {
  int W = 0, CPT = 1, RB = 0, nbox = 0, stages = 0, ctas = 0;
  size_t tile_bytes = 0;
};

// widest tile that leaves >= 2 stages, preferring the configuration with most resident warps
TilePlan plan_tile(int64_t R, size_t es, int only_w) {
  TilePlan best;
  int best_warps = 0;
  for (int W : {128, 64, 32}) {
    if (only_w && W != only_w) continue;
    if ((size_t)W * es < 16) continue;
    int nbox = (int)((R + 255) / 256);
    int RB = (int)((R + nbox - 1) / nbox);
    // every box must be a multiple of 128 bytes so that stacked boxes stay TMA-aligned
    int row_bytes = (int)(W * es);
    int m = 128;
    while (m > 1 && (row_bytes % m) != 0) m >>= 1;  // m = gcd(128, row_bytes) for power-of-two rows
    const int mult = 128 / m;
    RB = (RB + mult - 1) / mult * mult;
    if (RB > 256) continue;
    const size_t tile = (size_t)nbox * RB * row_bytes;
    if (2 * tile + kSlack > kSmemPerCtaMax) continue;
    int ctas = (int)(kSmemPerSm / (2 * tile + kSlack + kCtaReserve));
    if (ctas > 8) ctas = 8;
    if (ctas < 1) continue;
    const size_t per_cta = kSmemPerSm / ctas - kCtaReserve - kSlack;
    int stages = (int)(per_cta / tile);
    if (stages > TILE_MAX_STAGES) stages = TILE_MAX_STAGES;
    if (per_cta > kSmemPerCtaMax - kSlack) stages = (int)((kSmemPerCtaMax - kSlack) / tile);
    if (stages < 2) continue;
    // score: resident warps first; on a tie two CTAs per SM (their load / reduce phases interleave) beat
    // one, then the wider tile.  Measured on B200, f32: T=100 W=128 x2 CTAs 5990 GB/s, W=64 x4 5850,
    // W=32 4570; T=200 W=64 x2 CTAs 5010 GB/s (two-pass column kernel: 3500).
    const int warps = (ctas * W / 32) * 4 + (ctas >= 2 ? 2 : 0) + (W == 128 ? 1 : 0);
    if (warps > best_warps) {
      best_warps = warps;
      best.CPT = 1;  // measured on B200 (T=100 f32): 1 column/thread 5990 GB/s, 2 columns/thread 5455 GB/s
      best.W = W;
      best.RB = RB;
      best.nbox = nbox;
      best.stages = stages;
      best.ctas = ctas;
      best.tile_bytes = tile;
    }
  }
  return best;
}

template <typename In>
int launch_tile_t(const eoc_view& in, const ColumnParams& P, const TilePlan& plan, cudaStream_t s) {
  using A = typename std::conditional<std::is_same<In, float>::value, float, double>::type;
  CUtensorMap map;
  if (int rc = tma_make_map(in, plan.W, plan.RB, &map)) return rc;

  TileParams T;
  T.C = P;
  T.W = plan.W;
  T.RB = plan.RB;
  T.nbox = plan.nbox;
  T.stages = plan.stages;
  T.tile_bytes = (uint32_t)plan.tile_bytes;
  T.tiles_per_outer = (in.inner + plan.W - 1) / plan.W;
  T.num_tiles = T.tiles_per_outer * in.outer;
  const size_t smem = (size_t)plan.stages * plan.tile_bytes + kSlack;
  int64_t grid = (int64_t)sm_count() * plan.ctas;
  if (grid > T.num_tiles) grid = T.num_tiles;

  const bool rare = P.ops & ((1u << EOC_OP_MIN) | (1u << EOC_OP_MAX) | (1u << EOC_OP_PRODUCT) | (1u << EOC_OP_COUNT));
#define EOC_TILE_LAUNCH(GG, WW, CC)                                                                      \
  do {                                                                                                   \
    auto kern = k_tile<In, A, GG, WW, CC>;                                                               \
    EOC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
    kern<<<(unsigned)grid, WW / CC, smem, s>>>(map, T);                                                  \
  } while (0)
#define EOC_TILE_SHAPE(GG)                                        \
  do {                                                            \
    if (plan.W == 128 && plan.CPT == 2) EOC_TILE_LAUNCH(GG, 128, 2); \
    else if (plan.W == 128) EOC_TILE_LAUNCH(GG, 128, 1);          \
    else if (plan.W == 64 && plan.CPT == 2) EOC_TILE_LAUNCH(GG, 64, 2); \
    else if (plan.W == 64) EOC_TILE_LAUNCH(GG, 64, 1);            \
    else EOC_TILE_LAUNCH(GG, 32, 1);                              \
  } while (0)
  if (rare)
    EOC_TILE_SHAPE(G_MINMAX | G_RARE);
  else
    EOC_TILE_SHAPE(0u);
#undef EOC_TILE_SHAPE
#undef EOC_TILE_LAUNCH
  EOC_LAUNCH_CHECK();
  return EOC_OK;
}

// EOC_TILE=0 disables the tile kernel, EOC_TILE=1 forces it whenever a plan exists (tests / tuning)
int tile_env() {
  const char* e = getenv("EOC_TILE");
  if (!e || !e[0]) return -1;
  return e[0] == '0' ? 0 : 1;
}

}  // namespace

// Returns EOC_OK with *handled = true when the tile kernel was launched, *handled = false when the
// view does not qualify (the caller then uses the column kernel), or an error.
int launch_tile(const eoc_view& in, const ColumnParams& P, cudaStream_t s, bool* handled) {
  *handled = false;
  const int env = tile_env();
  if (env == 0) return EOC_OK;
  const size_t es = dtype_size(in.dtype);
  if (in.R < 2 || in.R > 0x7fffffffLL || in.outer * in.inner == 0) return EOC_OK;
  if (!tma_view_ok(in)) return EOC_OK;
  if (P.members && in.R > 0x7fffff) return EOC_OK;
  int only_w = 0;
  if (const char* e = getenv("EOC_TILE_W")) only_w = atoi(e);  // tuning knob: tile width
  TilePlan plan = plan_tile(in.R, es, only_w);
  if (plan.W == 0) return EOC_OK;
  if (const char* e = getenv("EOC_TILE_CPT"))  // tuning knob: columns per thread
    if (e[0] == '1' || (e[0] == '2' && plan.W >= 64)) plan.CPT = e[0] - '0';
  // too few resident columns cannot keep up with HBM: the two-pass column kernel is the better choice
  if (env != 1 && plan.ctas * plan.W < 128) return EOC_OK;
  // narrow cubes: not worth a persistent grid
  if (env != 1 && in.outer * in.inner < 4096) return EOC_OK;
  if (env != 1 && !tma_tiles_fill(in, 128)) return EOC_OK;  // short `inner`: mostly empty tiles, the column kernel is ahead
  *handled = true;
  switch (in.dtype) {
    case EOC_F32: return launch_tile_t<float>(in, P, plan, s);
    case EOC_F64: return launch_tile_t<double>(in, P, plan, s);
    case EOC_U8: return launch_tile_t<uint8_t>(in, P, plan, s);
    case EOC_I16: return launch_tile_t<int16_t>(in, P, plan, s);
    case EOC_U16: return launch_tile_t<uint16_t>(in, P, plan, s);
    case EOC_I32: return launch_tile_t<int32_t>(in, P, plan, s);
  }
  *handled = false;
  EOC_FAIL(EOC_ERR_INVALID, "unknown dtype %d", in.dtype);
}

}  // namespace eoc
