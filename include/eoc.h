/*
 * eoc.h -- C ABI of libeoc_sm100.so: B200-native reducers for openEO datacubes.
 *
 * This is the drop-in boundary for ONE path of Open-EO/openeo-processes-dask:
 * the array-level reducers that reduce_dimension / aggregate_temporal call over
 * the t / bands dimension, plus normalized_difference feeding such a reduce.
 * Each entry point names the reference code it replaces (paths relative to
 * openeo_processes_dask/process_implementations/ in the reference repo).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types cross this boundary
 *   - every function returns 0 on success or a negative eoc_status; the message
 *     of the last failure on the calling thread is eoc_last_error()
 *   - device entry points are asynchronous on the caller's cudaStream_t (passed
 *     as void*; NULL = legacy default stream), allocate nothing, and are
 *     re-entrant: any number of host threads may call concurrently
 *   - a cube axis reduction is described as a strided 3-D view
 *         [outer, R, inner]   element (o, r, i) at data[o*stride_outer + r*stride_R + i]
 *     with R the reduced axis and `inner` contiguous:
 *         t  of (t,b,y,x): outer=1,   R=T, inner=B*Y*X
 *         t  of (b,t,y,x): outer=B,   R=T, inner=Y*X
 *         b  of (t,b,y,x): outer=T,   R=B, inner=Y*X
 *         t  of (x,y,t,b): outer=X*Y, R=T, inner=B      (the reference's test fixture layout)
 *     Outputs are dense [outer, inner] (reducers), [outer, nq, inner] (quantiles, the
 *     q-axis lands where the reduced axis was -- math.py:384-385) or
 *     [outer, nseg, inner] (segmented reduce).
 *   - NaN is nodata (docs/decisions/handle-nodata-in-rastercubes.md in the reference)
 */
#ifndef EOC_H_
#define EOC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EOC_ABI_VERSION 3

typedef enum eoc_status {
  EOC_OK = 0,
  EOC_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, unknown dtype/op) */
  EOC_ERR_UNSUPPORTED = -2, /* valid request outside what the kernels implement (e.g. R too large) */
  EOC_ERR_CUDA = -3,        /* a CUDA runtime call failed; message has the CUDA error string */
  EOC_ERR_WORKSPACE = -4    /* workspace too small; query eoc_*_workspace_bytes */
} eoc_status;

typedef enum eoc_dtype {
  EOC_F32 = 0,
  EOC_F64 = 1,
  EOC_U8 = 2,
  EOC_I16 = 3,
  EOC_U16 = 4,
  EOC_I32 = 5
} eoc_dtype;

/* Reducer ids.  Bit (1u << id) in an ops mask selects the reducer; outs[id] receives it. */
typedef enum eoc_op {
  EOC_OP_SUM = 0,      /* math.py:390-398  _sum      np.nansum / np.sum   (all-NaN -> 0) */
  EOC_OP_MEAN = 1,     /* math.py:191-195  mean      np.nanmean / np.mean               */
  EOC_OP_MIN = 2,      /* math.py:118-122  _min      np.nanmin / np.min                 */
  EOC_OP_MAX = 3,      /* math.py:177-181  _max      np.nanmax / np.max                 */
  EOC_OP_COUNT = 4,    /* arrays.py:587-597 count(condition=None): not-NaN and not-inf; int64 */
  EOC_OP_VARIANCE = 5, /* math.py:205-209  variance  np.nanvar / np.var, ddof=1         */
  EOC_OP_SD = 6,       /* math.py:198-202  sd        np.nanstd / np.std, ddof=1         */
  EOC_OP_PRODUCT = 7,  /* math.py:401-409  product   np.nanprod / np.prod (all-NaN -> 1) */
  EOC_OP_COUNT_ALL = 8, /* arrays.py:598-599 count(condition=True): every element; value dtype */
  EOC_OP_NOT_NAN = 9,   /* number of not-NaN samples, int64: the n of mean/variance; lets partial
                           results of a sharded spatial reduce be merged (sum, n) / (n, mean, M2) */
  /* the remaining reducer-category processes */
  EOC_OP_FIRST = 10,    /* arrays.py:461-485 first: first not-NaN sample (ignore_nodata=0: sample 0)      */
  EOC_OP_LAST = 11,     /* arrays.py:488-495 last:  last not-NaN sample  (ignore_nodata=0: sample R-1)    */
  EOC_OP_ANY = 12,      /* logic.py:58-71 _any: 1 if a not-NaN sample is non-zero, else 0;
                           ignore_nodata=0: NaN when the column holds a NaN and no non-zero sample       */
  EOC_OP_ALL = 13       /* logic.py:74-87 _all: 1 if no sample equals zero (NaN is truthy like np.all), else 0;
                           ignore_nodata=0: NaN when the column holds a NaN and every sample is truthy    */
} eoc_op;
#define EOC_NUM_OPS 14

/* Value outputs use the compute dtype: EOC_F32 for EOC_F32 input, EOC_F64 for every other input
 * dtype (numpy promotes integer means/variances to float64; integer sums/extrema are exact in
 * float64 below 2^53 and are cast back by the host layer).  EOC_OP_COUNT and EOC_OP_NOT_NAN are
 * int64.  EOC_OP_ANY / EOC_OP_ALL are 0 / 1 / NaN in the compute dtype (the host layer turns them into the
 * bool or float64 array the reference returns). */

typedef struct eoc_view {
  const void* data;     /* device pointer (host pointer for the *_host entry points) */
  int32_t dtype;        /* eoc_dtype */
  int64_t outer, R, inner;
  int64_t stride_outer; /* in elements */
  int64_t stride_R;     /* in elements; inner stride is 1 */
} eoc_view;

int eoc_abi_version(void);
const char* eoc_last_error(void);

/* Device properties the host layer needs to size strips: SM count, total and free HBM bytes. */
int eoc_device_info(int32_t* sm_count, int64_t* total_bytes, int64_t* free_bytes);

/*
 * eoc_reduce -- streaming multi-output reduction over R.
 * Replaces the numpy calls of math.py:118-122,177-181,191-209,390-409 and arrays.py:595-599
 * when invoked by DataArray.reduce from cubes/reduce.py:26-34.
 *   ops            bit mask of eoc_op
 *   ignore_nodata  1: NaN samples are skipped (np.nan* functions); 0: any NaN -> NaN result
 *   ddof           delta degrees of freedom for variance/sd (the reference uses 1)
 *   outs[id]       device pointer, dense [outer, inner]; only entries selected by ops are read
 *   workspace      device scratch of at least eoc_reduce_workspace_bytes(in, ops) bytes (may be
 *                  NULL when that is 0)
 */
size_t eoc_reduce_workspace_bytes(const eoc_view* in, uint32_t ops);
int eoc_reduce(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
               void* workspace, size_t workspace_bytes, void* stream);

/*
 * eoc_select -- per-column order statistics over R (R <= EOC_SELECT_MAX_R; float32 device views that TMA can
 * address are served up to R = 1024, anything else beyond the limit is EOC_ERR_UNSUPPORTED: use eoc_select_general).
 * Replaces np.nanmedian / np.median (math.py:184-188) and np.nanquantile / np.quantile with
 * method="linear" (math.py:341-387).
 *   median_out   dense [outer, inner] in the compute dtype, or NULL
 *   nq, probs    number of probabilities and the HOST array of them (float64, as math.py:355-365)
 *   quant_out    dense [outer, nq, inner], or NULL when nq == 0
 *   quant_dtype  EOC_F64 (what numpy returns for float32 data and float64 probabilities) or EOC_F32
 */
#define EOC_SELECT_MAX_R 512
#define EOC_SELECT_MAX_Q 16
int eoc_select(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
               void* quant_out, int quant_dtype, void* stream);

/*
 * eoc_select_fused -- eoc_select plus streaming reducers of the SAME column in the same pass: the
 * samples are in registers for the sort anyway, so mean / sd / variance / min / max / sum / count
 * cost no extra HBM traffic (BASELINE config 3: quantiles([0.1,0.5,0.9]) + sd over t reads the cube
 * once instead of twice).  In the reference these are sibling reduce_dimension nodes, each a full
 * numpy pass (math.py:198-202 next to math.py:341-387).
 *   ops, ddof, outs   as in eoc_reduce (outs[id] dense [outer, inner]); ops == 0 is plain eoc_select
 */
int eoc_select_fused(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                     void* quant_out, int quant_dtype, uint32_t ops, int ddof, void* const* outs,
                     void* stream);

/*
 * eoc_select_general -- eoc_select / eoc_sort for columns of ANY length (R > EOC_SELECT_MAX_R; R <= 2^24): the
 * slower general path, since np.nanmedian / np.nanquantile / np.sort (math.py:184-188, 341-387,
 * arrays.py:560-584) take any length.  Columns are sorted in place in the caller's workspace (Batcher merge
 * exchange, NaN last) in batches of as many columns as fit; a workspace of about 64 MB keeps a batch inside L2.
 *   median_out, nq, probs, quant_out, quant_dtype   as in eoc_select
 *   sort_out, sort_descending, sort_nan_first       as in eoc_sort (sort_out may be NULL)
 *   workspace   device scratch, at least R * (4 for EOC_F32, else 8) bytes (one column);
 *               eoc_select_general_workspace_bytes(in) is the recommended size
 */
size_t eoc_select_general_workspace_bytes(const eoc_view* in);
int eoc_select_general(const eoc_view* in, int ignore_nodata, void* median_out, int nq, const double* probs,
                       void* quant_out, int quant_dtype, void* sort_out, int sort_descending, int sort_nan_first,
                       void* workspace, size_t workspace_bytes, void* stream);

/*
 * eoc_segmented_reduce -- reduce groups of R-indices ("segments") independently.
 * Replaces data.groupby_bins(t, bins).reduce(reducer) of cubes/aggregate.py:92-96 (a Python loop
 * over groups with a gather copy per group).  Segment s owns members[seg_offsets[s] ..
 * seg_offsets[s+1]) (indices into R, in the order they are reduced); an empty segment yields NaN
 * for every value reducer and 0 for EOC_OP_COUNT (xarray re-indexes missing groups to NaN).
 *   seg_offsets  device int32[nseg+1];  members  device int32[seg_offsets[nseg]]
 *   outs[id]     dense [outer, nseg, inner]
 */
int eoc_segmented_reduce(const eoc_view* in, int nseg, const int32_t* seg_offsets,
                         const int32_t* members, int max_seg_len, uint32_t ops, int ignore_nodata,
                         int ddof, void* const* outs, void* stream);

/*
 * eoc_nd_reduce -- normalized difference (a-b)/(a+b) fused into a reduction over R.
 * Replaces cubes/indices.py:49-52 (two band selections + math.py:412-414) followed by a
 * reduce_dimension over t.  a and b must have identical outer/R/inner (strides may differ).
 *   integer_wrap  how integer samples are treated (ignored for float inputs):
 *                 0  converted to the compute dtype first (float32 for u8/i16/u16 -- the "uint16 -> fp32"
 *                    cast a workflow applies before band math; float64 for i32), then float arithmetic;
 *                 1  numpy's arithmetic on the raw integers, i.e. what math.py:412-414 returns for an integer
 *                    cube: a - b and a + b WRAP in the sample dtype (uint16 1000, 2000 -> 64536 / 3000),
 *                    the true division is float64; compute dtype float64
 *   nd_out   optional dense [outer, R, inner] elementwise result in the compute dtype (NULL to skip);
 *            with ops == 0 this is the plain normalized_difference process.
 */
int eoc_nd_reduce(const eoc_view* a, const eoc_view* b, uint32_t ops, int ignore_nodata, int ddof,
                  void* const* outs, void* nd_out, int integer_wrap, void* stream);

/*
 * eoc_masked_reduce / eoc_masked_select / eoc_mask_canonical -- mask(data, mask, replacement) fused into
 * the reduce that follows it (cloud mask -> NaN -> temporal composite).
 * Replaces data.where(_not(mask), replacement) of cubes/mask.py:112 -- which materialises the masked cube
 * (one read + one write of everything) -- followed by the reducer's own read (cubes/reduce.py:26-34).
 *   mask         EOC_U8 view with the SAME outer / R / inner as `in` and its own strides; a stride of 0
 *                broadcasts (a (t,y,x) mask over the bands of a (t,bands,y,x) cube: stride_outer = 0 in the
 *                view [outer=bands, R=t, inner=y*x]).  A non-zero byte means "replace" -- the samples where
 *                the reference's mask is truthy and not NaN; eoc_mask_canonical builds such bytes (0 / 1)
 *                from a mask of any supported dtype (dense array of n elements).
 *   replacement  the value masked samples read as (the reference's default None -> NaN)
 *   masked_out   eoc_masked_reduce only: optional dense [outer, R, inner] output of the masked samples in
 *                the compute dtype -- with ops == 0 this is the plain `mask` process
 * eoc_masked_select covers float32 data, 32 < R <= 128, replacement = NaN and 4-byte aligned mask rows
 * (the composite case); anything else returns EOC_ERR_UNSUPPORTED and the caller materialises the masked
 * cube with eoc_masked_reduce(ops = 0) first.  The mask bytes must be 0 or 1 there.
 */
int eoc_masked_reduce(const eoc_view* in, const eoc_view* mask, double replacement, uint32_t ops,
                      int ignore_nodata, int ddof, void* const* outs, void* masked_out, void* stream);
int eoc_masked_select(const eoc_view* in, const eoc_view* mask, double replacement, int ignore_nodata,
                      void* median_out, int nq, const double* probs, void* quant_out, int quant_dtype,
                      void* stream);
int eoc_mask_canonical(const void* src, int dtype, int64_t n, uint8_t* dst, void* stream);

/*
 * eoc_rasterize_polygon / eoc_zonal_reduce -- aggregate_spatial (cubes/aggregate.py:256-323: xvec.zonal_stats with
 * method="iterate": per geometry a rasterio geometry_mask(all_touched=False) -- a pixel belongs to the polygon when
 * its CENTRE is inside -- and a reduce of the where()-masked cube over (y, x)).
 * eoc_rasterize_polygon: mask[y * W + x] = 1 where the point (xs[x], ys[y]) is inside by the even-odd rule.
 *   xs, ys         device float64 pixel-centre coordinates of the window (W and H of them)
 *   verts          device float64 (x, y) pairs of all rings one after the other (exterior rings, holes, the parts of a
 *                  multi-polygon: even-odd treats them alike); ring r owns vertices ring_offsets[r] .. ring_offsets[r+1]
 *                  (device int32[nrings + 1]); a ring may or may not repeat its first vertex
 * eoc_zonal_reduce: the streaming reducers of the samples under the mask, per plane.
 *   data           device pointer to the window's first sample; sample (p, y, x) at data[p*plane_stride + y*row_stride + x]
 *   mask           device uint8, row y at mask + y * mask_row_stride; non-zero = inside
 *   full_len       the number of samples the REFERENCE reduces over per plane: the whole raster (pixels outside the
 *                  polygon read NaN there), which decides ignore_nodata = 0 and count(condition=True)
 *   outs[id]       device arrays of `planes` values (dtypes as in eoc_reduce); float64 partial sums, shifted sums for
 *                  the variance: rtol 1e-6 against numpy's pairwise summation, not bit-exact
 *   workspace      eoc_zonal_reduce_workspace_bytes(planes, H) bytes of device scratch
 */
int eoc_rasterize_polygon(const double* xs, const double* ys, int64_t W, int64_t H, const double* verts,
                          const int32_t* ring_offsets, int nrings, uint8_t* mask, void* stream);
size_t eoc_zonal_reduce_workspace_bytes(int64_t planes, int64_t H);
int eoc_zonal_reduce(const void* data, int dtype, int64_t planes, int64_t plane_stride, int64_t row_stride, int64_t H,
                     int64_t W, const uint8_t* mask, int64_t mask_row_stride, int64_t full_len, uint32_t ops,
                     int ignore_nodata, int ddof, void* const* outs, void* workspace, size_t workspace_bytes, void* stream);

/*
 * eoc_scan -- cumulative sum / product / minimum / maximum along R, the "axis kept" array processes that
 * apply_dimension runs.  Replaces math.py:125-174: np.nancumsum / np.nancumprod /
 * np.minimum.accumulate / np.maximum.accumulate on a NaN-filled copy, then result[nan_mask] = nan.
 *   op             eoc_scan_op
 *   ignore_nodata  1: a NaN slot contributes the neutral element (0, 1, +inf, -inf) and reads NaN in the
 *                  result; 0: NaN propagates to every later position (np.cumsum / np.minimum.accumulate)
 *   out            dense [outer, R, inner] in the compute dtype (float32 for float32 input, else float64)
 */
typedef enum eoc_scan_op { EOC_SCAN_SUM = 0, EOC_SCAN_PRODUCT = 1, EOC_SCAN_MIN = 2, EOC_SCAN_MAX = 3 } eoc_scan_op;
int eoc_scan(const eoc_view* in, int op, int ignore_nodata, void* out, void* stream);

/*
 * eoc_sort -- every column sorted along R (R <= EOC_SELECT_MAX_R), the "axis kept" form of the selection
 * kernels: same networks, the whole parked column is written back instead of a few ranks.
 * Replaces np.sort(data, axis) / -np.sort(-data, axis) of arrays.py:560-584 (sort with nodata = True / False).
 *   descending  0: valid samples ascending; 1: descending
 *   nan_first   0: the NaNs of a column follow its valid samples (numpy's order; nodata = True);
 *               1: they come first (nodata = False)
 *   out         dense [outer, R, inner] in the compute dtype
 */
int eoc_sort(const eoc_view* in, int descending, int nan_first, void* out, void* stream);

/*
 * eoc_order -- the permutation that sorts every column along R ascending: np.argsort(data, kind="mergesort", axis)
 * of `order` (arrays.py:498-534).  Stable: equal samples (and -0.0 / +0.0) keep their index order; the NaNs come
 * last, in index order.  The host layer applies the reference's asc / nodata post-processing to it.
 *   out   dense int64 [outer, R, inner]; R <= EOC_SELECT_MAX_R
 * eoc_take -- out[o, j, i] = in[o, index[j], i]: np.take(data, order, axis) of `rearrange` (arrays.py:537-555).
 *   index   DEVICE array of n_index int64 positions along R (negative values count from the end, like numpy; the
 *           host layer checks the bounds); out dense [outer, n_index, inner] in the INPUT dtype
 */
int eoc_order(const eoc_view* in, void* out, void* stream);
int eoc_take(const eoc_view* in, const int64_t* index, int64_t n_index, void* out, void* stream);

/*
 * eoc_interpolate_linear -- array_interpolate_linear along R (arrays.py:405-458): runs of non-finite
 * samples inside a column are replaced by the line between their finite neighbours (np.interp in float64),
 * runs at either end become NaN, columns with fewer than two finite samples are returned unchanged.
 * Columns too long for a shared-memory tile (R > ~420 float32 / ~210 float64 samples) take a slower variant that
 * works in the output buffer (three sweeps instead of one); R < 2^22.
 *   x    device array of R float64 positions of the samples along R, or NULL for 0, 1, ..., R-1 (what the
 *        reference uses when apply_dimension calls it: it passes no labels)
 *   out  dense [outer, R, inner] in the compute dtype; float32 / float64 input only
 */
int eoc_interpolate_linear(const eoc_view* in, const double* x, void* out, void* stream);

/*
 * eoc_elementwise -- elementwise processes `apply` runs over a dense array of n samples (cubes/apply.py:24-39).
 *   EOC_ELEM_CLIP                params = {min, max}                                  math.py:332-338  np.clip
 *   EOC_ELEM_LINEAR_SCALE_RANGE  params = {inputMin, inputMax, outputMin, outputMax}  math.py:292-298
 * params is a HOST array of 4 doubles; out is dense, in the compute dtype (float32 for float32 input, else
 * float64); NaN passes through.
 *   EOC_ELEM_ABSOLUTE / SGN / SQRT / FLOOR / CEIL / TRUNC   (params unused)              math.py:212-222, 305-315
 *                                np.abs / np.sign / np.sqrt / np.floor / np.ceil / np.trunc in the compute dtype
 */
typedef enum eoc_elem_op {
  EOC_ELEM_CLIP = 0, EOC_ELEM_LINEAR_SCALE_RANGE = 1, EOC_ELEM_ABSOLUTE = 2, EOC_ELEM_SGN = 3, EOC_ELEM_SQRT = 4,
  EOC_ELEM_FLOOR = 5, EOC_ELEM_CEIL = 6, EOC_ELEM_TRUNC = 7
} eoc_elem_op;
int eoc_elementwise(const void* in, int dtype, int64_t n, int op, const double* params, void* out, void* stream);

/*
 * eoc_apply_tape -- a CHAIN of elementwise processes over dense arrays in one pass: what `apply` runs as a child
 * process graph, node by node, each node a full read and write of the cube (cubes/apply.py:24-39 with math.py:98-115,
 * 212-338).  The chain is an accumulator program of at most EOC_TAPE_MAX_OPS steps evaluated per sample:
 *   EOC_TAPE_LOAD             acc <- operand          EOC_TAPE_STORE_T0 / _T1   t0 / t1 <- acc
 *   binary codes              acc <- acc OP operand   (R-variants: operand OP acc)
 *   unary codes               acc <- f(acc)           (EOC_TAPE_ROUND / ROUND_NEG: imm = 10^|decimals|)
 * operand: 0 = imm, 1 = x, 2 = y, 3 = t0, 4 = t1.  acc starts as x; the result is acc after the last step.
 * Compute dtype = out_dtype: EOC_F64 for any input, EOC_F32 for float32 / uint8 / int16 / uint16 input (samples float32
 * holds exactly); every step rounds to it (numpy's operation by operation arithmetic on an array of that dtype,
 * immediates being weak python scalars).  + - * / sqrt abs sign floor ceil trunc round min max are exactly
 * numpy's; exp, ln, trigonometric / hyperbolic functions, power, arctan2 agree within 1e-6 relative (libm).
 *   x, y   dense device arrays of n samples of dtype `dtype` (y may be NULL when no step reads it)
 *   ops    HOST array of nops steps;  out  dense device array of n values in the compute dtype
 */
#define EOC_TAPE_MAX_OPS 24
typedef enum eoc_tape_code {
  EOC_TAPE_LOAD = 0, EOC_TAPE_STORE_T0 = 1, EOC_TAPE_STORE_T1 = 2,
  EOC_TAPE_ADD = 3, EOC_TAPE_SUB = 4, EOC_TAPE_RSUB = 5, EOC_TAPE_MUL = 6, EOC_TAPE_DIV = 7, EOC_TAPE_RDIV = 8,
  EOC_TAPE_MAXNUM = 9, EOC_TAPE_MINNUM = 10, EOC_TAPE_POW = 11, EOC_TAPE_RPOW = 12, EOC_TAPE_MOD = 13, EOC_TAPE_RMOD = 14,
  EOC_TAPE_ATAN2 = 15, EOC_TAPE_RATAN2 = 16,
  EOC_TAPE_ABS = 17, EOC_TAPE_SGN = 18, EOC_TAPE_SQRT = 19, EOC_TAPE_FLOOR = 20, EOC_TAPE_CEIL = 21, EOC_TAPE_TRUNC = 22,
  EOC_TAPE_ROUND = 23, EOC_TAPE_ROUND_NEG = 24, EOC_TAPE_EXP = 25, EOC_TAPE_LN = 26, EOC_TAPE_SIN = 27, EOC_TAPE_COS = 28,
  EOC_TAPE_TAN = 29, EOC_TAPE_ASIN = 30, EOC_TAPE_ACOS = 31, EOC_TAPE_ATAN = 32, EOC_TAPE_SINH = 33, EOC_TAPE_COSH = 34,
  EOC_TAPE_TANH = 35, EOC_TAPE_ASINH = 36, EOC_TAPE_ACOSH = 37, EOC_TAPE_ATANH = 38
} eoc_tape_code;
typedef struct eoc_tape_op {
  int32_t code;     /* eoc_tape_code */
  int32_t operand;  /* 0 imm, 1 x, 2 y, 3 t0, 4 t1 (ignored by unary codes and stores) */
  double imm;
} eoc_tape_op;
int eoc_apply_tape(const void* x, const void* y, int dtype, int64_t n, const eoc_tape_op* ops, int nops, void* out,
                   int out_dtype, void* stream);

/*
 * eoc_binary -- binary elementwise processes with numpy broadcasting: the overlap resolvers of merge_cubes
 * (cubes/merge.py:54-295: a process of x and y reduces the 2-long "cubes" dimension) and the arithmetic /
 * comparison / logic processes of apply chains.
 *   arithmetic  EOC_BIN_ADD / SUBTRACT / MULTIPLY / DIVIDE   math.py:98-115   out: the operands' dtype (integers
 *               wrap like numpy's), float64 for the true division of integers
 *   comparison  EOC_BIN_EQ / NEQ / GT / GTE / LT / LTE       comparison.py:45-127  out: float64 1 / 0, NaN where
 *               either operand is NaN (np.where(notnull(x) & notnull(y), x OP y, nan)); GTE is (x - y) >= 0
 *   logic       EOC_BIN_AND / OR / XOR                        logic.py:11-40   out: float64 1 / 0 / NaN with the
 *               reference's three-valued rules (and(NaN, 0) = 0, or(NaN, 1) = 1, xor(NaN, .) = NaN)
 * Both operands have the dtype `dtype` (the host layer promotes).  shape[ndim] is the broadcast result shape
 * (ndim <= EOC_BINARY_MAX_DIMS), a_strides / b_strides the operands' strides in ELEMENTS along it, 0 where
 * the operand is broadcast.  scalar_side 1 / 2: x / y is the host value `scalar` (its pointer and strides
 * are ignored), 0: both are arrays.  out is dense C order.
 */
#define EOC_BINARY_MAX_DIMS 8
typedef enum eoc_binary_op {
  EOC_BIN_ADD = 0, EOC_BIN_SUBTRACT = 1, EOC_BIN_MULTIPLY = 2, EOC_BIN_DIVIDE = 3,
  EOC_BIN_EQ = 4, EOC_BIN_NEQ = 5, EOC_BIN_GT = 6, EOC_BIN_GTE = 7, EOC_BIN_LT = 8, EOC_BIN_LTE = 9,
  EOC_BIN_AND = 10, EOC_BIN_OR = 11, EOC_BIN_XOR = 12
} eoc_binary_op;
int eoc_binary(const void* a, const void* b, int dtype, int ndim, const int64_t* shape, const int64_t* a_strides,
               const int64_t* b_strides, int op, double scalar, int scalar_side, void* out, void* stream);

/*
 * eoc_synth_fill -- deterministic synthetic samples (bench / test input generator).
 * out[p * strip_len + j] = f(seed, p * plane_stride + first_index + j) for p < planes, j < strip_len,
 * i.e. a y-strip of a (planes, Y*X) cube regenerates exactly the numbers of the full cube.
 * f: h = splitmix64(seed*0x9E3779B97F4A7C15 + idx); NaN iff (h & 0xFFFFFF) < nan_threshold (float
 * dtypes); value (h >> 40) % 10000.
 */
int eoc_synth_fill(void* out, int dtype, int64_t planes, int64_t plane_stride, int64_t first_index,
                   int64_t strip_len, uint64_t seed, uint32_t nan_threshold, void* stream);

/*
 * Host-buffer entry points: same semantics as eoc_reduce / eoc_select with HOST pointers in the
 * view and in outs.  The cube is cut into strips along `inner`, each strip is copied to the
 * device (cudaMemcpy2DAsync from the caller's buffer; pinned buffers copy at PCIe speed), reduced,
 * and its result copied back, with copy and compute overlapped on two streams.  This is the
 * device-array bridge a dask chunk goes through (numpy in -> numpy out).
 *   strip_bytes  device staging budget per strip (0 = default 256 MiB); two strips are in flight
 */
int eoc_reduce_host(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                    size_t strip_bytes);
int eoc_select_host(const eoc_view* in, int ignore_nodata, void* median_out, int nq,
                    const double* probs, void* quant_out, int quant_dtype, size_t strip_bytes);
int eoc_select_fused_host(const eoc_view* in, int ignore_nodata, void* median_out, int nq,
                          const double* probs, void* quant_out, int quant_dtype, uint32_t ops, int ddof,
                          void* const* outs, size_t strip_bytes);
/* The same strip pipeline for the other array processes (round 2): HOST pointers in the views, in seg_offsets /
 * members and in every output; semantics of eoc_segmented_reduce / eoc_nd_reduce / eoc_scan / eoc_sort.  This is
 * what aggregate_temporal, ndvi, cumsum ... of a numpy / dask chunk go through. */
int eoc_segmented_reduce_host(const eoc_view* in, int nseg, const int32_t* seg_offsets, const int32_t* members,
                              int max_seg_len, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                              size_t strip_bytes);
int eoc_nd_reduce_host(const eoc_view* a, const eoc_view* b, uint32_t ops, int ignore_nodata, int ddof,
                       void* const* outs, void* nd_out, int integer_wrap, size_t strip_bytes);
int eoc_scan_host(const eoc_view* in, int op, int ignore_nodata, void* out, size_t strip_bytes);
int eoc_sort_host(const eoc_view* in, int descending, int nan_first, void* out, size_t strip_bytes);
/* array_interpolate_linear / elementwise processes / recorded elementwise chains of a numpy chunk (arrays.py:405-458,
 * math.py:212-338, cubes/apply.py:24-39): the same strip pipeline; `x` (positions) and `params` are HOST arrays, `out`
 * a dense host array of the dtype the device entry point of the same name writes. */
int eoc_interpolate_linear_host(const eoc_view* in, const double* x, void* out, size_t strip_bytes);
int eoc_elementwise_host(const void* in, int dtype, int64_t n, int op, const double* params, void* out, size_t strip_bytes);
int eoc_apply_tape_host(const void* x, const void* y, int dtype, int64_t n, const eoc_tape_op* ops, int nops, void* out,
                        int out_dtype, size_t strip_bytes);
/* The *_host entry points keep, PER CALLING THREAD and until that thread exits, two lanes of device staging
 * (2 x (strip of inputs + strip of results), 256 MiB strips by default), for pageable input also two pinned host
 * staging strips (64 MiB each), two streams and a small table buffer -- the "allocates nothing" promise above holds
 * for the DEVICE entry points only.  eoc_host_release() frees the calling thread's staging (a dask worker pool of
 * 32 threads would otherwise hold 32 sets). */
int eoc_host_release(void);

/*
 * eoc_reduce_rows_static -- eoc_reduce for rows of contiguous samples (inner == 1, stride_R == 1: the spatial
 * reduction, the one case in which eoc_reduce needs a workspace) WITHOUT any allocation by anybody.
 * It has no counterpart in the reference; it was added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The per-(row, chunk) partials live in a file-scope __device__ array of EOC_ROWS_STATIC_BYTES inside the library
 * (every GPU that loads the library holds it until exit); the entry point calls no cudaMalloc / cudaFree and no
 * allocating library routine.  Same kernels, chunking and results as eoc_reduce with a caller workspace.
 *   limit     eoc_reduce_workspace_bytes(in, ops) <= EOC_ROWS_STATIC_BYTES, i.e. outer * chunks <= 8192 (about
 *             7000 rows on a B200); larger requests, other view shapes and first / last / any / all return
 *             EOC_ERR_UNSUPPORTED with a message naming the limit
 *   sharing   all calls on a device use the same array and it keeps the last call's contents; nothing in it has
 *             to start at zero (every slot read is written earlier in the same call), but calls that OVERLAP on
 *             the GPU -- different streams or unordered host threads -- race: keep them on one stream
 *   ops, ignore_nodata, ddof, outs, stream   as in eoc_reduce
 */
#define EOC_ROWS_STATIC_BYTES (512 * 1024)
int eoc_reduce_rows_static(const eoc_view* in, uint32_t ops, int ignore_nodata, int ddof, void* const* outs,
                           void* stream);

/*
 * eoc_selftest_nd_u16 -- exhaustive check of the unsigned fast path of eoc_nd_reduce: for all 2^32 pairs
 * (a, b) of uint16 samples the fused (a-b)/(a+b) must equal the IEEE-rounded float32 division bit for bit
 * (NaN for 0/0).  Adds the number of differing pairs to *mismatches (device pointer, caller zeroes it).
 */
int eoc_selftest_nd_u16(unsigned long long* mismatches, void* stream);

/* Number of kernel launches issued by this library on the calling thread since the last reset
 * (bench.py reports it as gpu_launches). */
int64_t eoc_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* EOC_H_ */
