/*
 * de_grid4d.h -- C ABI of libde_grid4d.so, the B200 (sm_100a) implementation of
 * DeepEarth's Grid4D multi-resolution hash-grid spacetime encoder.
 *
 * The reference has no FFI for this path: its boundary is two Python nn.Modules
 * (SURVEY.md section 8b).  Every entry point below therefore names the reference
 * Python symbol whose work it replaces; deepearth_b200/{hash_encoding,encoders,
 * geo2xyz}.py bind them through ctypes (see INTEGRATION.md for the stub a
 * reference maintainer would add).
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the
 *     argument name ends in _host (level descriptors are read on the host and
 *     travel to the kernels as launch parameters -- no device allocation, no
 *     global state, re-entrant);
 *   - `stream` is a cudaStream_t / CUstream passed as void*; nothing synchronises;
 *   - return value: 0 on success, a negative DE_E* code for argument errors,
 *     a positive cudaError_t for launch errors; de_last_error() gives the text
 *     for the calling thread;
 *   - there is no CPU implementation behind any of these.
 */
#ifndef DE_GRID4D_H
#define DE_GRID4D_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DE_API __attribute__((visibility("default")))
#else
#define DE_API
#endif

#define DE_ABI_VERSION 5

/* status codes (negative = caller error) */
#define DE_OK 0
#define DE_EINVAL (-1)      /* bad argument value                                  */
#define DE_EUNSUPPORTED (-2) /* combination the kernels are not instantiated for   */
#define DE_EALIGN (-3)      /* pointer / stride alignment requirement violated     */

/* element types */
#define DE_F32 0
#define DE_F16 1
#define DE_F64 2
#define DE_I32 3
#define DE_BF16 4

/* flags */
#define DE_INTERP_NEAREST 0u /* floor cell only: models/hash_encoding.py:57-72     */
#define DE_INTERP_LINEAR 1u  /* 2^D-corner multilinear (decomposed Grid4D mode)    */
#define DE_INTERP_MASK 1u
/* opt-in code paths that measured SLOWER than the default on B200 (profiles/r01c_ab_matrix.txt); kept for A/B runs */
#define DE_FLAG_STAGED_BWD 0x100u  /* backward: stage coordinates + gradient rows through TMA bulk copies + shared memory */
#define DE_FLAG_SMEM_COARSE 0x200u /* forward: serve dense coarse levels from shared-memory copies of their tables      */
#define DE_FLAG_NO_L2_HINTS 0x400u /* A/B: plain loads / reductions instead of evict_last tables + evict_first streams */
#define DE_FLAG_NO_REGROUP 0x800u  /* A/B: backward reads gradient-row slices in place instead of regrouping them first  */
#define DE_FLAG_GRAD_GROUPED 0x1000u /* backward: grad_out is already group-major (see de_grid4d_backward_group_width)            */
#define DE_FLAG_COHERENT_GATHER 0x2000u /* A/B: forward table gathers as GPU-scope relaxed loads (home L2 partition only)         */
#define DE_FLAG_NO_SMALL1D 0x4000u /* A/B: nearest-cell 1-D levels through the general kernels instead of the shared-memory path */
#define DE_FLAG_MASK (DE_FLAG_NO_SMALL1D | DE_INTERP_MASK | DE_FLAG_STAGED_BWD | DE_FLAG_SMEM_COARSE | DE_FLAG_NO_L2_HINTS | DE_FLAG_NO_REGROUP | DE_FLAG_GRAD_GROUPED | DE_FLAG_COHERENT_GATHER)

#define DE_MAX_DIMS 4
#define DE_MAX_LEVELS_PER_CALL 256

/*
 * One resolution level of one grid.  A "grid" is a D-dimensional hash grid over a
 * subset of the coordinate columns (xyz -> dim={0,1,2}; xyt -> {0,1,3}; yzt ->
 * {1,2,3}; xzt -> {0,2,3}; the reference's temporal encoder -> D=1, dim={3}).
 * Replaces one iteration of the level loop at models/hash_encoding.py:87-93 and
 * the nn.Embedding it indexes (:38-41).
 */
typedef struct de_level_desc {
    const void* table;   /* (table_size, F) row-major entries, dtype = table_dtype  */
    float* grad_table;   /* (table_size, F) fp32 accumulators; backward only        */
    uint32_t table_size; /* T                                                       */
    uint32_t resolution; /* integer R; coords are scaled by (float)R in fp32        */
    uint32_t out_col;    /* first column of this level in the feature row           */
    uint8_t n_dims;      /* D in 1..4                                               */
    uint8_t dim[DE_MAX_DIMS]; /* coordinate column feeding each grid dimension      */
    uint8_t dense;       /* 1: index = sum cell_d*(R+1)^d (requires (R+1)^D <= T)   */
    uint8_t mask_id;     /* 0 none, 1 mask_a (spatial), 2 mask_b (temporal)         */
    uint8_t reserved;
} de_level_desc;

DE_API int de_abi_version(void);
DE_API const char* de_last_error(void);

/* Number of interpolation corners per level for the given flags and D. */
DE_API int de_corners(uint32_t flags, int n_dims);

/*
 * Hash / dense cell indices -- replaces MultiResolutionHashEncoding.hash_coords
 * (models/hash_encoding.py:47-74) for every level in one launch.
 * indices layout: [level][corner][point] int64, corner count = de_corners().
 * All levels must share n_dims.
 */
DE_API int de_grid4d_hash_indices(const de_level_desc* levels_host, int n_levels,
                           const float* coords, int64_t n_points, int coord_stride,
                           uint32_t flags, int64_t* indices, void* stream);

/*
 * Forward encode -- replaces MultiResolutionHashEncoding.forward
 * (models/hash_encoding.py:76-96) and, with levels from several grids plus the two
 * row masks, the feature half of Grid4DEncoder.forward (models/encoders.py:68-82).
 *   out[p, out_col + f] = mask(p) * out_scale * sum_k w_k(p) * table[idx_k(p)][f]
 * coords: (n_points, coord_stride) fp32; out: rows of out_stride floats.
 * mask_a / mask_b: optional (n_points,) bytes, 0 zeroes the level's features
 * (models/encoders.py:73-79); pass NULL for none.
 * levels_per_group: how many consecutive levels one CTA handles (0 = auto); the
 * launch walks groups in order so a group's tables stay L2 resident.
 */
DE_API int de_grid4d_forward(const de_level_desc* levels_host, int n_levels,
                      const float* coords, int64_t n_points, int coord_stride,
                      int table_dtype, int n_features, float out_scale,
                      const uint8_t* mask_a, const uint8_t* mask_b,
                      float* out, int64_t out_stride,
                      uint32_t flags, int levels_per_group, void* stream);

/*
 * Backward -- replaces autograd's aten::embedding_dense_backward for the tables
 * (SURVEY.md section 3.2): grad_table[idx_k(p)][f] += mask(p) * w_k(p) *
 * grad_out[p, out_col + f], accumulated with vector fp32 reductions.  grad_table
 * buffers must be zeroed (or hold the running sum) by the caller.  Coordinates
 * receive no gradient, as in the reference (graph cut at hash_encoding.py:61).
 * workspace: device scratch of de_grid4d_backward_workspace_bytes() bytes, 16 B aligned, or NULL.
 */
DE_API int de_grid4d_backward(const de_level_desc* levels_host, int n_levels,
                       const float* coords, int64_t n_points, int coord_stride,
                       int n_features,
                       const uint8_t* mask_a, const uint8_t* mask_b,
                       const float* grad_out, int64_t grad_stride,
                       uint32_t flags, int levels_per_group,
                       void* workspace, size_t workspace_bytes, void* stream);

/*
 * Scratch the backward wants for this call (0 = none).  Coarse levels with few live
 * entries take thousands of reductions per entry; with a workspace the kernel spreads
 * them over private copies ("replicas") that a small kernel folds into grad_table
 * afterwards.  The call zeroes what it uses; pass NULL / 0 to run without replicas.
 */
DE_API size_t de_grid4d_backward_workspace_bytes(const de_level_desc* levels_host, int n_levels,
                                                 int n_features, int64_t n_points, uint32_t flags);
/*
 * Same, for an explicit levels_per_group.  Besides the replicas the scratch holds the gradient rows re-laid
 * group-major when a level group's slice of a row is <= 64 B (large tables => few levels per group): every
 * walk over the points then streams contiguous bytes instead of one 128 B HBM line per slice.  A smaller
 * workspace is accepted; the call then skips what does not fit.
 */
DE_API size_t de_grid4d_backward_workspace_bytes_ex(const de_level_desc* levels_host, int n_levels,
                                                    int n_features, int64_t n_points, uint32_t flags,
                                                    int levels_per_group);

/*
 * Producer-side regrouping.  When this returns W > 0, de_grid4d_backward for the same arguments would re-lay the
 * gradient rows as  g[c / W][p][c % W]  (c = feature column, W = levels-per-group * n_features floats) before
 * walking them.  A producer that can write that layout itself -- de_mlp_head_backward with gx_stride = -W --
 * passes the buffer with DE_FLAG_GRAD_GROUPED and grad_stride = W, and the pass is skipped.  0 = not applicable
 * (levels of several dimensionalities, first column not 0, small inputs, wide slices): use the row-major layout.
 */
DE_API int de_grid4d_backward_group_width(const de_level_desc* levels_host, int n_levels, int n_features,
                                          int64_t n_points, uint32_t flags, int levels_per_group);

/*
 * Sort-by-cell + warp-aggregated scatter: the alternative backward for densely colliding (coarse) levels.  Opt-in; the
 * default backward spreads such levels over gradient replicas instead (A/B in profiles/).
 *   de_grid4d_sort_points: perm[i] = index of the i-th point when the points are ordered by the cell they occupy on a
 *     2^bits-per-axis grid over coordinate columns dims_host[0 .. n_dims) (x-major; coordinates outside [0, 1) clamp to the
 *     border cells).  Counting sort: bits * n_dims <= 16; workspace = de_grid4d_sort_workspace_bytes() bytes.
 *   de_grid4d_backward_sorted: same contract as de_grid4d_backward for 3-D DE_INTERP_LINEAR levels, n_features in
 *     {1, 2, 4}; a warp walks 32 consecutive points of `perm` for one level, the lanes of a run that shares a cell add their
 *     contributions with a segmented shuffle scan and one lane issues the reduction.  Any permutation gives the same
 *     gradients (up to fp32 summation order); only a cell-ordered one makes it fast.
 */
DE_API size_t de_grid4d_sort_workspace_bytes(int n_dims, int bits);
DE_API int de_grid4d_sort_points(const float* coords, int64_t n_points, int coord_stride, const uint8_t* dims_host, int n_dims,
                          int bits, int32_t* perm, void* workspace, size_t workspace_bytes, void* stream);
DE_API int de_grid4d_backward_sorted(const de_level_desc* levels_host, int n_levels, const float* coords, int64_t n_points,
                              int coord_stride, int n_features, const uint8_t* mask_a, const uint8_t* mask_b,
                              const float* grad_out, int64_t grad_stride, uint32_t flags, const int32_t* perm, void* stream);

/*
 * de_grid4d_sort_points for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and result as de_grid4d_sort_points minus the workspace: the bucket counters live in a 256 KiB static
 * device array inside the library (2^16 buckets x 4 B, i.e. the full bits * n_dims <= 16 range of the sort; held by every GPU
 * that loads the library until the process exits).  Sizes beyond DE_SORT_NOALLOC_MAX_BUCKETS buckets return
 * DE_EUNSUPPORTED.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the first call per
 * device looks the array's address up (cudaGetSymbolAddress with that device current) -- a call with n_points = 0 does just
 * that and can serve as warm-up.  ALL calls on a device share the array (it is zeroed at the start of each call, on `stream`,
 * and keeps the last call's counters afterwards): two calls whose kernels overlap on the GPU -- different streams, or
 * several host threads -- race; order them on one stream or with events.  perm: as for de_grid4d_sort_points (the order of
 * points inside one cell is unspecified in both).
 */
#define DE_SORT_NOALLOC_MAX_BUCKETS 65536
DE_API int de_grid4d_sort_points_noalloc(const float* coords, int64_t n_points, int coord_stride, const uint8_t* dims_host, int n_dims,
                                  int bits, int32_t* perm, void* stream);

/*
 * fp32 master tables -> fp16 working tables, dst = half(src * scale), one launch
 * for all levels (level i: src = levels_host[i].table (fp32), dst =
 * dst_tables[i]).  dst_tables_host: host array of n_levels device pointers.
 */
DE_API int de_tables_to_f16(const de_level_desc* levels_host, int n_levels, int n_features,
                     void* const* dst_tables_host, float scale, void* stream);

/*
 * Geodetic prologue -- replaces GeospatialConverter.geodetic_to_xyz position
 * branch (geospatial/geo2xyz.py:146-160) + xyz_to_norm with a pre-fitted bounding
 * box (:327-332, utils.py:54-77) + the spec's time normalisation
 * (SPECIFICATIONS.MD:765), fused into one pass.
 *   geo: (n,3) [lat deg, lon deg, alt m] fp64 or fp32 (geo_dtype)
 *   t:   (n,) fp64 or fp32 (t_dtype) or NULL (t column written as 0)
 *   bbox_min_host/bbox_span_host: 3 doubles each; t_range_host: {t_min, t_max}
 *   xyzt: (n,4) fp32 out; ecef: optional (n,3) fp64 out (NULL to skip)
 */
DE_API int de_geo2xyzt_norm(const void* geo, int geo_dtype, const void* t, int t_dtype,
                     const double* bbox_min_host, const double* bbox_span_host,
                     const double* t_range_host,
                     float* xyzt, double* ecef, int64_t n_points, void* stream);

/*
 * Timestamps from date columns -- replaces the per-row `datetime(...).timestamp()` loop of the spec's preprocessor
 * (SPECIFICATIONS.MD:740-768, `_process_temporal`).  components: (n, n_cols) rows, `stride` elements apart, of
 * [year, month, day, hour, minute, second] (n_cols in 1..6; missing month / day / hour take the defaults, minute and
 * second 0), dtype DE_I32 or DE_F64 (fractional seconds give microseconds, int((second % 1) * 1e6)).  Times are UTC.
 * timestamps: optional (n,) fp64 epoch seconds; t_norm: optional (n,) fp32 (ts - t_min) / (t_max - t_min) with
 * t_range_host = {t_min, t_max} (required when t_norm is given).
 */
DE_API int de_time_from_components(const void* components, int dtype, int64_t n, int64_t stride, int n_cols,
                            int default_month, int default_day, int default_hour, const double* t_range_host,
                            double* timestamps, float* t_norm, void* stream);

/*
 * Bounding box of ECEF positions of a geodetic batch -- replaces
 * GeospatialConverter.update_bbox / BoundingBox.from_points
 * (geo2xyz.py:305-320, data_structures.py:95-100).  minmax: 6 doubles on the
 * device {min_x,min_y,min_z,max_x,max_y,max_z}; the caller initialises them to
 * +inf/-inf (or a running box) -- the kernel folds the batch in with atomics.
 */
DE_API int de_geo_bbox(const void* geo, int geo_dtype, double* minmax, int64_t n_points,
                void* stream);

/*
 * Tensor-core MLP head -- replaces Grid4DEncoder.projection (models/encoders.py:47-52:
 * Linear -> LayerNorm -> GELU -> Linear) at the 64-wide size of BASELINE.json configs[4]:
 *   y = W2 * act( LN( W1 * x + b1 ) ) + b2        hidden = out_dim = 64
 * One kernel: bf16 tcgen05.mma with fp32 accumulators in tensor memory; the hidden row never
 * leaves the SM.  x: (n, in_dim) fp32 features (the encoder's output), in_dim in {32,48,64,96,128};
 * w1_bf16: (64, in_dim), w2_bf16: (64, 64) bf16 in nn.Linear.weight layout; biases / LayerNorm
 * vectors fp32 or NULL; layernorm 0/1; activation 0 none, 1 ReLU, 2 GELU (erf form).
 */
DE_API int de_mlp_head_supported(int in_dim, int hidden, int out_dim);
DE_API int de_mlp_head_forward(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                        const void* w1_bf16, const float* b1,
                        const float* ln_gamma, const float* ln_beta, float ln_eps, int layernorm,
                        int activation, const void* w2_bf16, const float* b2,
                        int hidden, int out_dim, float* y, int64_t y_stride, void* stream);

/*
 * Backward of the same head, one tcgen05 kernel: recomputes the hidden row, returns grad_x (n, in_dim) -- the
 * gradient the hash-grid backward consumes -- and ACCUMULATES the parameter gradients (caller zeroes them):
 * grad_w1 (64, in_dim), grad_b1, grad_gamma, grad_beta, grad_w2 (64, 64), grad_b2 (each (64,), any may be NULL
 * except the two weight gradients).  w1t_bf16 = W1 transposed (in_dim, 64), w2t_bf16 = W2 transposed (hidden, out).
 * Replaces autograd's LinearBackward / GeluBackward / NativeLayerNormBackward chain for models/encoders.py:47-52.
 * gx_stride > 0: grad_x rows are gx_stride floats apart.  gx_stride = -W (W a power of two >= 4 dividing in_dim):
 * grad_x is written group-major, grad_x[c / W][p][c % W] -- the layout de_grid4d_backward_group_width() describes.
 */
DE_API int de_mlp_head_backward(const float* x, const float* grad_y, int64_t n_points, int in_dim,
                         int64_t x_stride, int64_t gy_stride,
                         const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                         const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps,
                         int layernorm, int activation, int hidden, int out_dim,
                         float* grad_x, int64_t gx_stride, float* grad_w1, float* grad_b1,
                         float* grad_gamma, float* grad_beta, float* grad_w2, float* grad_b2, void* stream);

/*
 * The projection at its real width -- replaces Grid4DEncoder.projection (models/encoders.py:47-52) for hidden_dim = 768
 * (models/configs.py:44), 2048 (models/deepearth_integrated.py:69-74) or any multiple of 128:
 *   a = GELU( LN( W1 x + b1 ) )   one tcgen05 kernel: x (n, in_dim) fp32 read once, a (n, hidden) bf16 written once; the
 *                                 first Linear runs twice per 128-row tile in 128-column TMEM chunks (pass 1 = LayerNorm
 *                                 statistics, pass 2 = normalise + GELU), so the pre-activation never exists in memory;
 *   y = W2 a + b2                 persistent TMA (cp.async.bulk.tensor, 128 B swizzle) + tcgen05.mma GEMM: CTA pairs
 *                                 (cta_group::2, 256 x 256 tiles), double-buffered TMEM accumulators, fp32 output through
 *                                 TMA tensor stores.
 * in_dim in {32, 48, 64, 96, 128}; out_dim a multiple of 8.  w1_bf16 (hidden, in_dim), w2_bf16 (out_dim, hidden): bf16 in
 * nn.Linear.weight layout.  hidden_bf16: (n, hidden) bf16 device buffer (the activations the backward needs);
 * row_stats: optional (n, 2) fp32 {mean, rstd} of the pre-LayerNorm rows.
 */
DE_API int de_projection_supported(int in_dim, int hidden, int out_dim);
DE_API int de_projection_forward(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                          const void* w1_bf16, const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps,
                          const void* w2_bf16, const float* b2, int hidden, int out_dim,
                          void* hidden_bf16, float* row_stats, float* y, int64_t y_stride, void* stream);
/* Same, with the output dtype stated: y_dtype DE_F32 or DE_BF16 (y_stride in elements).  DE_BF16 is what nn.Linear returns
 * under torch.autocast(bfloat16) -- the mode the reference trains in (hpc/train_distrbuted.py:284): the GEMM's epilogue rounds
 * acc + b2 once, so the result equals the fp32 output rounded to bf16, without the 6 bytes per element of a separate cast. */
DE_API int de_projection_forward_ex(const float* x, int64_t n_points, int in_dim, int64_t x_stride,
                             const void* w1_bf16, const float* b1, const float* ln_gamma, const float* ln_beta, float ln_eps,
                             const void* w2_bf16, const float* b2, int hidden, int out_dim,
                             void* hidden_bf16, float* row_stats, void* y, int y_dtype, int64_t y_stride, void* stream);
/*
 * Backward of de_projection_forward -- replaces autograd's AddmmBackward / GeluBackward / NativeLayerNormBackward chain for
 * models/encoders.py:47-52 at hidden_dim 768 / 2048.  grad_x (n, in_dim) fp32 is the gradient the hash grids consume; the
 * parameter gradients are ACCUMULATED (caller zeroes them; any but the two weight gradients may be NULL).
 * w1t_bf16 = W1 transposed (in_dim, hidden), w2t_bf16 = W2 transposed (hidden, out_dim); hidden_bf16 / row_stats are what the
 * forward wrote.  workspace: de_projection_backward_workspace_bytes() bytes, 16-byte aligned (bf16 copies of grad_y and x, and
 * d(loss)/d(activation), overwritten in place by d(loss)/d(pre-activation)).
 * Kernels: cast + column sums (db2), GEMM da = gy W2, proj_dpre_kernel (GELU' and LayerNorm backward with the first Linear
 * recomputed in tensor memory; dgamma, dbeta, db1), two split-K GEMMs over the points (dW2, dW1), GEMM dx = dpre W1.
 */
DE_API size_t de_projection_backward_workspace_bytes(int64_t n_points, int in_dim, int hidden, int out_dim);
DE_API int de_projection_backward(const float* x, const float* grad_y, int64_t n_points, int in_dim, int64_t x_stride,
                           int64_t gy_stride, const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                           const float* b1, const float* ln_gamma, const float* ln_beta, int hidden, int out_dim,
                           const void* hidden_bf16, const float* row_stats,
                           float* grad_x, int64_t gx_stride, float* grad_w1, float* grad_b1, float* grad_gamma,
                           float* grad_beta, float* grad_w2, float* grad_b2,
                           void* workspace, size_t workspace_bytes, void* stream);
/* Same, with the dtype of grad_y stated: DE_F32, or DE_BF16 (autocast: the gradient of a bf16 y) -- then grad_y (rows
 * gy_stride elements apart, a multiple of 8) is consumed in place by the GEMMs: no fp32 -> bf16 copy pass, db2 by a column-sum
 * kernel over the bf16 rows.  Same workspace size. */
DE_API int de_projection_backward_ex(const float* x, const void* grad_y, int gy_dtype, int64_t n_points, int in_dim, int64_t x_stride,
                              int64_t gy_stride, const void* w1_bf16, const void* w1t_bf16, const void* w2t_bf16,
                              const float* b1, const float* ln_gamma, const float* ln_beta, int hidden, int out_dim,
                              const void* hidden_bf16, const float* row_stats,
                              float* grad_x, int64_t gx_stride, float* grad_w1, float* grad_b1, float* grad_gamma,
                              float* grad_beta, float* grad_w2, float* grad_b2,
                              void* workspace, size_t workspace_bytes, void* stream);

/*
 * The GEMM kernel on its own: C (M, N) = A . B^T + bias, bf16 operands, fp32 accumulation in tensor memory.
 *   transposed = 0: A (M, K), B (N, K) row-major (K contiguous: nn.Linear layout);
 *   transposed = 1: A (K, M), B (K, N) row-major -- the contraction runs over rows (weight gradients, K = points);
 *   out_dtype DE_F32 / DE_BF16; accumulate = 1: C (fp32) += product, the K range split over CTAs (red.global.add).
 */
DE_API int de_gemm_bf16(const void* a_bf16, int64_t lda, const void* b_bf16, int64_t ldb, int64_t m, int n, int64_t k,
                 const float* bias, void* c, int64_t ldc, int out_dtype, int transposed, int accumulate, void* stream);

/*
 * Device-wide L2 fetch granularity hint (cudaLimitMaxL2FetchGranularity) of the current device: 32, 64 or
 * 128 bytes.  The encoder reads gradient rows in pieces of G*F*4 bytes per level group; with the
 * default 128 B granularity the L2 over-fetches the rest of each line from HBM.  Application-level
 * knob (it affects every kernel of the process on that device); get returns the value in *bytes.
 */
DE_API int de_set_l2_fetch_granularity(int bytes);
DE_API int de_get_l2_fetch_granularity(int* bytes);

/*
 * L2 set-aside for persisting accesses (cudaLimitPersistingL2CacheSize) of the current device.  The table
 * gathers and gradient reductions carry an L2::evict_last policy; the set-aside is the share of L2 such lines
 * may hold against the streamed coordinates / feature rows.  bytes < 0 asks for the device maximum
 * (cudaDeviceProp::persistingL2CacheMaxSize); get returns the current value and the maximum.
 */
DE_API int de_set_persisting_l2(long long bytes);
DE_API int de_get_persisting_l2(long long* bytes, long long* max_bytes);

/*
 * Dense-field evaluation -- the encoder over a regular lat/lon grid at fixed altitude and time, inference only.
 * Replaces the query construction + preprocessing + encoder forward of `visualize_embedding_field`
 * (SPECIFICATIONS.MD:1963-2088: np.linspace latitudes x np.linspace longitudes, latitude-major point order) and of the
 * species distribution maps (core/plant_data_loader.py:543-549).  No coordinate tensor exists: each CTA generates the
 * geodetic points of its 16 x 16 patch, runs the de_geo2xyzt_norm arithmetic on them and encodes them; neighbouring points
 * share a warp so coarse-level gathers coalesce.  Point p = i_lat * n_lon + i_lon.
 */
typedef struct de_field_desc {
    double lat_start, lat_stop;   /* np.linspace(lat_start, lat_stop, n_lat), degrees */
    double lon_start, lon_stop;
    double altitude;              /* metres, one value for the whole grid             */
    double time;                  /* raw time value, normalised with t_range_host     */
    int32_t n_lat, n_lon;
    int32_t has_time;             /* 0: t column = 0                                  */
    int32_t reserved;
} de_field_desc;
/* Normalised coordinates of the grid, (n_lat * n_lon, 4) fp32 -- bit-identical to de_geo2xyzt_norm on the same points. */
DE_API int de_field_xyzt(const de_field_desc* field_host, const double* bbox_min_host, const double* bbox_span_host,
                  const double* t_range_host, float* xyzt, void* stream);
/*
 * Features of the grid: out[p, out_col + f] as de_grid4d_forward would produce for the same points (bit-identical).
 * Levels must be 3-D with DE_INTERP_LINEAR and n_features in {1, 2, 4}; other level lists return DE_EUNSUPPORTED
 * (use de_field_xyzt + de_grid4d_forward).  xyzt_out: optional (n_lat * n_lon, 4) copy of the generated coordinates.
 */
DE_API int de_grid4d_field_forward(const de_level_desc* levels_host, int n_levels, const de_field_desc* field_host,
                            const double* bbox_min_host, const double* bbox_span_host, const double* t_range_host,
                            int table_dtype, int n_features, float out_scale, float* out, int64_t out_stride,
                            float* xyzt_out, uint32_t flags, int levels_per_group, void* stream);

/*
 * SpatialTemporalEmbedding -- replaces SpatialTemporalEmbedding.forward (models/cross_modal_fusion.py:130-164), the other
 * consumer of xyzt in the integrated model:
 *   out[m, :D/4]      = tokens + embed_x[(pos[m,0] * R).long()]   (:145-152; negative indices wrap, as PyTorch indexing)
 *   out[m, D/4:D/2]   = tokens + embed_y[(pos[m,1] * R).long()]
 *   out[m, D/2:]      = tokens + W2 . GELU(t[m] * w1 + b1) + b2    (:155-158, fp32)
 *   (+ modality[c] inside the bracket when given, :161-162; summation order of the reference)
 * tokens/out: (n, dim) fp32 contiguous; pos: (n, >=2) rows pos_stride floats apart or NULL; t: (n,) or NULL;
 * embed_x/embed_y: (resolution, dim/4); t_w1, t_b1, t_b2: (dim/2,); t_w2: (dim/2, dim/2) row-major (nn.Linear.weight);
 * modality: (dim,) or NULL.  dim must be a multiple of 4.  error_flag: optional device int, bit 0 is set when an index
 * falls outside [-resolution, resolution) (the reference raises IndexError there); such rows are written as NaN.
 */
DE_API int de_st_embedding_forward(const float* tokens, int64_t n_tokens, int dim,
                            const float* pos, int64_t pos_stride, const float* t,
                            const float* embed_x, const float* embed_y, int resolution,
                            const float* t_w1, const float* t_b1, const float* t_w2, const float* t_b2,
                            const float* modality, float* out, int* error_flag, void* stream);
/*
 * Backward of the same module -- replaces autograd's IndexBackward / AddmmBackward / GeluBackward chain for
 * models/cross_modal_fusion.py:143-164.  grad_out (n, dim) is also the gradient of `tokens` (identity).  Parameter
 * gradients are ACCUMULATED (caller zeroes them); any may be NULL.  scratch: (n, dim/2) floats, required when t is given.
 * Positions receive no gradient (the spatial ones are cut by .long(); temporal positions are data).
 */
DE_API int de_st_embedding_backward(const float* grad_out, int64_t n_tokens, int dim,
                             const float* pos, int64_t pos_stride, const float* t, int resolution,
                             const float* t_w1, const float* t_b1, const float* t_w2,
                             float* grad_embed_x, float* grad_embed_y, float* grad_t_w1, float* grad_t_b1,
                             float* grad_t_w2, float* grad_t_b2, float* grad_modality,
                             float* scratch, void* stream);

/*
 * The same two calls with the temporal MLP's second Linear (2 (dim/2)^2 flop per token) on the tensor cores at fp32 accuracy:
 * every fp32 operand is split into three bf16 parts (8 + 8 + 8 mantissa bits) and the six significant part products are
 * accumulated in fp32 in tensor memory by ONE TMA + tcgen05 GEMM over six operand segments (csrc/st_embed_tc.cu) -- the
 * reference runs this layer as an fp32 cuBLAS SGEMM.  dim must be a multiple of 128 (de_st_embedding_tc_supported); results
 * agree with the fp32 entry points to fp32 rounding (same tolerances in tests/test_st_embed.py).  workspace:
 * de_st_embedding_tc_workspace_bytes(n_tokens, dim, backward) bytes, 256-byte aligned (split operands of one slab of tokens;
 * bounded: <= ~1.1 GiB at dim 2048); the backward needs no (n, dim/2) scratch.
 */
DE_API int de_st_embedding_tc_supported(int dim);
DE_API size_t de_st_embedding_tc_workspace_bytes(int64_t n_tokens, int dim, int backward);
DE_API int de_st_embedding_forward_tc(const float* tokens, int64_t n_tokens, int dim,
                               const float* pos, int64_t pos_stride, const float* t,
                               const float* embed_x, const float* embed_y, int resolution,
                               const float* t_w1, const float* t_b1, const float* t_w2, const float* t_b2,
                               const float* modality, float* out, int* error_flag,
                               void* workspace, size_t workspace_bytes, void* stream);
DE_API int de_st_embedding_backward_tc(const float* grad_out, int64_t n_tokens, int dim,
                                const float* pos, int64_t pos_stride, const float* t, int resolution,
                                const float* t_w1, const float* t_b1, const float* t_w2,
                                float* grad_embed_x, float* grad_embed_y, float* grad_t_w1, float* grad_t_b1,
                                float* grad_t_w2, float* grad_t_b2, float* grad_modality,
                                void* workspace, size_t workspace_bytes, void* stream);

/* fp64 sum of an fp32 buffer into *result (device, caller-zeroed): bench/e2e checksum. */
DE_API int de_sum_f32(const float* x, int64_t n, double* result, void* stream);

/*
 * Gather / reduction peak probes used to define the "gather roofline"
 * (SURVEY.md section 8d).  table: working set of n_entries entries of entry_bytes
 * (4, 8 or 16); each of n_threads threads issues `per_thread` independent random
 * accesses (xorshift addresses, optional x-pair pattern).  mode 0 = load,
 * 1 = red.add.f32 vector.  sink: >= 4 bytes.
 */
DE_API int de_probe_gather(void* table, int64_t n_entries, int entry_bytes, int mode,
                    int64_t n_threads, int per_thread, int pair_pattern,
                    void* sink, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DE_GRID4D_H */
