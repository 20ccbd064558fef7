// launch.h -- host-side launcher declarations shared by the translation units of libde_grid4d.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace de {

struct EncodeParams;
struct SmemPlan;

cudaError_t launch_forward(const EncodeParams& P, int D, bool linear, bool half_tables, int F, cudaStream_t s);
cudaError_t launch_backward(const EncodeParams& P, int D, bool linear, int F, cudaStream_t s);
cudaError_t launch_indices(const EncodeParams& P, int D, bool linear, cudaStream_t s);

// lane-pair layout (grid4d_pair_kernels.cu): multilinear mode, F in {1,2,4,8}
bool pair_layout_supported(bool linear, int F);
cudaError_t launch_forward_pair(const EncodeParams& P, int D, bool half_tables, int F, bool pow2, cudaStream_t s);
cudaError_t launch_backward_pair(const EncodeParams& P, int D, int F, bool pow2, cudaStream_t s);

// TMA-staged backward (grid4d_staged_kernels.cu): group columns contiguous, rows 16 B aligned
size_t staged_smem_bytes(const EncodeParams& P, int F);
cudaError_t launch_backward_staged(const EncodeParams& P, int D, int F, bool pow2, cudaStream_t s);

// dense coarse levels served from shared memory (grid4d_smem_kernels.cu)
bool smem_layout_supported(bool half_tables, int F);
cudaError_t launch_forward_smem(const EncodeParams& P, const SmemPlan& S, int D, bool half_tables, int F, bool pow2,
                                cudaStream_t s);

// nearest-cell 1-D levels whose live rows fit shared memory (grid4d_small1d_kernels.cu)
constexpr int kSmall1dMaxLevels = 16;
struct Small1dPlan {
    uint32_t off[kSmall1dMaxLevels];      // byte offset of the level's rows (forward) / accumulators (backward), 16 B aligned
    uint32_t live[kSmall1dMaxLevels];     // rows kept in shared memory (0 = level served from global memory only)
    uint32_t copies[kSmall1dMaxLevels];   // backward: private accumulator copies (power of two)
    uint32_t total_bytes;
};
cudaError_t launch_small1d_forward(const EncodeParams& P, const Small1dPlan& S, bool half_tables, int F, int n_ctas, cudaStream_t s);
cudaError_t launch_small1d_backward(const EncodeParams& P, const Small1dPlan& S, int F, int n_ctas, cudaStream_t s);

// sort-by-cell + warp-aggregated backward (grid4d_sorted_kernels.cu; opt-in alternative to the gradient replicas)
struct SortParams {
    const float* coords;
    long long n;
    int coord_stride;
    int n_dims, bits;             // key = cells on a 2^bits-per-axis grid over columns dim[0..n_dims), x-major
    unsigned char dim[4];
    unsigned int* counts;         // 2^(bits * n_dims) counters (workspace)
    int* perm;                    // out: perm[i] = index of the i-th point in cell order
};
cudaError_t launch_sort_points(const SortParams& S, int n_sms, cudaStream_t s);
cudaError_t launch_backward_sorted(const EncodeParams& P, const int* perm, int F, bool pow2, int n_sms, cudaStream_t s);

struct CastParams {
    const float* src[64];
    void* dst[64];
    long long n_elems[64];
    int n_levels;
    float scale;
};
cudaError_t launch_cast_f16(const CastParams& P, cudaStream_t s);

// folds gradient replicas back into their tables: dst[i] += sum_r rep[r * stride + i]
struct ReplicaFoldParams {
    float* dst[64];
    const float* rep[64];
    unsigned n_floats[64];
    unsigned count[64];    // replicas in the workspace (rep_count - 1)
    unsigned stride[64];
    int n;
};
cudaError_t launch_replica_fold(const ReplicaFoldParams& P, cudaStream_t s);
// gradient rows -> group-major slices dst[g][p][0..W) = src[p][c0 + g*W ..) (misc.cu)
cudaError_t launch_regroup(const float* src, long long stride, int c0, int n_cols, int W, float* dst, long long n, int n_sms,
                           cudaStream_t s);

// tensor-core MLP head (mlp_head.cu): y = W2 * act(LN(W1 * x + b1)) + b2, hidden = out = 64
struct MlpHeadParams {
    const float* x;               // (n, in_dim) fp32 features, rows x_stride floats apart (16 B aligned)
    long long n;
    long long x_stride;
    const void* w1;               // (64, in_dim) bf16 row-major (nn.Linear.weight layout)
    const float* b1;              // (64,) or NULL
    const float* ln_gamma;        // (64,) or NULL
    const float* ln_beta;         // (64,) or NULL
    float ln_eps;
    int layernorm;                // 0 / 1
    int activation;               // 0 none, 1 ReLU, 2 GELU (erf)
    const void* w2;               // (64, 64) bf16 row-major
    const float* b2;              // (64,) or NULL
    float* y;                     // (n, 64) fp32, rows y_stride floats apart (16 B aligned)
    long long y_stride;
    int n_tiles;
};
struct MlpHeadBwdParams {
    const float* x;               // (n, in_dim) fp32 features (forward input)
    const float* gy;              // (n, 64) fp32 gradient of the head output
    long long n;
    long long x_stride, gy_stride;
    const void* w1;               // (64, in_dim) bf16
    const void* w1t;              // (in_dim, 64) bf16 = W1 transposed
    const void* w2t;              // (64 hidden, 64 out) bf16 = W2 transposed
    const float* b1;
    const float* ln_gamma;
    const float* ln_beta;
    float ln_eps;
    int layernorm;
    int activation;
    float* dx;                    // (n, in_dim) fp32 out
    long long dx_stride;
    int dx_group_w;               // 0: rows of dx_stride floats; W (power of two >= 4): group-major dx[c / W][p][c % W], the
                                  // layout de_grid4d_backward reads with DE_FLAG_GRAD_GROUPED (ping-pong kernel only)
    float* dw1;                   // (64, in_dim) fp32, accumulated (caller zeroes)
    float* db1;                   // (64,) accumulated, or NULL
    float* dgamma;                // (64,) accumulated, or NULL
    float* dbeta;                 // (64,) accumulated, or NULL
    float* dw2;                   // (64, 64) accumulated
    float* db2;                   // (64,) accumulated, or NULL
    int n_tiles;
};
// wide projection (proj_kernels.cu): hidden = GELU(LN(W1 x + b1)) as bf16, then a TMA / tcgen05 GEMM
struct ProjHiddenParams {
    const float* x;               // (n, in_dim) fp32 features, rows x_stride floats apart (16 B aligned)
    long long n;
    long long x_stride;
    const void* w1;               // (hidden, in_dim) bf16 row-major
    const float* b1;              // (hidden,)
    const float* ln_gamma;        // (hidden,)
    const float* ln_beta;         // (hidden,)
    float ln_eps;
    int hidden;                   // multiple of 128
    void* hidden_out;             // (n, hidden) bf16, rows hidden_stride elements apart
    long long hidden_stride;
    float* stats_out;             // (n, 2) {mean, rstd} of the pre-LayerNorm row, or NULL
    int n_tiles;
};
struct GemmParams {
    long long M;                  // rows of C
    int N;
    long long K;                  // contraction length
    int k_splits;                 // filled by the launcher
    int prefetch;                 // filled by the launcher: L2 prefetch cursor (DE_GEMM_PREFETCH=1; measured slower, off by default)
    void* c;                      // (M, N) fp32 or bf16, rows ldc elements apart
    long long ldc;
    const float* bias;            // (N,) or NULL
    int vec_ok;
    int bias_vec;                 // filled by the launcher: bias is 16-byte aligned
    int tma_store;                // filled by the launcher: fp32 C leaves through shared memory and TMA tensor stores
    int full_n;                   // filled by the launcher: DE_GEMM_FULL_N=1 keeps every MMA 256 columns wide (A/B switch)
    // optional: C = acc + bias + residual (fp32 store only; rows ldr floats apart)
    const float* residual;
    long long ldr;
    // optional (fp32 store only): C(r, c) *= GELU'(row_t[r] * col_w[c] + col_b[c]) -- the temporal MLP's backward (st_embed_tc.cu);
    // the erf evaluations run in the epilogue warps, under the next tile's mainloop
    const float* gg_row_t;
    const float* gg_col_w;
    const float* gg_col_b;
    // optional operand segments (split-precision products, st_embed_tc): the contraction runs over `n_seg` segments of
    // seg_blocks 64-element k-blocks each (K = n_seg * seg_blocks * 64); segment s multiplies operand part seg_a[s] of A
    // with part seg_b[s] of B.  Parts lie side by side in the operands' COLUMNS, seg_stride_a / _b elements apart: along K
    // for K-major operands (transposed = 0), along M / N for row-contracted ones (transposed = 1).  0 segments = off.
    int n_seg, seg_blocks;
    long long seg_stride_a, seg_stride_b;
    unsigned char seg_a[8], seg_b[8];
    long long a_rows, a_cols, b_rows, b_cols;     // true extents of the operand matrices when segments are used (tensor maps)
};
struct ProjDpreParams {
    const float* x;               // (n, in_dim) fp32 features (forward input)
    long long n;
    long long x_stride;
    const void* w1;               // (hidden, in_dim) bf16
    const float* b1;
    const float* ln_gamma;
    const float* ln_beta;
    const float* stats;           // (n, 2) {mean, rstd} saved by the forward
    const void* dact;             // (n, hidden) bf16: d(loss)/d(activation) = gy . W2, rows dact_stride elements apart
    long long dact_stride;
    void* dpre_out;               // (n, hidden) bf16 contiguous; may alias dact (each thread reads its piece before writing it)
    void* x_bf16_out;             // (n, in_dim) bf16 copy of the features, or NULL
    float* dgamma;                // (hidden,) accumulated; any of the three may be NULL
    float* dbeta;
    float* db1;
    int hidden;
    int n_tiles;
};
cudaError_t launch_proj_dpre(ProjDpreParams P, int in_dim, int n_sms, cudaStream_t s);
cudaError_t launch_transpose_bf16(const void* src, long long ld_src, void* dst, long long ld_dst, long long rows, int cols,
                                  cudaStream_t s);
cudaError_t launch_cast_colsum(const float* src, long long ld_src, void* dst_bf16, long long ld_dst, long long n_rows, int n_cols,
                               float* colsum, int n_sms, cudaStream_t s);
cudaError_t launch_colsum_bf16(const void* src_bf16, long long ld_src, long long n_rows, int n_cols, float* colsum, int n_sms,
                               cudaStream_t s);
bool proj_hidden_supported(int in_dim, int hidden);
cudaError_t launch_proj_hidden(ProjHiddenParams P, int in_dim, int n_sms, cudaStream_t s);
cudaError_t launch_gemm_bf16(const void* a, long long lda, const void* b, long long ldb, int transposed, GemmParams P, int epi,
                             int n_sms, cudaStream_t s);
bool mlp_head_supported(int in_dim, int hidden, int out_dim);
cudaError_t launch_mlp_head_bwd(MlpHeadBwdParams P, int in_dim, int n_sms, cudaStream_t s);        // mlp_head_bwd2.cu
cudaError_t launch_mlp_head(MlpHeadParams P, int in_dim, int n_sms, cudaStream_t s);

// dense-field evaluation over a regular lat/lon grid (grid4d_field_kernels.cu; SPECIFICATIONS.MD:1963-2088)
struct FieldParams {
    double lat_start, lat_stop, lat_step;     // np.linspace(lat_start, lat_stop, n_lat): step = (stop - start) / (n - 1)
    double lon_start, lon_stop, lon_step;
    double altitude, time;
    double bbox_min[3], bbox_span[3];
    double t_min, t_span;
    float* xyzt_out;                          // optional (n_lat * n_lon, 4): the generated normalised coordinates
    int n_lat, n_lon;
    int has_time;
};
int field_max_group(int F);
cudaError_t launch_field_forward(EncodeParams P, const FieldParams& Fp, bool half_tables, int F, bool pow2, cudaStream_t s);
cudaError_t launch_field_xyzt(const FieldParams& Fp, float* xyzt, cudaStream_t s);

// SpatialTemporalEmbedding (st_embed.cu; models/cross_modal_fusion.py:130-164)
struct StEmbedParams {
    const float* tokens;          // (n, D)
    const float* pos;             // (n, >= 2) spatial positions, rows pos_stride floats apart, or NULL
    const float* t;               // (n,) temporal positions, or NULL
    const float* ex;              // (R, D/4) spatial_embed_x
    const float* ey;              // (R, D/4) spatial_embed_y
    const float* w1;              // (D/2,)  temporal_embed.0.weight[:, 0]
    const float* b1;              // (D/2,)
    const float* w2;              // (D/2, D/2) temporal_embed.2.weight
    const float* b2;              // (D/2,)
    const float* mod;             // (D,) modality embedding, or NULL
    float* out;                   // (n, D)
    int* err;                     // device flag: bit 0 set when a spatial index falls outside [-R, R); may be NULL
    long long n;
    long long pos_stride;
    int D, R;
};
struct StEmbedBwdParams {
    const float* g;               // (n, D) gradient of the output (= gradient of the tokens)
    const float* pos;
    const float* t;
    const float* w1;
    const float* b1;
    const float* w2;
    float* dpre;                  // (n, D/2) scratch (temporal branch)
    float* d_ex;                  // accumulated (caller zeroes); any may be NULL
    float* d_ey;
    float* d_w1;
    float* d_b1;
    float* d_w2;
    float* d_b2;
    float* d_mod;
    long long n;
    long long pos_stride;
    int D, R;
};
cudaError_t launch_st_embed_forward(const StEmbedParams& P, cudaStream_t s);
cudaError_t launch_st_embed_backward(const StEmbedBwdParams& P, cudaStream_t s);
cudaError_t launch_st_embed_spatial(const StEmbedParams& P, int c_end, cudaStream_t s);
cudaError_t launch_st_embed_backward_common(const StEmbedBwdParams& P, cudaStream_t s);
// temporal MLP on the tensor cores (st_embed_tc.cu): dim / 2 a multiple of 64
bool st_embed_tc_supported(int D);
size_t st_embed_tc_workspace_bytes(long long n, int D, bool backward);
cudaError_t launch_st_embed_forward_tc(const StEmbedParams& P, void* ws, int n_sms, cudaStream_t s);
cudaError_t launch_st_embed_backward_tc(const StEmbedBwdParams& P, void* ws, int n_sms, cudaStream_t s);

cudaError_t launch_geo2xyzt(const void* geo, bool geo_f64, const void* t, bool t_f64, const double* bbox_min,
                            const double* bbox_span, const double* t_range, float* xyzt, double* ecef,
                            long long n, cudaStream_t s);
cudaError_t launch_time_components(const void* comp, bool is_f64, long long n, long long stride, int n_cols, int def_month,
                                   int def_day, int def_hour, double t_min, double t_span, double* ts_out, float* tn_out,
                                   cudaStream_t s);
cudaError_t launch_geo_bbox(const void* geo, bool geo_f64, double* minmax, long long n, cudaStream_t s);
cudaError_t launch_sum_f32(const float* x, long long n, double* result, cudaStream_t s);
cudaError_t launch_probe(void* table, long long n_entries, int entry_bytes, int mode, long long n_threads,
                         int per_thread, int pair_pattern, void* sink, cudaStream_t s);

}  // namespace de
