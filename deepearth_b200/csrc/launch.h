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

struct CastParams {
    const float* src[64];
    void* dst[64];
    long long n_elems[64];
    int n_levels;
    float scale;
};
cudaError_t launch_cast_f16(const CastParams& P, cudaStream_t s);

cudaError_t launch_geo2xyzt(const void* geo, bool geo_f64, const void* t, bool t_f64, const double* bbox_min,
                            const double* bbox_span, const double* t_range, float* xyzt, double* ecef,
                            long long n, cudaStream_t s);
cudaError_t launch_geo_bbox(const void* geo, bool geo_f64, double* minmax, long long n, cudaStream_t s);
cudaError_t launch_sum_f32(const float* x, long long n, double* result, cudaStream_t s);
cudaError_t launch_probe(void* table, long long n_entries, int entry_bytes, int mode, long long n_threads,
                         int per_thread, int pair_pattern, void* sink, cudaStream_t s);

}  // namespace de
