// grid4d_small1d_kernels.cu -- nearest-cell 1-D levels with a small live set, served from shared memory (sm_100a).
//
// The reference's temporal encoder (/root/reference/models/encoders.py:33-39: 8 levels, R = 4 .. 512, T = 2^17) hashes a
// single coordinate with prime[0] = 1 (models/hash_encoding.py:64), so the index IS the cell number: for t in [0, 1]
// a level touches entries 0 .. R only -- at most 513 of its 131072 rows.  Through the general kernels those levels were
// the least efficient part of the reference geometry (ncu r01i: backward 0.27 ms for a third of the reductions, all of
// them landing on <= 513 addresses per level in the L2 atomic units, with gradient replicas to soften it).
// Here a CTA keeps the live rows in shared memory:
//   forward:  the live rows of every level are copied in once per CTA; a gather is a shared-memory load;
//   backward: the CTA accumulates its whole tile of points with shared-memory atomics (a few private copies per level,
//             picked by point slot, so the lanes of a warp never collide) and flushes each level once.
// Indices outside the live range (coordinates outside [0, 1]) take the ordinary global path, so results are exactly the
// general kernels': features bit-identical (pure copy), gradients equal up to fp32 summation order.
#include "grid4d_common.cuh"
#include "launch.h"

namespace de {

namespace {

__device__ __forceinline__ uint32_t index_1d(const LevelParams& lv, float t) {
    const float s = __fmul_rn(t, lv.res_f);            // hash_encoding.py:58
    const long long g = __float2ll_rd(s);              // :61
    if (lv.pow2) return static_cast<uint32_t>(g) & (lv.T - 1u);
    long long r = g % static_cast<long long>(lv.T);    // :72, Python floor-mod (prime[0] = 1)
    if (r < 0) r += lv.T;
    return static_cast<uint32_t>(r);
}

__device__ __forceinline__ float coord_1d(const EncodeParams& P, const LevelParams& lv, long long p) {
    return __ldg(P.coords + p * P.coord_stride + lv.dim[0]);
}

}  // namespace

template <typename TT, int F>
__global__ void __launch_bounds__(kThreads, 4)
small1d_forward_kernel(const __grid_constant__ EncodeParams P, const __grid_constant__ Small1dPlan S) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ LevelParams s_lv[kSmall1dMaxLevels];
    const int G = P.group_size;                       // power of two >= n_levels (<= 16)
    if (threadIdx.x < P.n_levels) s_lv[threadIdx.x] = P.levels[threadIdx.x];
    for (int l = 0; l < P.n_levels; ++l) {
        const int n = static_cast<int>(S.live[l]) * F;
        TT* dst = reinterpret_cast<TT*>(s_raw + S.off[l]);
        const TT* src = static_cast<const TT*>(P.levels[l].table);
        for (int i = threadIdx.x; i < n; i += kThreads) dst[i] = src[i];
    }
    __syncthreads();
    const int lane_level = threadIdx.x & (G - 1);
    if (lane_level >= P.n_levels) return;
    const LevelParams lv = s_lv[lane_level];
    const uint32_t live = S.live[lane_level];
    const TT* s_tab = reinterpret_cast<const TT*>(s_raw + S.off[lane_level]);
    const int p_step = kThreads / G;
    const long long p_end = min(P.n_points, (static_cast<long long>(blockIdx.x) + 1) * P.points_per_cta);
    // four points per trip: the coordinate loads of all four are in flight before the first dependent shared-memory read
    for (long long p = static_cast<long long>(blockIdx.x) * P.points_per_cta + threadIdx.x / G; p < p_end; p += 4 * p_step) {
        float t[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) t[u] = (p + u * p_step < p_end) ? coord_1d(P, lv, p + u * p_step) : 0.0f;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long pu = p + u * p_step;
            if (pu >= p_end) break;
            const uint32_t idx = index_1d(lv, t[u]);
            bool has_mask;
            const float m = row_mask(P, lv, pu, has_mask);
            float v[F];
            if (idx < live) {
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    if constexpr (sizeof(TT) == 4) v[f] = reinterpret_cast<const float*>(s_tab)[idx * F + f];
                    else v[f] = __half2float(reinterpret_cast<const __half*>(s_tab)[idx * F + f]);
                }
            } else {
                EntryLoader<TT, F>::load(lv.table, idx, v);
            }
#pragma unroll
            for (int f = 0; f < F; ++f) {
                if (P.out_scale != 1.0f) v[f] = __fmul_rn(v[f], P.out_scale);
                if (has_mask) v[f] = __fmul_rn(v[f], m);           // encoders.py:73-79
            }
            store_row<F>(P.out + pu * P.out_stride + lv.out_col, v, P.vec_rows != 0);
        }
    }
}

template <int F>
__global__ void __launch_bounds__(kThreads, 4)
small1d_backward_kernel(const __grid_constant__ EncodeParams P, const __grid_constant__ Small1dPlan S) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    float* s_acc = reinterpret_cast<float*>(s_raw);
    __shared__ LevelParams s_lv[kSmall1dMaxLevels];
    if (threadIdx.x < P.n_levels) s_lv[threadIdx.x] = P.levels[threadIdx.x];
    for (int i = threadIdx.x; i < static_cast<int>(S.total_bytes / 4); i += kThreads) s_acc[i] = 0.0f;
    __syncthreads();
    const int G = P.group_size;
    const int lane_level = threadIdx.x & (G - 1);
    if (lane_level < P.n_levels) {
        const LevelParams lv = s_lv[lane_level];
        const uint32_t live = S.live[lane_level];
        const int p_step = kThreads / G;
        const int slot = threadIdx.x / G;
        // private copy of the level's live rows for this point slot: the lanes of one warp that share a level differ in slot
        float* mine = s_acc + S.off[lane_level] / 4 + static_cast<size_t>(slot & (S.copies[lane_level] - 1)) * live * F;
        const long long p_end = min(P.n_points, (static_cast<long long>(blockIdx.x) + 1) * P.points_per_cta);
        for (long long p = static_cast<long long>(blockIdx.x) * P.points_per_cta + slot; p < p_end; p += 4 * p_step) {
            float t[4], go[4][F];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long pu = p + u * p_step;
                t[u] = 0.0f;
#pragma unroll
                for (int f = 0; f < F; ++f) go[u][f] = 0.0f;
                if (pu < p_end) {
                    t[u] = coord_1d(P, lv, pu);
                    load_row<F>(P.grad_out + pu * P.out_stride + lv.out_col, go[u], P.vec_rows != 0);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long pu = p + u * p_step;
                if (pu >= p_end) break;
                bool has_mask;
                const float m = row_mask(P, lv, pu, has_mask);
                if (has_mask && m == 0.0f) continue;
                const uint32_t idx = index_1d(lv, t[u]);
                if (idx < live) {
#pragma unroll
                    for (int f = 0; f < F; ++f) atomicAdd(mine + idx * F + f, go[u][f]);
                } else {
                    red_add_entry<F>(lv.grad, idx, go[u]);
                }
            }
        }
    }
    __syncthreads();
    for (int l = 0; l < P.n_levels; ++l) {
        const int n = static_cast<int>(S.live[l]) * F;
        const int copies = S.copies[l];
        const float* base = s_acc + S.off[l] / 4;
        float* grad = P.levels[l].grad;
        for (int i = threadIdx.x; i < n; i += kThreads) {
            float a = 0.0f;
            for (int c = 0; c < copies; ++c) a += base[c * n + i];
            if (a != 0.0f) red_add_f32(grad + i, a);
        }
    }
}

cudaError_t launch_small1d_forward(const EncodeParams& P, const Small1dPlan& S, bool half_tables, int F, int n_ctas, cudaStream_t s) {
#define DE_S1F(TT, FF) small1d_forward_kernel<TT, FF><<<n_ctas, kThreads, S.total_bytes, s>>>(P, S)
    if (half_tables) {
        switch (F) { case 1: DE_S1F(__half, 1); break; case 2: DE_S1F(__half, 2); break; case 4: DE_S1F(__half, 4); break;
                     case 8: DE_S1F(__half, 8); break; default: return cudaErrorInvalidValue; }
    } else {
        switch (F) { case 1: DE_S1F(float, 1); break; case 2: DE_S1F(float, 2); break; case 4: DE_S1F(float, 4); break;
                     case 8: DE_S1F(float, 8); break; default: return cudaErrorInvalidValue; }
    }
#undef DE_S1F
    return cudaGetLastError();
}

cudaError_t launch_small1d_backward(const EncodeParams& P, const Small1dPlan& S, int F, int n_ctas, cudaStream_t s) {
    switch (F) {
        case 1: small1d_backward_kernel<1><<<n_ctas, kThreads, S.total_bytes, s>>>(P, S); break;
        case 2: small1d_backward_kernel<2><<<n_ctas, kThreads, S.total_bytes, s>>>(P, S); break;
        case 4: small1d_backward_kernel<4><<<n_ctas, kThreads, S.total_bytes, s>>>(P, S); break;
        case 8: small1d_backward_kernel<8><<<n_ctas, kThreads, S.total_bytes, s>>>(P, S); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace de
