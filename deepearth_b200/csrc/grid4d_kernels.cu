// grid4d_kernels.cu -- forward / backward / index kernels of the Grid4D hash-grid encoder (sm_100a).
//
// Work decomposition (DESIGN.md "kernels"):
//   * a launch carries up to 64 levels that share the grid dimensionality D;
//   * levels are cut into GROUPS of G consecutive levels; blockIdx.x = group * n_tiles + tile, so the
//     hardware's in-order CTA dispatch walks one group's point tiles before the next group's --
//     the G tables (and, in backward, the G gradient tables) being hammered at any moment stay L2
//     resident even when the whole encoder's tables do not fit the 126 MB L2;
//   * inside a CTA, thread t handles level (t % G) of point (t / G): the G lanes of a point write
//     G*F consecutive floats of its feature row (one full 32 B sector for G=4, F=2), coordinates are
//     one float4 per point served to those lanes from L1, and every lane keeps 2^D independent
//     table gathers (or vector reductions) in flight.
// The gathers / reductions are the roofline: everything else is a handful of integer ops.
#include "grid4d_common.cuh"
#include "launch.h"

namespace de {

template <int D, bool LINEAR>
struct Corners {
    static constexpr int value = LINEAR ? (1 << D) : 1;
};

__device__ __forceinline__ bool setup_thread(const EncodeParams& P, LevelParams* s_lv, int& lane_level,
                                             long long& p0, int& p_step, LevelParams& lv) {
    const int G = P.group_size;
    const int group = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - group * P.n_tiles;
    const int first = group * G;
    if (threadIdx.x < G && first + threadIdx.x < P.n_levels) s_lv[threadIdx.x] = P.levels[first + threadIdx.x];
    __syncthreads();
    lane_level = threadIdx.x & (G - 1);
    if (first + lane_level >= P.n_levels) return false;
    lv = s_lv[lane_level];
    p_step = kThreads / G;
    p0 = static_cast<long long>(tile) * P.points_per_cta + threadIdx.x / G;
    return true;
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <int D, bool LINEAR, typename TT, int F>
__global__ void __launch_bounds__(kThreads, LINEAR ? 1 : 8)
grid4d_forward_kernel(const __grid_constant__ EncodeParams P) {
    __shared__ LevelParams s_lv[kMaxGroup];
    int lane_level, p_step;
    long long p0;
    LevelParams lv;
    if (!setup_thread(P, s_lv, lane_level, p0, p_step, lv)) return;
    constexpr int C = Corners<D, LINEAR>::value;
    const long long p_end = min(P.n_points, (static_cast<long long>(blockIdx.x % P.n_tiles) + 1) * P.points_per_cta);

    for (long long p = p0; p < p_end; p += p_step) {
        Cell<D> cell;
        load_cell<D>(P, lv, p, cell);
        bool has_mask;
        const float m = row_mask(P, lv, p, has_mask);
        uint32_t idx[C];
#pragma unroll
        for (int k = 0; k < C; ++k) idx[k] = corner_index<D>(lv, cell, k);
        float* dst = P.out + p * P.out_stride + lv.out_col;

        if constexpr (F > 0) {
            float v[C][F];
#pragma unroll
            for (int k = 0; k < C; ++k) EntryLoader<TT, F>::load(lv.table, idx[k], v[k]);  // all gathers first
            float acc[F];
            if constexpr (LINEAR) {
#pragma unroll
                for (int f = 0; f < F; ++f) acc[f] = 0.0f;
#pragma unroll
                for (int k = 0; k < C; ++k) {
                    const float wk = corner_weight<D>(cell, k);
#pragma unroll
                    for (int f = 0; f < F; ++f) acc[f] = __fmaf_rn(wk, v[k][f], acc[f]);
                }
            } else {
#pragma unroll
                for (int f = 0; f < F; ++f) acc[f] = v[0][f];
            }
#pragma unroll
            for (int f = 0; f < F; ++f) {
                if (P.out_scale != 1.0f) acc[f] = __fmul_rn(acc[f], P.out_scale);
                if (has_mask) acc[f] = __fmul_rn(acc[f], m);   // encoders.py:73-79
            }
            store_row<F>(dst, acc, P.vec_rows != 0);
        } else {
            // runtime feature count: scalar element loads (rare shapes such as F=3)
            const int nf = P.n_features;
            float wk[C];
#pragma unroll
            for (int k = 0; k < C; ++k) wk[k] = LINEAR ? corner_weight<D>(cell, k) : 1.0f;
            for (int f = 0; f < nf; ++f) {
                float acc = 0.0f;
#pragma unroll
                for (int k = 0; k < C; ++k) {
                    const float val = load_elem<TT>(lv.table, static_cast<unsigned long long>(idx[k]) * nf + f);
                    acc = LINEAR ? __fmaf_rn(wk[k], val, acc) : val;
                }
                if (P.out_scale != 1.0f) acc = __fmul_rn(acc, P.out_scale);
                if (has_mask) acc = __fmul_rn(acc, m);
                __stcs(dst + f, acc);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// backward: grad_table[idx_k] += mask * w_k * grad_out[p, out_col:out_col+F]
// ------------------------------------------------------------------------------------------------
template <int D, bool LINEAR, int F>
__global__ void __launch_bounds__(kThreads, LINEAR ? 1 : 8)
grid4d_backward_kernel(const __grid_constant__ EncodeParams P) {
    __shared__ LevelParams s_lv[kMaxGroup];
    int lane_level, p_step;
    long long p0;
    LevelParams lv;
    if (!setup_thread(P, s_lv, lane_level, p0, p_step, lv)) return;
    constexpr int C = Corners<D, LINEAR>::value;
    const long long p_end = min(P.n_points, (static_cast<long long>(blockIdx.x % P.n_tiles) + 1) * P.points_per_cta);
    const GradTarget gt = grad_target(lv, blockIdx.x % P.n_tiles);

    for (long long p = p0; p < p_end; p += p_step) {
        bool has_mask;
        const float m = row_mask(P, lv, p, has_mask);
        if (has_mask && m == 0.0f) continue;   // masked rows contribute exact zeros
        Cell<D> cell;
        load_cell<D>(P, lv, p, cell);
        const float* src = P.grad_out + p * P.out_stride + lv.out_col;
        if constexpr (F > 0) {
            float go[F];
            load_row<F>(src, go, P.vec_rows != 0);
#pragma unroll
            for (int k = 0; k < C; ++k) {
                const uint32_t idx = corner_index<D>(lv, cell, k);
                float val[F];
                if constexpr (LINEAR) {
                    const float wk = corner_weight<D>(cell, k);
#pragma unroll
                    for (int f = 0; f < F; ++f) val[f] = __fmul_rn(wk, go[f]);
                } else {
#pragma unroll
                    for (int f = 0; f < F; ++f) val[f] = go[f];
                }
                red_add_entry<F>(idx < gt.live ? gt.base : lv.grad, idx, val);
            }
        } else {
            const int nf = P.n_features;
#pragma unroll
            for (int k = 0; k < C; ++k) {
                const uint32_t idx = corner_index<D>(lv, cell, k);
                const float wk = LINEAR ? corner_weight<D>(cell, k) : 1.0f;
                float* dst = (idx < gt.live ? gt.base : lv.grad) + static_cast<unsigned long long>(idx) * nf;
                for (int f = 0; f < nf; ++f) red_add_f32(dst + f, LINEAR ? __fmul_rn(wk, __ldcs(src + f)) : __ldcs(src + f));
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// indices[level][corner][point] (int64) -- the public hash_coords and the bit-exactness tests
// ------------------------------------------------------------------------------------------------
template <int D, bool LINEAR>
__global__ void __launch_bounds__(kThreads, LINEAR ? 1 : 8)
grid4d_indices_kernel(const __grid_constant__ EncodeParams P) {
    __shared__ LevelParams s_lv[kMaxGroup];
    int lane_level, p_step;
    long long p0;
    LevelParams lv;
    if (!setup_thread(P, s_lv, lane_level, p0, p_step, lv)) return;
    constexpr int C = Corners<D, LINEAR>::value;
    const long long p_end = min(P.n_points, (static_cast<long long>(blockIdx.x % P.n_tiles) + 1) * P.points_per_cta);
    const int level = (blockIdx.x / P.n_tiles) * P.group_size + lane_level;
    for (long long p = p0; p < p_end; p += p_step) {
        Cell<D> cell;
        load_cell<D>(P, lv, p, cell);
#pragma unroll
        for (int k = 0; k < C; ++k)
            P.indices[(static_cast<long long>(level) * C + k) * P.n_points + p] = corner_index<D>(lv, cell, k);
    }
}

// ------------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------------
template <typename K>
static cudaError_t launch(K kernel, const EncodeParams& P, cudaStream_t stream) {
    const long long blocks = static_cast<long long>(P.n_groups) * P.n_tiles;
    kernel<<<static_cast<unsigned>(blocks), kThreads, 0, stream>>>(P);
    return cudaGetLastError();
}

template <int D, bool LINEAR, typename TT>
static cudaError_t fwd_f(const EncodeParams& P, int F, cudaStream_t s) {
    switch (F) {
        case 1: return launch(grid4d_forward_kernel<D, LINEAR, TT, 1>, P, s);
        case 2: return launch(grid4d_forward_kernel<D, LINEAR, TT, 2>, P, s);
        case 4: return launch(grid4d_forward_kernel<D, LINEAR, TT, 4>, P, s);
        case 8: return launch(grid4d_forward_kernel<D, LINEAR, TT, 8>, P, s);
        default: return launch(grid4d_forward_kernel<D, LINEAR, TT, 0>, P, s);
    }
}

template <int D, bool LINEAR>
static cudaError_t fwd_t(const EncodeParams& P, int F, bool half_tables, cudaStream_t s) {
    return half_tables ? fwd_f<D, LINEAR, __half>(P, F, s) : fwd_f<D, LINEAR, float>(P, F, s);
}

cudaError_t launch_forward(const EncodeParams& P, int D, bool linear, bool half_tables, int F, cudaStream_t s) {
    switch (D * 2 + (linear ? 1 : 0)) {
        case 2: return fwd_t<1, false>(P, F, half_tables, s);
        case 3: return fwd_t<1, true>(P, F, half_tables, s);
        case 4: return fwd_t<2, false>(P, F, half_tables, s);
        case 5: return fwd_t<2, true>(P, F, half_tables, s);
        case 6: return fwd_t<3, false>(P, F, half_tables, s);
        case 7: return fwd_t<3, true>(P, F, half_tables, s);
        case 8: return fwd_t<4, false>(P, F, half_tables, s);
        case 9: return fwd_t<4, true>(P, F, half_tables, s);
        default: return cudaErrorInvalidValue;
    }
}

template <int D, bool LINEAR>
static cudaError_t bwd_f(const EncodeParams& P, int F, cudaStream_t s) {
    switch (F) {
        case 1: return launch(grid4d_backward_kernel<D, LINEAR, 1>, P, s);
        case 2: return launch(grid4d_backward_kernel<D, LINEAR, 2>, P, s);
        case 4: return launch(grid4d_backward_kernel<D, LINEAR, 4>, P, s);
        case 8: return launch(grid4d_backward_kernel<D, LINEAR, 8>, P, s);
        default: return launch(grid4d_backward_kernel<D, LINEAR, 0>, P, s);
    }
}

cudaError_t launch_backward(const EncodeParams& P, int D, bool linear, int F, cudaStream_t s) {
    switch (D * 2 + (linear ? 1 : 0)) {
        case 2: return bwd_f<1, false>(P, F, s);
        case 3: return bwd_f<1, true>(P, F, s);
        case 4: return bwd_f<2, false>(P, F, s);
        case 5: return bwd_f<2, true>(P, F, s);
        case 6: return bwd_f<3, false>(P, F, s);
        case 7: return bwd_f<3, true>(P, F, s);
        case 8: return bwd_f<4, false>(P, F, s);
        case 9: return bwd_f<4, true>(P, F, s);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_indices(const EncodeParams& P, int D, bool linear, cudaStream_t s) {
    switch (D * 2 + (linear ? 1 : 0)) {
        case 2: return launch(grid4d_indices_kernel<1, false>, P, s);
        case 3: return launch(grid4d_indices_kernel<1, true>, P, s);
        case 4: return launch(grid4d_indices_kernel<2, false>, P, s);
        case 5: return launch(grid4d_indices_kernel<2, true>, P, s);
        case 6: return launch(grid4d_indices_kernel<3, false>, P, s);
        case 7: return launch(grid4d_indices_kernel<3, true>, P, s);
        case 8: return launch(grid4d_indices_kernel<4, false>, P, s);
        case 9: return launch(grid4d_indices_kernel<4, true>, P, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
