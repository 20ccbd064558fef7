// grid4d_pair_kernels.cu -- multilinear (2^D-corner) forward / backward kernels, "lane-pair" layout (sm_100a).
//
// Why this layout (profiles/r01b_peaks_gather.json, measured on B200):
//   * the SM's L1 -> L2 path moves one 32 B sector packet per ~1.3 cycles; a gather or a vector
//     reduction costs one packet per DISTINCT sector an instruction touches, not per lane;
//   * the two x-neighbour corners of a cell are neighbouring table entries (prime[0] = 1,
//     hash_encoding.py:64; dense stride 1), i.e. the same sector 3/4 (8 B grads) to 7/8 (4 B fp16 entries)
//     of the time -- but only lanes of the SAME instruction are merged into one packet.
// So two adjacent lanes share one (point, level): lane bit 0 is the x offset of the corners the lane
// owns (2^(D-1) of them).  Random red.add.v2.f32 runs at ~219 G lane-ops/s, lane pairs on adjacent
// entries at ~430 G/s; 4 B gathers go from ~290 (random) to ~550 G/s.
//
// Thread t of a 256-thread CTA: ox = t & 1, level = (t >> 1) % G of the CTA's level group, point slot
// = t / (2G).  CTAs are ordered group-major (blockIdx.x = group * n_tiles + tile) so that only the G
// tables (gradient tables) of one group are hot in L2 at a time.
//
// Arithmetic contract: per-corner index and weight are exactly the oracle's (grid4d_common.cuh header);
// the forward sum runs over the lane's corners in ascending order and then adds the partner lane's
// partial sum, so features agree with the oracle to fp32 rounding (tests: 1e-5 of max|ref|), not bit-exactly.
#include "grid4d_common.cuh"
#include "grid4d_pair.cuh"
#include "launch.h"

namespace de {

namespace {

struct PairThread {
    int ox, lane_level, p_step;
    long long p0, p_end;
    unsigned pair_mask;
};

__device__ __forceinline__ bool pair_setup(const EncodeParams& P, LevelParams* s_lv, PairThread& t, LevelParams& lv) {
    const int G = P.group_size;
    const int group = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - group * P.n_tiles;
    const int first = group * G;
    if (threadIdx.x < G && first + threadIdx.x < P.n_levels) s_lv[threadIdx.x] = P.levels[first + threadIdx.x];
    __syncthreads();
    t.ox = threadIdx.x & 1;
    t.lane_level = (threadIdx.x >> 1) & (G - 1);
    if (first + t.lane_level >= P.n_levels) return false;   // both lanes of a pair leave together
    lv = s_lv[t.lane_level];
    t.p_step = kThreads / (2 * G);
    t.p0 = static_cast<long long>(tile) * P.points_per_cta + threadIdx.x / (2 * G);
    t.p_end = min(P.n_points, (static_cast<long long>(tile) + 1) * P.points_per_cta);
    t.pair_mask = 3u << (threadIdx.x & 30);
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <int D, typename TT, int F, bool POW2>
__global__ void __launch_bounds__(kThreads, 8)
grid4d_forward_pair_kernel(const __grid_constant__ EncodeParams P) {
    __shared__ LevelParams s_lv[kMaxGroup];
    PairThread t;
    LevelParams lv;
    if (!pair_setup(P, s_lv, t, lv)) return;
    constexpr int J = PairGeom<D>::kCombos;
    const L2Policies pol = make_l2_policies(P.l2_hints != 0);

    for (long long p = t.p0; p < t.p_end; p += t.p_step) {
        float x[D];
        load_coords_stream<D>(P, lv, p, x, pol.stream);
        long long g[D];
        float w[D];
        cell_of<D>(lv, x, g, w);
        CellTerms<D, POW2> ct;
        ct.set(lv, g, t.ox);
        bool has_mask;
        const float m = row_mask(P, lv, p, has_mask);

        uint32_t idx[J];
#pragma unroll
        for (int j = 0; j < J; ++j) idx[j] = ct.index(lv, j);
        float v[J][F];
#pragma unroll
        for (int j = 0; j < J; ++j) load_entry_hint<TT, F>(lv.table, idx[j], v[j], pol.keep, P.l2_hints == 2);   // all gathers in flight
        float acc[F];
#pragma unroll
        for (int f = 0; f < F; ++f) acc[f] = 0.0f;
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const float wk = pair_weight<D>(w, t.ox, j);
#pragma unroll
            for (int f = 0; f < F; ++f) acc[f] = __fmaf_rn(wk, v[j][f], acc[f]);
        }
#pragma unroll
        for (int f = 0; f < F; ++f) {
            acc[f] = __fadd_rn(acc[f], __shfl_xor_sync(t.pair_mask, acc[f], 1));   // partner's corners (commutative: both lanes agree)
            if (P.out_scale != 1.0f) acc[f] = __fmul_rn(acc[f], P.out_scale);
            if (has_mask) acc[f] = __fmul_rn(acc[f], m);                            // encoders.py:73-79
        }
        float* dst = P.out + p * P.out_stride + lv.out_col;
        pair_store<F>(dst, acc, t.ox, P.vec_rows != 0);
    }
}

// ------------------------------------------------------------------------------------------------
// backward: grad_table[idx_k] += mask * w_k * grad_out[p, out_col:out_col+F]; the next point's
// coordinates and gradient row are fetched before the current point's reductions are issued.
// ------------------------------------------------------------------------------------------------
template <int D, int F, bool POW2>
__global__ void __launch_bounds__(kThreads)
grid4d_backward_pair_kernel(const __grid_constant__ EncodeParams P) {
    __shared__ LevelParams s_lv[kMaxGroup];
    PairThread t;
    LevelParams lv;
    if (!pair_setup(P, s_lv, t, lv)) return;
    constexpr int J = PairGeom<D>::kCombos;
    const bool vec = P.vec_rows != 0;
    const GradTarget gt = grad_target(lv, blockIdx.x % P.n_tiles);
    const L2Policies pol = make_l2_policies(P.l2_hints != 0);

    float x_n[D], go_n[F];
    if (t.p0 < t.p_end) {
        load_coords_stream<D>(P, lv, t.p0, x_n, pol.stream);
        load_row_stream<F>(P.grad_out + t.p0 * P.out_stride + lv.out_col, go_n, vec, pol.stream);
    }
    for (long long p = t.p0; p < t.p_end; p += t.p_step) {
        float x[D], go[F];
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = x_n[d];
#pragma unroll
        for (int f = 0; f < F; ++f) go[f] = go_n[f];
        const long long pn = p + t.p_step;
        if (pn < t.p_end) {
            load_coords_stream<D>(P, lv, pn, x_n, pol.stream);
            load_row_stream<F>(P.grad_out + pn * P.out_stride + lv.out_col, go_n, vec, pol.stream);
        }
        bool has_mask;
        const float m = row_mask(P, lv, p, has_mask);
        if (has_mask && m == 0.0f) continue;   // masked rows contribute exact zeros
        long long g[D];
        float w[D];
        cell_of<D>(lv, x, g, w);
        CellTerms<D, POW2> ct;
        ct.set(lv, g, t.ox);
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const uint32_t idx = ct.index(lv, j);
            const float wk = pair_weight<D>(w, t.ox, j);
            float val[F];
#pragma unroll
            for (int f = 0; f < F; ++f) val[f] = __fmul_rn(wk, go[f]);
            red_add_entry_hint<F>(idx < gt.live ? gt.base : lv.grad, idx, val, pol.keep);
        }
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

template <int D, typename TT, bool POW2>
static cudaError_t fwd_pair_f(const EncodeParams& P, int F, cudaStream_t s) {
    switch (F) {
        case 1: return launch(grid4d_forward_pair_kernel<D, TT, 1, POW2>, P, s);
        case 2: return launch(grid4d_forward_pair_kernel<D, TT, 2, POW2>, P, s);
        case 4: return launch(grid4d_forward_pair_kernel<D, TT, 4, POW2>, P, s);
        case 8: return launch(grid4d_forward_pair_kernel<D, TT, 8, POW2>, P, s);
        default: return cudaErrorInvalidValue;
    }
}

template <int D>
static cudaError_t fwd_pair_d(const EncodeParams& P, bool half_tables, int F, bool pow2, cudaStream_t s) {
    if (half_tables) return pow2 ? fwd_pair_f<D, __half, true>(P, F, s) : fwd_pair_f<D, __half, false>(P, F, s);
    return pow2 ? fwd_pair_f<D, float, true>(P, F, s) : fwd_pair_f<D, float, false>(P, F, s);
}

bool pair_layout_supported(bool linear, int F) { return linear && (F == 1 || F == 2 || F == 4 || F == 8); }

cudaError_t launch_forward_pair(const EncodeParams& P, int D, bool half_tables, int F, bool pow2, cudaStream_t s) {
    switch (D) {
        case 1: return fwd_pair_d<1>(P, half_tables, F, pow2, s);
        case 2: return fwd_pair_d<2>(P, half_tables, F, pow2, s);
        case 3: return fwd_pair_d<3>(P, half_tables, F, pow2, s);
        case 4: return fwd_pair_d<4>(P, half_tables, F, pow2, s);
        default: return cudaErrorInvalidValue;
    }
}

template <int D, bool POW2>
static cudaError_t bwd_pair_f(const EncodeParams& P, int F, cudaStream_t s) {
    switch (F) {
        case 1: return launch(grid4d_backward_pair_kernel<D, 1, POW2>, P, s);
        case 2: return launch(grid4d_backward_pair_kernel<D, 2, POW2>, P, s);
        case 4: return launch(grid4d_backward_pair_kernel<D, 4, POW2>, P, s);
        case 8: return launch(grid4d_backward_pair_kernel<D, 8, POW2>, P, s);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_backward_pair(const EncodeParams& P, int D, int F, bool pow2, cudaStream_t s) {
    switch (D * 2 + (pow2 ? 1 : 0)) {
        case 2: return bwd_pair_f<1, false>(P, F, s);
        case 3: return bwd_pair_f<1, true>(P, F, s);
        case 4: return bwd_pair_f<2, false>(P, F, s);
        case 5: return bwd_pair_f<2, true>(P, F, s);
        case 6: return bwd_pair_f<3, false>(P, F, s);
        case 7: return bwd_pair_f<3, true>(P, F, s);
        case 8: return bwd_pair_f<4, false>(P, F, s);
        case 9: return bwd_pair_f<4, true>(P, F, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
