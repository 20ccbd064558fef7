// grid4d_sorted_kernels.cu -- the north_star's alternative for densely colliding levels (sm_100a): order the points by
// coarse cell first, then scatter with WARP-AGGREGATED reductions.  Opt-in (de_grid4d_sort_points +
// de_grid4d_backward_sorted); the default path handles the same levels with gradient replicas, and
// profiles/ holds the A/B between the two.
//
//   sort:      counting sort of the points by the cell they occupy on a 2^bits-per-axis grid over the level list's own
//              coordinate columns (x-major key) -- histogram, one-CTA exclusive scan, scatter; perm[i] = i-th point in
//              cell order.  The order inside a cell is whatever the scatter's atomic cursor produced: any order is valid.
//   backward:  a warp takes 32 CONSECUTIVE sorted points of ONE level (the pair kernel: 16 levels of one point).  Points
//              of one cell sit next to each other, so whole runs of lanes address the same 2^D corners: the lanes of a run
//              add their contributions with a segmented shuffle scan (5 steps) and only the last lane of each run issues
//              the reduction -- up to 32x fewer red.global operations on the contended rows.
// Index and weight arithmetic are the pair kernel's (grid4d_pair.cuh); gradients equal the default path's up to fp32
// summation order.
#include "grid4d_common.cuh"
#include "grid4d_pair.cuh"
#include "launch.h"

namespace de {

namespace {

__device__ __forceinline__ uint32_t sort_key(const float* __restrict__ row, const SortParams& S) {
    uint32_t key = 0;
    const float cells = static_cast<float>(1u << S.bits);
#pragma unroll 4
    for (int d = 0; d < S.n_dims; ++d) {
        float c = __ldg(row + S.dim[d]) * cells;
        c = fminf(fmaxf(c, 0.0f), cells - 1.0f);                  // outside [0, 1): first / last cell (NaN -> 0)
        key = (key << S.bits) | static_cast<uint32_t>(c);
    }
    return key;
}

}  // namespace

__global__ void __launch_bounds__(256) sort_histogram_kernel(const SortParams S) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < S.n; p += stride)
        atomicAdd(S.counts + sort_key(S.coords + p * S.coord_stride, S), 1u);
}

// exclusive scan of n_buckets counters by one CTA (n_buckets <= 2^16: 64 per thread)
__global__ void __launch_bounds__(1024) sort_scan_kernel(uint32_t* counts, uint32_t n_buckets) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_buckets; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n_buckets ? counts[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = s_warp[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (threadIdx.x >= o) w += y;
            }
            s_warp[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t warp_off = (threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u;
        const uint32_t carry = s_carry;
        if (i < n_buckets) counts[i] = carry + warp_off + x - v;               // exclusive
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + warp_off + x;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) sort_scatter_kernel(const SortParams S) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < S.n; p += stride) {
        const uint32_t pos = atomicAdd(S.counts + sort_key(S.coords + p * S.coord_stride, S), 1u);
        S.perm[pos] = static_cast<int>(p);
    }
}

cudaError_t launch_sort_points(const SortParams& S, int n_sms, cudaStream_t s) {
    if (S.n == 0) return cudaSuccess;
    const uint32_t n_buckets = 1u << (S.bits * S.n_dims);
    cudaError_t e = cudaMemsetAsync(S.counts, 0, static_cast<size_t>(n_buckets) * 4, s);
    if (e != cudaSuccess) return e;
    long long blocks = (S.n + 255) / 256;
    if (blocks > n_sms * 8) blocks = n_sms * 8;
    sort_histogram_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(S);
    sort_scan_kernel<<<1, 1024, 0, s>>>(S.counts, n_buckets);
    sort_scatter_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(S);
    return cudaGetLastError();
}

// ---- warp-aggregated backward --------------------------------------------------------------------------------------
// blockIdx.y = level; a warp walks 32 sorted points per trip.  D = 3, multilinear, F in {1, 2, 4}.
template <int F, bool POW2>
__global__ void __launch_bounds__(kThreads)
grid4d_backward_sorted_kernel(const __grid_constant__ EncodeParams P, const int* __restrict__ perm) {
    constexpr int D = 3;
    __shared__ LevelParams s_lv;
    if (threadIdx.x == 0) s_lv = P.levels[blockIdx.y];
    __syncthreads();
    const LevelParams lv = s_lv;
    const int lane = threadIdx.x & 31;
    const long long warp_global = static_cast<long long>(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
    const long long n_warps = static_cast<long long>(gridDim.x) * (kThreads / 32);
    const bool vec = P.vec_rows != 0;
    for (long long base = warp_global * 32; base < P.n_points; base += n_warps * 32) {
        const long long i = base + lane;
        const bool live = i < P.n_points;
        long long g[D] = {0, 0, 0};
        float w[D] = {0.f, 0.f, 0.f};
        float go[F];
#pragma unroll
        for (int f = 0; f < F; ++f) go[f] = 0.0f;
        if (live) {
            const long long p = __ldg(perm + i);
            float x[D];
            load_coords<D>(P, lv, p, x);
            cell_of<D>(lv, x, g, w);
            load_row<F>(P.grad_out + p * P.out_stride + lv.out_col, go, vec);
            bool has_mask;
            const float m = row_mask(P, lv, p, has_mask);
            if (has_mask && m == 0.0f) {
#pragma unroll
                for (int f = 0; f < F; ++f) go[f] = 0.0f;
            }
        }
        // runs of lanes in the same cell: a lane starts a run when its cell differs from its left neighbour's
        bool head = lane == 0;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const long long left = __shfl_up_sync(0xffffffffu, g[d], 1);
            head = head || (left != g[d]);
        }
        const bool left_live = __shfl_up_sync(0xffffffffu, live ? 1 : 0, 1) != 0;
        head = head || (live != left_live);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const int run_start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
        const bool tail = lane == 31 || ((heads >> (lane + 1)) & 1u);
        // the lanes of a run share cell and level, hence the corner indices; their weights differ
        CellTerms<D, POW2> c0, c1;
        c0.set(lv, g, 0);
        c1.set(lv, g, 1);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int ox = k & 1, j = k >> 1;
            const float wk = pair_weight<D>(w, ox, j);
            float val[F];
#pragma unroll
            for (int f = 0; f < F; ++f) val[f] = __fmul_rn(wk, go[f]);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    const float up = __shfl_up_sync(0xffffffffu, val[f], o);
                    if (lane - o >= run_start) val[f] += up;
                }
            }
            if (tail && live) {
                const uint32_t idx = ox ? c1.index(lv, j) : c0.index(lv, j);
                red_add_entry<F>(lv.grad, idx, val);
            }
        }
    }
}

cudaError_t launch_backward_sorted(const EncodeParams& P, const int* perm, int F, bool pow2, int n_sms, cudaStream_t s) {
    long long bx = (P.n_points + kThreads - 1) / kThreads;
    const long long cap = static_cast<long long>(n_sms) * 8 / (P.n_levels > 0 ? P.n_levels : 1) + 1;
    if (bx > cap) bx = cap;
    if (bx < 1) bx = 1;
    const dim3 grid(static_cast<unsigned>(bx), static_cast<unsigned>(P.n_levels));
#define DE_SORTED(FF)                                                                                             \
    if (pow2) grid4d_backward_sorted_kernel<FF, true><<<grid, kThreads, 0, s>>>(P, perm);                        \
    else grid4d_backward_sorted_kernel<FF, false><<<grid, kThreads, 0, s>>>(P, perm)
    switch (F) {
        case 1: DE_SORTED(1); break;
        case 2: DE_SORTED(2); break;
        case 4: DE_SORTED(4); break;
        default: return cudaErrorInvalidValue;
    }
#undef DE_SORTED
    return cudaGetLastError();
}

}  // namespace de
