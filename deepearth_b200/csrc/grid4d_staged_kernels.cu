// grid4d_staged_kernels.cu -- multilinear backward with TMA-staged inputs (sm_100a).
//
// Why (profiles/r01b_kernels_c2.txt, r01b_stalls_c2.txt): the lane-pair backward kernel spends ~65 % of its
// warp-stall samples on the first use of the streamed inputs (coordinates + gradient row).  Those loads
// enter the SM's LSU queue BEHIND the vector reductions the same warps keep issuing, so their latency is
// the drain time of that queue and a one-iteration register prefetch does not cover it: the L1->L2
// reduction path ran at 0.47 packets/clk/SM against 0.74 measured by the probe (r01b_peaks_gather.json).
//
// Here the streamed inputs never touch the LSU load path: per stage, one bulk-async copy (TMA engine,
// cp.async.bulk global -> shared, mbarrier complete_tx) brings the coordinates of TP points, and one copy
// per point brings that point's G*F gradient floats of the CTA's level group.  A ring of S stages keeps
// S-1 tiles in flight while every thread does nothing but index arithmetic and red.global.add.v2.f32.
//
// Thread mapping is the lane-pair layout of grid4d_pair_kernels.cu (lane bit 0 = x offset of the corners
// the lane owns, then level within the group, then point slot); arithmetic contract unchanged.
#include "grid4d_common.cuh"
#include "grid4d_pair.cuh"
#include "launch.h"

namespace de {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
// global -> shared bulk copy through the TMA engine; bytes % 16 == 0, both addresses 16 B aligned
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

constexpr int kMaxStages = 8;

}  // namespace

// dynamic shared memory per stage: float4 coords[TP] | float grad[TP][G*F]
template <int D, int F, bool POW2>
__global__ void __launch_bounds__(kThreads)
grid4d_backward_staged_kernel(const __grid_constant__ EncodeParams P) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ LevelParams s_lv[kMaxGroup];
    __shared__ __align__(8) uint64_t s_full[kMaxStages];

    const int G = P.group_size;
    const int group = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - group * P.n_tiles;
    const int first = group * G;
    const int n_in = min(G, P.n_levels - first);                // levels of this group (last group may be short)
    const int S = P.n_stages;
    const int TP = P.stage_points;
    const int GF = G * F;
    const uint32_t row_bytes = static_cast<uint32_t>(n_in) * F * 4u;
    const size_t stage_bytes = static_cast<size_t>(TP) * (16 + GF * 4);

    if (threadIdx.x < n_in) s_lv[threadIdx.x] = P.levels[first + threadIdx.x];
    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&s_full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const long long cta_p0 = static_cast<long long>(tile) * P.points_per_cta;
    const long long cta_p1 = min(P.n_points, cta_p0 + P.points_per_cta);
    const int n_k = static_cast<int>((cta_p1 - cta_p0 + TP - 1) / TP);
    const int col0 = s_lv[0].out_col;   // group columns are contiguous (host-checked)

    auto fill = [&](int k) {
        const int s = k % S;
        unsigned char* st = smem_raw + s * stage_bytes;
        const long long pb = cta_p0 + static_cast<long long>(k) * TP;
        const int rows = static_cast<int>(min(static_cast<long long>(TP), cta_p1 - pb));
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(&s_full[s], static_cast<uint32_t>(rows) * (16u + row_bytes));
            bulk_g2s(st, reinterpret_cast<const float4*>(P.coords) + pb, static_cast<uint32_t>(rows) * 16u, &s_full[s]);
        }
        float* gs = reinterpret_cast<float*>(st + static_cast<size_t>(TP) * 16);
        for (int r = threadIdx.x; r < rows; r += kThreads)
            bulk_g2s(gs + static_cast<size_t>(r) * GF, P.grad_out + (pb + r) * P.out_stride + col0, row_bytes, &s_full[s]);
    };

    for (int k = 0; k < S - 1 && k < n_k; ++k) fill(k);

    const int ox = threadIdx.x & 1;
    const int lane_level = (threadIdx.x >> 1) & (G - 1);
    const bool active = lane_level < n_in;
    const int p_step = kThreads / (2 * G);
    const int slot0 = threadIdx.x / (2 * G);
    LevelParams lv = s_lv[active ? lane_level : 0];
    constexpr int J = PairGeom<D>::kCombos;
    const GradTarget gt = grad_target(lv, tile);

    for (int k = 0; k < n_k; ++k) {
        // refill the stage consumed in iteration k-1 (every thread passed the __syncthreads below)
        if (k + S - 1 < n_k) fill(k + S - 1);
        const int s = k % S;
        mbar_wait(&s_full[s], static_cast<uint32_t>(k / S) & 1u);
        const unsigned char* st = smem_raw + s * stage_bytes;
        const float4* cs = reinterpret_cast<const float4*>(st);
        const float* gs = reinterpret_cast<const float*>(st + static_cast<size_t>(TP) * 16);
        const long long pb = cta_p0 + static_cast<long long>(k) * TP;
        const int rows = static_cast<int>(min(static_cast<long long>(TP), cta_p1 - pb));
        if (active) {
#pragma unroll 2
            for (int slot = slot0; slot < rows; slot += p_step) {
                const float4 c4 = cs[slot];
                float x[D], go[F];
#pragma unroll
                for (int d = 0; d < D; ++d) x[d] = pick4(c4, lv.dim[d]);
                const float* gp = gs + static_cast<size_t>(slot) * GF + lane_level * F;
                if constexpr (F == 2) {
                    const float2 v = *reinterpret_cast<const float2*>(gp);
                    go[0] = v.x; go[1] = v.y;
                } else if constexpr (F == 4) {
                    const float4 v = *reinterpret_cast<const float4*>(gp);
                    go[0] = v.x; go[1] = v.y; go[2] = v.z; go[3] = v.w;
                } else {
#pragma unroll
                    for (int f = 0; f < F; ++f) go[f] = gp[f];
                }
                bool has_mask;
                const float m = row_mask(P, lv, pb + slot, has_mask);
                if (has_mask && m == 0.0f) continue;   // masked rows contribute exact zeros
                long long g[D];
                float w[D];
                cell_of<D>(lv, x, g, w);
                CellTerms<D, POW2> ct;
                ct.set(lv, g, ox);
#pragma unroll
                for (int j = 0; j < J; ++j) {
                    const uint32_t idx = ct.index(lv, j);
                    const float wk = pair_weight<D>(w, ox, j);
                    float val[F];
#pragma unroll
                    for (int f = 0; f < F; ++f) val[f] = __fmul_rn(wk, go[f]);
                    red_add_entry<F>(idx < gt.live ? gt.base : lv.grad, idx, val);
                }
            }
        }
        __syncthreads();   // stage s fully consumed: it may be refilled at the top of the next iteration
    }
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
template <int D, int F, bool POW2>
static cudaError_t launch_staged(const EncodeParams& P, size_t smem, cudaStream_t s) {
    auto kernel = grid4d_backward_staged_kernel<D, F, POW2>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    const long long blocks = static_cast<long long>(P.n_groups) * P.n_tiles;
    kernel<<<static_cast<unsigned>(blocks), kThreads, smem, s>>>(P);
    return cudaGetLastError();
}

template <int D, bool POW2>
static cudaError_t staged_f(const EncodeParams& P, int F, size_t smem, cudaStream_t s) {
    switch (F) {
        case 1: return launch_staged<D, 1, POW2>(P, smem, s);
        case 2: return launch_staged<D, 2, POW2>(P, smem, s);
        case 4: return launch_staged<D, 4, POW2>(P, smem, s);
        case 8: return launch_staged<D, 8, POW2>(P, smem, s);
        default: return cudaErrorInvalidValue;
    }
}

size_t staged_smem_bytes(const EncodeParams& P, int F) {
    return static_cast<size_t>(P.n_stages) * P.stage_points * (16 + P.group_size * F * 4);
}

cudaError_t launch_backward_staged(const EncodeParams& P, int D, int F, bool pow2, cudaStream_t s) {
    if (P.n_stages < 2 || P.n_stages > kMaxStages) return cudaErrorInvalidValue;
    const size_t smem = staged_smem_bytes(P, F);
    switch (D * 2 + (pow2 ? 1 : 0)) {
        case 2: return staged_f<1, false>(P, F, smem, s);
        case 3: return staged_f<1, true>(P, F, smem, s);
        case 4: return staged_f<2, false>(P, F, smem, s);
        case 5: return staged_f<2, true>(P, F, smem, s);
        case 6: return staged_f<3, false>(P, F, smem, s);
        case 7: return staged_f<3, true>(P, F, smem, s);
        case 8: return staged_f<4, false>(P, F, smem, s);
        case 9: return staged_f<4, true>(P, F, smem, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
