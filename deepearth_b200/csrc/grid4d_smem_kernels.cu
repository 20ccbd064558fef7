// grid4d_smem_kernels.cu -- multilinear forward for dense COARSE levels whose live cells fit shared memory.
//
// A dense level only ever touches its (R+1)^D cells; for the coarsest levels that is a few KB to ~144 KB
// (R=16: 4913 cells, R=32: 35937 cells; fp16 F=2 -> 19.6 KB and 143.7 KB).  Every point reads 2^D
// entries of each of them, so on the L2 path these levels cost exactly as many L1->L2 packets as a fine
// level although their whole table would fit on the SM.  Here one persistent CTA per SM copies the live
// part of up to 4 such tables ("shared-memory group", <= ~200 KB) into shared memory once and then streams
// its share of the points: 16 B of coordinates in, 2^D shared-memory gathers per level, F floats out.
// Nothing but the coordinate / feature streams touches L2.
//
// Thread mapping is the lane-pair layout (grid4d_pair.cuh); arithmetic contract unchanged.  Coordinates
// outside [0,1] can index past the live cells (the index wraps mod T exactly like the reference's hash):
// those rare corners are read from the global table instead.
#include "grid4d_common.cuh"
#include "grid4d_pair.cuh"
#include "launch.h"

namespace de {

namespace {

template <typename TT, int F>
__device__ __forceinline__ void smem_entry(const unsigned char* base, uint32_t idx, float (&v)[F]) {
    if constexpr (sizeof(TT) == 2) {
        const __half2* q = reinterpret_cast<const __half2*>(base) + static_cast<size_t>(idx) * (F / 2);
        if constexpr (F == 2) {
            const float2 a = __half22float2(q[0]);
            v[0] = a.x; v[1] = a.y;
        } else if constexpr (F == 4) {
            const uint2 raw = *reinterpret_cast<const uint2*>(q);
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
            const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
        } else {
            static_assert(F == 8, "fp16 shared-memory tables: F in {2,4,8}");
            const uint4 raw = *reinterpret_cast<const uint4*>(q);
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
            const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
            const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
            const float2 d = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
        }
    } else {
        const float* q = reinterpret_cast<const float*>(base) + static_cast<size_t>(idx) * F;
        if constexpr (F == 1) {
            v[0] = q[0];
        } else if constexpr (F == 2) {
            const float2 a = *reinterpret_cast<const float2*>(q);
            v[0] = a.x; v[1] = a.y;
        } else {
            static_assert(F == 4 || F == 8, "fp32 shared-memory tables: F in {1,2,4,8}");
#pragma unroll
            for (int h = 0; h < F / 4; ++h) {
                const float4 a = reinterpret_cast<const float4*>(q)[h];
                v[4 * h] = a.x; v[4 * h + 1] = a.y; v[4 * h + 2] = a.z; v[4 * h + 3] = a.w;
            }
        }
    }
}

}  // namespace

template <int D, typename TT, int F, bool POW2>
__global__ void __launch_bounds__(kSmemThreads, 1)
grid4d_forward_smem_kernel(const __grid_constant__ EncodeParams P, const __grid_constant__ SmemPlan S) {
    extern __shared__ __align__(16) unsigned char s_tab[];
    __shared__ LevelParams s_lv[kSmemGroupLevels];

    int sg = 0;
    while (sg + 1 < S.n_sgroups && static_cast<int>(blockIdx.x) >= S.cta_begin[sg + 1]) ++sg;
    const int cta = blockIdx.x - S.cta_begin[sg];
    const int n_cta = S.cta_count[sg];
    const int n_in = S.count[sg];
    const int Gs = S.gs[sg];
    if (threadIdx.x < n_in) s_lv[threadIdx.x] = P.levels[S.first[sg] + threadIdx.x];
    __syncthreads();

    // ---- stage the live cells of the group's tables --------------------------------------------
    constexpr int EB = static_cast<int>(sizeof(TT)) * F;   // entry bytes (multiple of 4, host-checked)
    for (int l = 0; l < n_in; ++l) {
        const uint32_t words = S.live[sg][l] * (EB / 4);
        const uint32_t* src = static_cast<const uint32_t*>(s_lv[l].table);
        uint32_t* dst = reinterpret_cast<uint32_t*>(s_tab + S.off[sg][l]);
        for (uint32_t i = threadIdx.x; i < words; i += kSmemThreads) dst[i] = __ldg(src + i);
    }
    __syncthreads();

    const int ox = threadIdx.x & 1;
    const int lane_level = (threadIdx.x >> 1) & (Gs - 1);
    if (lane_level >= n_in) return;                        // both lanes of a pair leave together
    const LevelParams lv = s_lv[lane_level];
    const unsigned char* tab = s_tab + S.off[sg][lane_level];
    const uint32_t live = S.live[sg][lane_level];
    const unsigned pair_mask = 3u << (threadIdx.x & 30);
    const int p_step = kSmemThreads / (2 * Gs);
    const long long stride = static_cast<long long>(n_cta) * p_step;
    constexpr int J = PairGeom<D>::kCombos;
    constexpr int U = 4;                                   // points in flight per thread

    for (long long p0 = static_cast<long long>(cta) * p_step + threadIdx.x / (2 * Gs); p0 < P.n_points; p0 += U * stride) {
        float x[U][D];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long p = p0 + u * stride;
            if (p < P.n_points) load_coords<D>(P, lv, p, x[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long p = p0 + u * stride;
            if (p >= P.n_points) break;                    // uniform within a pair
            long long g[D];
            float w[D];
            cell_of<D>(lv, x[u], g, w);
            CellTerms<D, POW2> ct;
            ct.set(lv, g, ox);
            bool has_mask;
            const float m = row_mask(P, lv, p, has_mask);
            float acc[F];
#pragma unroll
            for (int f = 0; f < F; ++f) acc[f] = 0.0f;
#pragma unroll
            for (int j = 0; j < J; ++j) {
                const uint32_t idx = ct.index(lv, j);
                float v[F];
                if (idx < live) smem_entry<TT, F>(tab, idx, v);
                else EntryLoader<TT, F>::load(lv.table, idx, v);   // out-of-range coordinates only
                const float wk = pair_weight<D>(w, ox, j);
#pragma unroll
                for (int f = 0; f < F; ++f) acc[f] = __fmaf_rn(wk, v[f], acc[f]);
            }
#pragma unroll
            for (int f = 0; f < F; ++f) {
                acc[f] = __fadd_rn(acc[f], __shfl_xor_sync(pair_mask, acc[f], 1));
                if (P.out_scale != 1.0f) acc[f] = __fmul_rn(acc[f], P.out_scale);
                if (has_mask) acc[f] = __fmul_rn(acc[f], m);
            }
            pair_store<F>(P.out + p * P.out_stride + lv.out_col, acc, ox, P.vec_rows != 0);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
template <int D, typename TT, int F, bool POW2>
static cudaError_t launch_smem(const EncodeParams& P, const SmemPlan& S, cudaStream_t s) {
    auto kernel = grid4d_forward_smem_kernel<D, TT, F, POW2>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(S.smem_bytes));
    if (e != cudaSuccess) return e;
    kernel<<<static_cast<unsigned>(S.n_ctas), kSmemThreads, S.smem_bytes, s>>>(P, S);
    return cudaGetLastError();
}

template <int D, bool POW2>
static cudaError_t smem_tf(const EncodeParams& P, const SmemPlan& S, bool half_tables, int F, cudaStream_t s) {
    if (half_tables) {
        switch (F) {
            case 2: return launch_smem<D, __half, 2, POW2>(P, S, s);
            case 4: return launch_smem<D, __half, 4, POW2>(P, S, s);
            case 8: return launch_smem<D, __half, 8, POW2>(P, S, s);
            default: return cudaErrorInvalidValue;
        }
    }
    switch (F) {
        case 1: return launch_smem<D, float, 1, POW2>(P, S, s);
        case 2: return launch_smem<D, float, 2, POW2>(P, S, s);
        case 4: return launch_smem<D, float, 4, POW2>(P, S, s);
        case 8: return launch_smem<D, float, 8, POW2>(P, S, s);
        default: return cudaErrorInvalidValue;
    }
}

bool smem_layout_supported(bool half_tables, int F) {
    return half_tables ? (F == 2 || F == 4 || F == 8) : (F == 1 || F == 2 || F == 4 || F == 8);
}

cudaError_t launch_forward_smem(const EncodeParams& P, const SmemPlan& S, int D, bool half_tables, int F, bool pow2,
                                cudaStream_t s) {
    switch (D * 2 + (pow2 ? 1 : 0)) {
        case 2: return smem_tf<1, false>(P, S, half_tables, F, s);
        case 3: return smem_tf<1, true>(P, S, half_tables, F, s);
        case 4: return smem_tf<2, false>(P, S, half_tables, F, s);
        case 5: return smem_tf<2, true>(P, S, half_tables, F, s);
        case 6: return smem_tf<3, false>(P, S, half_tables, F, s);
        case 7: return smem_tf<3, true>(P, S, half_tables, F, s);
        case 8: return smem_tf<4, false>(P, S, half_tables, F, s);
        case 9: return smem_tf<4, true>(P, S, half_tables, F, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
