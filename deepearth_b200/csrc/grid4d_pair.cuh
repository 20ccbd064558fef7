// grid4d_pair.cuh -- lane-pair building blocks shared by the multilinear kernels (pair / staged / coarse).
// Lane bit 0 (ox) is the x offset of the 2^(D-1) corners a lane owns; see grid4d_pair_kernels.cu.
#pragma once
#include "grid4d_common.cuh"

namespace de {

template <int D>
struct PairGeom {
    static constexpr int kCombos = 1 << (D - 1);  // corners owned by one lane
};

// Integer cell terms of one (point, level): term[d][o] = (g_d + o) * mul_d for d >= 1, cell0 = g_0 + ox.
template <int D, bool POW2>
struct CellTerms;

template <int D>
struct CellTerms<D, true> {
    uint32_t cell0;
    uint32_t term[D > 1 ? D - 1 : 1][2];
    __device__ __forceinline__ void set(const LevelParams& lv, const long long (&g)[D], int ox) {
        cell0 = static_cast<uint32_t>(g[0]) + static_cast<uint32_t>(ox);
#pragma unroll
        for (int d = 1; d < D; ++d) {
            const uint32_t c = static_cast<uint32_t>(g[d]);
            term[d - 1][0] = c * lv.mul[d];
            term[d - 1][1] = (c + 1u) * lv.mul[d];
        }
    }
    // combo j: bit (d-1) of j is the offset along dimension d
    __device__ __forceinline__ uint32_t index(const LevelParams& lv, int j) const {
        uint32_t h = cell0;
        if (lv.dense) {
#pragma unroll
            for (int d = 1; d < D; ++d) h += term[d - 1][(j >> (d - 1)) & 1];
        } else {
#pragma unroll
            for (int d = 1; d < D; ++d) h ^= term[d - 1][(j >> (d - 1)) & 1];
        }
        return h & (lv.T - 1u);   // T | 2^32: equals the int64 floor-mod of the reference
    }
};

template <int D>
struct CellTerms<D, false> {
    long long cell0;
    long long term[D > 1 ? D - 1 : 1][2];
    __device__ __forceinline__ void set(const LevelParams& lv, const long long (&g)[D], int ox) {
        cell0 = g[0] + ox;
#pragma unroll
        for (int d = 1; d < D; ++d) {
            const unsigned long long m = lv.mul[d];
            term[d - 1][0] = static_cast<long long>(static_cast<unsigned long long>(g[d]) * m);
            term[d - 1][1] = static_cast<long long>(static_cast<unsigned long long>(g[d] + 1) * m);
        }
    }
    __device__ __forceinline__ uint32_t index(const LevelParams& lv, int j) const {
        long long h = cell0;
        if (lv.dense) {
#pragma unroll
            for (int d = 1; d < D; ++d) h += term[d - 1][(j >> (d - 1)) & 1];
        } else {
#pragma unroll
            for (int d = 1; d < D; ++d) h ^= term[d - 1][(j >> (d - 1)) & 1];
        }
        long long r = h % static_cast<long long>(lv.T);   // hash_encoding.py:72, Python floor-mod
        if (r < 0) r += lv.T;
        return static_cast<uint32_t>(r);
    }
};

template <int D>
__device__ __forceinline__ void cell_of(const LevelParams& lv, const float (&x)[D], long long (&g)[D], float (&w)[D]) {
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float s = __fmul_rn(x[d], lv.res_f);   // hash_encoding.py:58
        const float fl = floorf(s);                  // :61
        g[d] = __float2ll_rd(s);
        w[d] = __fsub_rn(s, fl);
    }
}

// weight of the lane's corner (ox, combo j): ((u_0 * u_1) * u_2) ... -- the oracle's product order
template <int D>
__device__ __forceinline__ float pair_weight(const float (&w)[D], int ox, int j) {
    float wk = ox ? w[0] : __fsub_rn(1.0f, w[0]);
#pragma unroll
    for (int d = 1; d < D; ++d) {
        const float u = ((j >> (d - 1)) & 1) ? w[d] : __fsub_rn(1.0f, w[d]);
        wk = __fmul_rn(wk, u);
    }
    return wk;
}

template <int D>
__device__ __forceinline__ void load_coords(const EncodeParams& P, const LevelParams& lv, long long p, float (&x)[D]) {
    if (P.vec_coords) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(P.coords) + p);
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = pick4(v, lv.dim[d]);
    } else {
        const float* row = P.coords + p * P.coord_stride;
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = __ldg(row + lv.dim[d]);
    }
}

// same, as a stream: no L1 allocation, evict-first in L2
template <int D>
__device__ __forceinline__ void load_coords_stream(const EncodeParams& P, const LevelParams& lv, long long p, float (&x)[D], uint64_t pol) {
    if (P.vec_coords) {
        const float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(P.coords) + p, pol);
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = pick4(v, lv.dim[d]);
    } else {
        const float* row = P.coords + p * P.coord_stride;
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = ldg_stream_f1(row + lv.dim[d], pol);
    }
}

template <int F>
__device__ __forceinline__ void load_row_stream(const float* src, float (&v)[F], bool vec, uint64_t pol) {
    if constexpr (F == 2) {
        if (vec) { const float2 x = ldg_stream_f2(src, pol); v[0] = x.x; v[1] = x.y; return; }
    } else if constexpr (F == 4 || F == 8) {
        if (vec) {
#pragma unroll
            for (int h = 0; h < F / 4; ++h) {
                const float4 x = ldg_stream_f4(src + 4 * h, pol);
                v[4 * h] = x.x; v[4 * h + 1] = x.y; v[4 * h + 2] = x.z; v[4 * h + 3] = x.w;
            }
            return;
        }
    }
#pragma unroll
    for (int f = 0; f < F; ++f) v[f] = ldg_stream_f1(src + f, pol);
}

// The two lanes of a pair hold identical sums; they split the F features of the level so that
// consecutive lanes write consecutive floats of the feature row.
template <int F>
__device__ __forceinline__ void pair_store(float* dst, const float (&acc)[F], int ox, bool vec_rows) {
    if constexpr (F == 1) {
        if (ox == 0) __stcs(dst, acc[0]);
    } else if constexpr (F == 2) {
        __stcs(dst + ox, ox ? acc[1] : acc[0]);
    } else if constexpr (F == 4) {
        const float a = ox ? acc[2] : acc[0], b = ox ? acc[3] : acc[1];
        if (vec_rows) __stcs(reinterpret_cast<float2*>(dst) + ox, make_float2(a, b));
        else { __stcs(dst + 2 * ox, a); __stcs(dst + 2 * ox + 1, b); }
    } else {
        static_assert(F == 8, "pair kernels are instantiated for F in {1,2,4,8}");
        const float a = ox ? acc[4] : acc[0], b = ox ? acc[5] : acc[1];
        const float c = ox ? acc[6] : acc[2], d = ox ? acc[7] : acc[3];
        if (vec_rows) __stcs(reinterpret_cast<float4*>(dst) + ox, make_float4(a, b, c, d));
        else { float* q = dst + 4 * ox; __stcs(q, a); __stcs(q + 1, b); __stcs(q + 2, c); __stcs(q + 3, d); }
    }
}

}  // namespace de
