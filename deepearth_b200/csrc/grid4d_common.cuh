// grid4d_common.cuh -- device-side building blocks shared by the Grid4D kernels (sm_100a).
//
// Arithmetic contract (must match oracle/grid4d_oracle.c, which restates
// /root/reference/models/hash_encoding.py:47-96):
//   s_d = fl32(c_d * fl32(R));  g_d = (int64) floor(s_d);  w_d = s_d - floor(s_d)
//   hashed: h = XOR_d (g_d + o_d) * prime_d  (wrapping int64), dense: h = SUM_d (g_d+o_d)*(R+1)^d
//   idx = floor_mod(h, T)
// Every fp32 operation that decides an index or a weight is a single IEEE operation written with
// an explicit intrinsic (__fmul_rn / __fsub_rn / __fmaf_rn) so the compiler cannot contract it.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace de {

constexpr int kMaxLaunchLevels = 64;  // levels carried by one launch (kernel parameter space)
constexpr int kMaxGroup = 32;         // levels handled by one CTA
constexpr int kThreads = 256;

// hash_encoding.py:64 -- the int64 primes; their low 32 bits are what a power-of-two table sees.
__device__ __constant__ const uint32_t kPrime32[4] = {1u, 2654435761u, 805459861u, 3674653429u};
__device__ __constant__ const uint64_t kPrime64[4] = {1ull, 2654435761ull, 805459861ull, 3674653429ull};

struct LevelParams {
    const void* table;  // (T,F) entries
    float* grad;        // (T,F) fp32 accumulators (backward)
    unsigned long long out_col;   // float offset of the level's slice inside a row (or, for regrouped gradient
                                  // rows, plane offset + position inside the slice: can exceed 2^32)
    uint32_t T;
    float res_f;        // (float)R
    uint32_t res_p1;    // R+1: dense stride base
    uint32_t pad0;
    uint8_t dim[4];
    uint8_t dense;
    uint8_t pow2;       // T is a power of two: 32-bit hash fast path is bit-identical
    uint8_t mask_id;
    uint8_t pad;
    uint32_t mul[4];    // per-dimension index multiplier: prime_d (hashed) or (R+1)^d (dense); mul[0] == 1
    // backward, contended coarse levels: CTA tile t accumulates into replica (t % rep_count); replica 0 is
    // `grad` itself, replicas 1.. live in the caller's workspace and are folded into `grad` afterwards
    float* rep_base;
    uint32_t rep_count;   // 0 / 1 = not replicated
    uint32_t rep_stride;  // floats between replicas
    uint32_t rep_live;    // entries a replica holds (indices >= rep_live go to `grad`: out-of-range coordinates)
    uint32_t pad2;
};

// Gradient table a CTA tile accumulates into, and the bound below which it applies.
struct GradTarget {
    float* base;
    uint32_t live;
};
__device__ __forceinline__ GradTarget grad_target(const LevelParams& lv, int tile) {
    GradTarget t{lv.grad, 0u};
    if (lv.rep_count > 1) {
        const uint32_t r = static_cast<uint32_t>(tile) % lv.rep_count;
        if (r) {
            t.base = lv.rep_base + static_cast<size_t>(r - 1) * lv.rep_stride;
            t.live = lv.rep_live;
        }
    }
    return t;
}

struct EncodeParams {
    const float* coords;
    const uint8_t* mask_a;
    const uint8_t* mask_b;
    float* out;             // forward: features; hash_indices: unused
    const float* grad_out;  // backward
    long long* indices;     // hash_indices kernel
    long long n_points;
    long long out_stride;   // row stride of out / grad_out in floats
    int coord_stride;
    int n_levels;           // levels in this launch
    int group_size;         // levels per CTA (power of two)
    int n_groups;
    int n_tiles;            // point tiles per group
    int points_per_cta;
    int n_features;         // runtime F (used when the template F is 0)
    int vec_coords;         // coords are 16 B-aligned float4 rows
    int vec_rows;           // feature / grad rows are aligned for F-wide vector access
    int n_stages;           // staged kernels: ring depth
    int stage_points;       // staged kernels: points per stage (multiple of the points one CTA pass covers)
    int l2_hints;           // pair kernels: evict_last on tables, evict_first on streams
    float out_scale;
    LevelParams levels[kMaxLaunchLevels];
};

// ---- shared-memory coarse levels (grid4d_smem_kernels.cu) ---------------------------------------
constexpr int kSmemThreads = 512;
constexpr int kSmemGroupLevels = 4;   // levels one CTA keeps in shared memory
constexpr int kMaxSmemGroups = 8;     // shared-memory groups per launch

struct SmemPlan {
    int n_sgroups;
    int n_ctas;
    unsigned smem_bytes;              // dynamic shared memory per CTA (largest group)
    int first[kMaxSmemGroups];        // index of the group's first level in EncodeParams::levels
    int count[kMaxSmemGroups];        // levels in the group (<= kSmemGroupLevels)
    int gs[kMaxSmemGroups];           // power of two >= count: lanes per point = 2 * gs
    int cta_begin[kMaxSmemGroups];    // first blockIdx.x of the group's CTAs
    int cta_count[kMaxSmemGroups];
    uint32_t off[kMaxSmemGroups][kSmemGroupLevels];    // byte offset of each staged table (16 B aligned)
    uint32_t live[kMaxSmemGroups][kSmemGroupLevels];   // staged entries: (R+1)^D
};

__device__ __forceinline__ float pick4(const float4& v, int d) {
    return d == 0 ? v.x : (d == 1 ? v.y : (d == 2 ? v.z : v.w));
}

// Per-(point, level) cell: integer cell coordinates (low 32 bits + full 64) and fractional weights.
template <int D>
struct Cell {
    long long g[D];
    float w[D];
};

template <int D>
__device__ __forceinline__ void load_cell(const EncodeParams& P, const LevelParams& lv, long long p,
                                          Cell<D>& c) {
    float x[D];
    if (P.vec_coords) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(P.coords) + p);
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = pick4(v, lv.dim[d]);
    } else {
        const float* row = P.coords + p * P.coord_stride;
#pragma unroll
        for (int d = 0; d < D; ++d) x[d] = __ldg(row + lv.dim[d]);
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float s = __fmul_rn(x[d], lv.res_f);  // hash_encoding.py:58
        const float fl = floorf(s);                 // :61
        c.g[d] = __float2ll_rd(s);
        c.w[d] = __fsub_rn(s, fl);
    }
}

// Index of corner `k` (bit d = +1 offset along grid dim d).
template <int D>
__device__ __forceinline__ uint32_t corner_index(const LevelParams& lv, const Cell<D>& c, int k) {
    if (lv.pow2) {
        // T | 2^32, so arithmetic mod 2^32 followed by & (T-1) equals the int64 floor-mod.
        uint32_t h = 0;
        uint32_t stride = 1;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const uint32_t cell = static_cast<uint32_t>(c.g[d]) + ((k >> d) & 1);
            if (lv.dense) {
                h += cell * stride;
                stride *= lv.res_p1;
            } else {
                h ^= cell * kPrime32[d];
            }
        }
        return h & (lv.T - 1u);
    }
    long long h = 0;
    long long stride = 1;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const long long cell = c.g[d] + ((k >> d) & 1);
        if (lv.dense) {
            h += cell * stride;
            stride *= static_cast<long long>(lv.res_p1);
        } else {
            h ^= static_cast<long long>(static_cast<unsigned long long>(cell) * kPrime64[d]);
        }
    }
    long long r = h % static_cast<long long>(lv.T);  // hash_encoding.py:72, Python floor-mod
    if (r < 0) r += lv.T;
    return static_cast<uint32_t>(r);
}

template <int D>
__device__ __forceinline__ float corner_weight(const Cell<D>& c, int k) {
    float wk = 1.0f;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float u = ((k >> d) & 1) ? c.w[d] : __fsub_rn(1.0f, c.w[d]);
        wk = (d == 0) ? u : __fmul_rn(wk, u);
    }
    return wk;
}

__device__ __forceinline__ float row_mask(const EncodeParams& P, const LevelParams& lv, long long p,
                                          bool& has_mask) {
    has_mask = false;
    if (lv.mask_id == 1 && P.mask_a) {
        has_mask = true;
        return __ldg(P.mask_a + p) ? 1.0f : 0.0f;
    }
    if (lv.mask_id == 2 && P.mask_b) {
        has_mask = true;
        return __ldg(P.mask_b + p) ? 1.0f : 0.0f;
    }
    return 1.0f;
}

// ---- L2 residency hints -----------------------------------------------------------------------
// The working set that must stay in L2 is the level group's tables (forward) or gradient tables (backward);
// coordinates, gradient rows and feature rows stream through once per group.  With plain loads the
// streams evict the tables: profiles/r01c_kernels_c3.txt shows 81 GB read + 41 GB written to HBM by a
// backward whose algorithmic HBM traffic is ~25 GB.  Table accesses therefore carry an evict_last
// policy and the streams an evict_first one (createpolicy + .L2::cache_hint).
struct L2Policies {
    uint64_t keep;     // tables / gradient tables
    uint64_t stream;   // coordinates, gradient rows
};
__device__ __forceinline__ L2Policies make_l2_policies(bool hints) {
    L2Policies p;
    if (hints) {
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p.keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p.stream));
    } else {
        asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p.keep));
        p.stream = p.keep;
    }
    return p;
}
__device__ __forceinline__ uint32_t ldg_hint_u32(const void* a, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.nc.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ uint2 ldg_hint_u64(const void* a, uint64_t pol) {
    uint2 v;
    asm volatile("ld.global.nc.L2::cache_hint.v2.b32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ uint4 ldg_hint_u128(const void* a, uint64_t pol) {
    uint4 v;
    asm volatile("ld.global.nc.L2::cache_hint.v4.b32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a), "l"(pol));
    return v;
}
// streaming loads: no L1 allocation, evict-first in L2
__device__ __forceinline__ float4 ldg_stream_f4(const void* a, uint64_t pol) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ float2 ldg_stream_f2(const void* a, uint64_t pol) {
    float2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f32 {%0, %1}, [%2], %3;" : "=f"(v.x), "=f"(v.y) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ float ldg_stream_f1(const void* a, uint64_t pol) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(a), "l"(pol));
    return v;
}

// Experiment (DE_FLAG_COHERENT_GATHER): the same 4 / 8 byte gathers as GPU-scope relaxed loads, which are served by the
// line's home L2 partition instead of being cached a second time in the requesting SM's partition.
__device__ __forceinline__ uint32_t ldg_coherent_u32(const void* a, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ uint2 ldg_coherent_u64(const void* a, uint64_t pol) {
    uint2 v;
    asm volatile("ld.relaxed.gpu.global.L2::cache_hint.v2.b32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(a), "l"(pol));
    return v;
}

// F-wide table entry with an L2 policy (F * sizeof(TT) in {2,4,8,16,32} bytes)
template <typename TT, int F>
__device__ __forceinline__ void load_entry_hint(const void* t, uint32_t i, float (&v)[F], uint64_t pol, bool coherent = false) {
    constexpr int B = static_cast<int>(sizeof(TT)) * F;
    const char* a = static_cast<const char*>(t) + static_cast<size_t>(i) * B;
    uint32_t w[B >= 4 ? B / 4 : 1];
    if constexpr (B == 2) {
        unsigned short h;
        asm volatile("ld.global.nc.L2::cache_hint.b16 %0, [%1], %2;" : "=h"(h) : "l"(a), "l"(pol));
        w[0] = h;
    } else if constexpr (B == 4) {
        w[0] = coherent ? ldg_coherent_u32(a, pol) : ldg_hint_u32(a, pol);
    } else if constexpr (B == 8) {
        const uint2 x = coherent ? ldg_coherent_u64(a, pol) : ldg_hint_u64(a, pol);
        w[0] = x.x; w[1] = x.y;
    } else if constexpr (B == 16) {
        const uint4 x = ldg_hint_u128(a, pol);
        w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w;
    } else {
        static_assert(B == 32, "entry size");
        const uint4 x = ldg_hint_u128(a, pol), y = ldg_hint_u128(a + 16, pol);
        w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w; w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
    }
    if constexpr (sizeof(TT) == 4) {
#pragma unroll
        for (int f = 0; f < F; ++f) v[f] = __uint_as_float(w[f]);
    } else if constexpr (F == 1) {
        v[0] = __half2float(__ushort_as_half(static_cast<unsigned short>(w[0])));
    } else {
#pragma unroll
        for (int f = 0; f < F / 2; ++f) {
            const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&w[f]));
            v[2 * f] = x.x; v[2 * f + 1] = x.y;
        }
    }
}

// ---- table entry loads: one vector load per entry where the entry is 2/4/8/16/32 bytes ----------
template <typename TT, int F>
struct EntryLoader;

template <>
struct EntryLoader<float, 1> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[1]) {
        v[0] = __ldg(static_cast<const float*>(t) + i);
    }
};
template <>
struct EntryLoader<float, 2> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[2]) {
        const float2 x = __ldg(static_cast<const float2*>(t) + i);
        v[0] = x.x; v[1] = x.y;
    }
};
template <>
struct EntryLoader<float, 4> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[4]) {
        const float4 x = __ldg(static_cast<const float4*>(t) + i);
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
    }
};
template <>
struct EntryLoader<float, 8> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[8]) {
        const float4 a = __ldg(static_cast<const float4*>(t) + 2ull * i);
        const float4 b = __ldg(static_cast<const float4*>(t) + 2ull * i + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
};
template <>
struct EntryLoader<__half, 1> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[1]) {
        const unsigned short raw = __ldg(static_cast<const unsigned short*>(t) + i);
        v[0] = __half2float(__ushort_as_half(raw));
    }
};
template <>
struct EntryLoader<__half, 2> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[2]) {
        const unsigned int raw = __ldg(static_cast<const unsigned int*>(t) + i);
        const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&raw));
        v[0] = x.x; v[1] = x.y;
    }
};
template <>
struct EntryLoader<__half, 4> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[4]) {
        const uint2 raw = __ldg(static_cast<const uint2*>(t) + i);
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
};
template <>
struct EntryLoader<__half, 8> {
    static __device__ __forceinline__ void load(const void* t, uint32_t i, float (&v)[8]) {
        const uint4 raw = __ldg(static_cast<const uint4*>(t) + i);
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
        const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
        const float2 d = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
    }
};

template <typename TT>
__device__ __forceinline__ float load_elem(const void* t, unsigned long long e) {
    return __ldg(static_cast<const float*>(t) + e);
}
template <>
__device__ __forceinline__ float load_elem<__half>(const void* t, unsigned long long e) {
    return __half2float(__ushort_as_half(__ldg(static_cast<const unsigned short*>(t) + e)));
}

// ---- streaming row stores / loads of F consecutive floats ---------------------------------------
template <int F>
__device__ __forceinline__ void store_row(float* dst, const float (&v)[F], bool vec) {
    if constexpr (F == 2) {
        if (vec) { __stcs(reinterpret_cast<float2*>(dst), make_float2(v[0], v[1])); return; }
    } else if constexpr (F == 4) {
        if (vec) { __stcs(reinterpret_cast<float4*>(dst), make_float4(v[0], v[1], v[2], v[3])); return; }
    } else if constexpr (F == 8) {
        if (vec) {
            __stcs(reinterpret_cast<float4*>(dst), make_float4(v[0], v[1], v[2], v[3]));
            __stcs(reinterpret_cast<float4*>(dst) + 1, make_float4(v[4], v[5], v[6], v[7]));
            return;
        }
    }
#pragma unroll
    for (int f = 0; f < F; ++f) __stcs(dst + f, v[f]);
}

template <int F>
__device__ __forceinline__ void load_row(const float* src, float (&v)[F], bool vec) {
    if constexpr (F == 2) {
        if (vec) {
            const float2 x = __ldcs(reinterpret_cast<const float2*>(src));
            v[0] = x.x; v[1] = x.y;
            return;
        }
    } else if constexpr (F == 4) {
        if (vec) {
            const float4 x = __ldcs(reinterpret_cast<const float4*>(src));
            v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
            return;
        }
    } else if constexpr (F == 8) {
        if (vec) {
            const float4 a = __ldcs(reinterpret_cast<const float4*>(src));
            const float4 b = __ldcs(reinterpret_cast<const float4*>(src) + 1);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            return;
        }
    }
#pragma unroll
    for (int f = 0; f < F; ++f) v[f] = __ldcs(src + f);
}

// ---- vector fp32 reductions into the gradient table (REDG.E.ADD.F32x2 / F32x4 on sm_100a) -------
// Written as red.global PTX on purpose: the gradient pointer travels through shared memory (LevelParams),
// so the compiler cannot prove it is a global address and atomicAdd() becomes a GENERIC `ATOM ... RZ`
// -- an atomic WITH return: a 32 B response per sector back to the SM (45 GB per C2 backward in
// profiles/r01b_kernels_c2.txt) and a scoreboard wait.  REDG is fire-and-forget.
__device__ __forceinline__ void red_add_f32(float* p, float a) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}
__device__ __forceinline__ void red_add_v2(float* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void red_add_v2_hint(float* p, float a, float b, uint64_t pol) {
    asm volatile("red.global.add.L2::cache_hint.v2.f32 [%0], {%1, %2}, %3;" ::"l"(p), "f"(a), "f"(b), "l"(pol) : "memory");
}
__device__ __forceinline__ void red_add_v4_hint(float* p, float a, float b, float c, float d, uint64_t pol) {
    asm volatile("red.global.add.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "l"(pol) : "memory");
}
__device__ __forceinline__ void red_add_f32_hint(float* p, float a, uint64_t pol) {
    asm volatile("red.global.add.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(p), "f"(a), "l"(pol) : "memory");
}
template <int F>
__device__ __forceinline__ void red_add_entry_hint(float* grad, uint32_t idx, const float (&v)[F], uint64_t pol) {
    float* dst = grad + static_cast<unsigned long long>(idx) * F;
    if constexpr (F == 2) {
        red_add_v2_hint(dst, v[0], v[1], pol);
    } else if constexpr (F == 4) {
        red_add_v4_hint(dst, v[0], v[1], v[2], v[3], pol);
    } else if constexpr (F == 8) {
        red_add_v4_hint(dst, v[0], v[1], v[2], v[3], pol);
        red_add_v4_hint(dst + 4, v[4], v[5], v[6], v[7], pol);
    } else {
#pragma unroll
        for (int f = 0; f < F; ++f) red_add_f32_hint(dst + f, v[f], pol);
    }
}

template <int F>
__device__ __forceinline__ void red_add_entry(float* grad, uint32_t idx, const float (&v)[F]) {
    float* dst = grad + static_cast<unsigned long long>(idx) * F;
    if constexpr (F == 2) {
        red_add_v2(dst, v[0], v[1]);
    } else if constexpr (F == 4) {
        red_add_v4(dst, v[0], v[1], v[2], v[3]);
    } else if constexpr (F == 8) {
        red_add_v4(dst, v[0], v[1], v[2], v[3]);
        red_add_v4(dst + 4, v[4], v[5], v[6], v[7]);
    } else {
#pragma unroll
        for (int f = 0; f < F; ++f) red_add_f32(dst + f, v[f]);
    }
}

}  // namespace de
