// mlp_common.cuh -- tcgen05 / TMEM / mbarrier building blocks shared by the MLP-head kernels (sm_100a).
// Bit layouts follow cute::UMMA::SmemDescriptor / InstrDescriptor (CUTLASS 4.x, cute/arch/mma_sm100_desc.hpp).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace de {
namespace mlp {

constexpr int kMlpThreads = 128;
constexpr int kTileM = 128;
constexpr int kH = 64;            // hidden width = output width
constexpr int kTmemCols = 128;    // D1: columns [0,64), D2: columns [64,128)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- UMMA canonical layout, K-major, SWIZZLE_NONE ----------------------------------------------
// core matrix = 8 rows x 16 bytes (8 bf16), stored as 128 contiguous bytes; core matrices adjacent along K are
// 128 B apart (leading byte offset), 8-row groups are (Kt/8)*128 B apart (stride byte offset).
__device__ __forceinline__ uint32_t canon_chunk_offset(int row, int k8, int kt) {
    return static_cast<uint32_t>((row >> 3) * (kt >> 3) * 128 + k8 * 128 + (row & 7) * 16);
}

// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);             // [0,14)  start address >> 4
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;   // [16,30) leading byte offset >> 4
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;   // [32,46) stride byte offset >> 4
    d |= 1ull << 46;                                                // [46,48) descriptor version 1 (Blackwell)
    return d;                                                       // base offset 0, layout type 0 = no swizzle
}

// instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate, bf16 A and B, both K-major, M=128
__device__ __forceinline__ constexpr uint32_t umma_idesc(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(kTileM >> 4) << 24);
}

// general form: M x N tile, either operand K-major (0) or MN-major (1: bits 15 / 16)
__host__ __device__ __forceinline__ constexpr uint32_t idesc_bf16(int m, int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn ? (1u << 15) : 0u) | (b_mn ? (1u << 16) : 0u) |
           (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}


__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
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

// 32 consecutive fp32 accumulator columns of this thread's TMEM lane (= tile row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// 16 columns, split into issue and wait so that the next piece's load is in flight while this one is consumed.  The wait
// names the destination registers as read-write operands: nothing may read them before it.
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}

// 32 contiguous bytes of a feature / gradient row (streamed: evict-first).  One 256-bit load (sm_100: LDG.256) when the
// address is 32 B aligned: one L1 request per 32 B sector instead of two half-used ones -- the head kernels issue
// 48 such loads per thread per tile, and the L1 request rate of those loads was ~1/3 of a tile's time (ncu r01g).
__device__ __forceinline__ void ld_stream_32B(const float* p, float4& a, float4& b, bool aligned32) {
    if (aligned32) {
        asm volatile("ld.global.cs.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
    } else {
        a = __ldcs(reinterpret_cast<const float4*>(p));
        b = __ldcs(reinterpret_cast<const float4*>(p) + 1);
    }
}

__device__ __forceinline__ void st_stream_32B(float* p, const float4& a, const float4& b, bool aligned32) {
    if (aligned32) {
        asm volatile("st.global.cs.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                     ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
    } else {
        __stcs(reinterpret_cast<float4*>(p), a);
        __stcs(reinterpret_cast<float4*>(p) + 1, b);
    }
}

// 32 contiguous, 32-byte aligned bytes as ONE request per lane (a whole sector): the projection kernels move their bf16
// rows 16 columns at a time per thread, and two 16-byte accesses per piece doubled the L1 wavefronts of that traffic
// (B200 r02s: hidden kernel 1.63 -> 1.35 ms, dpre kernel 3.63 -> 2.72 ms per 2^20 x 768).  Loads bypass L1 allocation
// (streamed once; the per-column constants are what should stay there).
__device__ __forceinline__ void ld_nc_32B(const void* p, uint32_t (&r)[8]) {
    asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
}
__device__ __forceinline__ void st_32B(void* p, const uint32_t (&r)[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// Row reductions over the 64 hidden units with four independent accumulator chains: a CTA keeps only 8..16 warps
// resident, so a 64-long dependent FADD chain per statistic would sit exposed (ncu r01f: one issue per 12 cycles per warp).
__device__ __forceinline__ float sum64(const float (&v)[kH]) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int j = 0; j < kH; j += 4) { a0 += v[j]; a1 += v[j + 1]; a2 += v[j + 2]; a3 += v[j + 3]; }
    return (a0 + a1) + (a2 + a3);
}
__device__ __forceinline__ float sumsq_centered64(const float (&v)[kH], float mean) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int j = 0; j < kH; j += 4) {
        const float d0 = v[j] - mean, d1 = v[j + 1] - mean, d2 = v[j + 2] - mean, d3 = v[j + 3] - mean;
        a0 = fmaf(d0, d0, a0); a1 = fmaf(d1, d1, a1); a2 = fmaf(d2, d2, a2); a3 = fmaf(d3, d3, a3);
    }
    return (a0 + a1) + (a2 + a3);
}
__device__ __forceinline__ float dot64(const float (&a)[kH], const float (&b)[kH]) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int j = 0; j < kH; j += 4) {
        a0 = fmaf(a[j], b[j], a0); a1 = fmaf(a[j + 1], b[j + 1], a1); a2 = fmaf(a[j + 2], b[j + 2], a2); a3 = fmaf(a[j + 3], b[j + 3], a3);
    }
    return (a0 + a1) + (a2 + a3);
}

// GELU (erf form, nn.GELU()): Phi(z) by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, far below the bf16 rounding of
// the hidden row), 1 - Phi(|z|) = 0.5 t (a1 + t (a2 + ...)) exp(-z^2/2), t = 1 / (1 + p |z| / sqrt 2): one MUFU.RCP and
// one MUFU.EX2 -- and the exponential is the one the derivative Phi(z) + z phi(z) needs.  ~16 instructions against
// ~35 for erff + expf; the tail z -> -inf keeps its relative accuracy (no 1 + erf cancellation).
__device__ __forceinline__ void gelu_parts(float z, float& cdf, float& e) {
    const float x = fabsf(z) * 0.70710678118654752f;
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, x, 1.0f)));     // MUFU.RCP (__frcp_rn is a Newton sequence)
    e = __expf(-x * x);
    float p = fmaf(t, 1.061405429f, -1.453152027f);
    p = fmaf(t, p, 1.421413741f);
    p = fmaf(t, p, -0.284496736f);
    p = fmaf(t, p, 0.254829592f);
    const float q = 0.5f * (p * t) * e;                 // 1 - Phi(|z|)
    cdf = z >= 0.0f ? 1.0f - q : q;
}
__device__ __forceinline__ float gelu_fwd(float z) {
    float cdf, e;
    gelu_parts(z, cdf, e);
    return z * cdf;
}
// a = gelu(z), g = d gelu / dz
__device__ __forceinline__ void gelu_fwd_bwd(float z, float& a, float& g) {
    float cdf, e;
    gelu_parts(z, cdf, e);
    a = z * cdf;
    g = fmaf(z * 0.3989422804014327f, e, cdf);
}


// Column sums over the 32 lanes of a warp for 32 register values each: afterwards lane l holds sum_over_lanes v[l].
// Butterfly with halving payload: 16 + 8 + 4 + 2 + 1 = 31 shuffles (a per-column butterfly would take 160).
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
    for (int w = 16; w >= 1; w >>= 1) {
        const bool up = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; ++i) {
            const float send = up ? v[i] : v[i + w];
            const float keep = up ? v[i + w] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
        }
    }
    return v[0];
}

// The same for 16 values: lane l returns the sum over all 32 lanes of v[l & 15] (15 + 1 shuffles).
__device__ __forceinline__ float warp_transpose_sum16(float (&v)[16], int lane) {
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1) {
        const bool up = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; ++i) {
            const float send = up ? v[i] : v[i + w];
            const float keep = up ? v[i + w] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

// ---- packed fp32 x 2 arithmetic (sm_100: FFMA2 / FADD2 / FMUL2, one issue slot for two lanes' worth of flops) ----------
// The LayerNorm / GELU epilogues of the wide projection are instruction-issue bound; three quarters of their instructions
// are fp32 pipe operations.  Each half of a pair is an ordinary IEEE operation: results equal the scalar code's bit for bit.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 v;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
    return v;
}
__device__ __forceinline__ f32x2 pk2u(uint32_t lo, uint32_t hi) {
    f32x2 v;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(lo), "r"(hi));
    return v;
}
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 splat2(float x) { return pk2(x, x); }

// GELU on a pair.  q = 1 - Phi(|z|) with the polynomial of gelu_parts (its 0.5 folded into the coefficients: exact), the
// exponential as one ex2.approx.ftz of -z^2/2 * log2 e (the non-ftz __expf wraps MUFU.EX2 in a range test and two FMULs
// per element; flushing exp(-z^2/2) < 2^-126 to zero changes nothing at the bf16 precision of the outputs).
__device__ __forceinline__ void gelu_tail2(f32x2 z, f32x2& az, f32x2& q, f32x2& e) {
    float zl, zh;
    upk2(z, zl, zh);
    az = pk2(fabsf(zl), fabsf(zh));
    const f32x2 x = mul2(az, splat2(0.70710678118654752f));
    float dl, dh;
    upk2(fma2(splat2(0.3275911f), x, splat2(1.0f)), dl, dh);
    float tl, th;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(tl) : "f"(dl));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(th) : "f"(dh));
    const f32x2 t = pk2(tl, th);
    float al, ah, el, eh;
    upk2(mul2(mul2(x, x), splat2(-1.4426950408889634f)), al, ah);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(el) : "f"(al));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eh) : "f"(ah));
    e = pk2(el, eh);
    f32x2 p = fma2(t, splat2(0.5f * 1.061405429f), splat2(0.5f * -1.453152027f));
    p = fma2(t, p, splat2(0.5f * 1.421413741f));
    p = fma2(t, p, splat2(0.5f * -0.284496736f));
    p = fma2(t, p, splat2(0.5f * 0.254829592f));
    q = mul2(mul2(p, t), e);
}
// gelu(z) = z Phi(z) = max(z, 0) - |z| q   (z >= 0: z - z q;  z < 0: z q)
__device__ __forceinline__ f32x2 gelu_fwd2(f32x2 z) {
    f32x2 az, q, e;
    gelu_tail2(z, az, q, e);
    float zl, zh;
    upk2(z, zl, zh);
    return fma2(az, mul2(q, splat2(-1.0f)), pk2(fmaxf(zl, 0.0f), fmaxf(zh, 0.0f)));
}
// a = gelu(z), g = d gelu / dz = Phi(z) + z phi(z), Phi(z) = 0.5 + copysign(0.5 - q, z)
__device__ __forceinline__ void gelu_fwd_bwd2(f32x2 z, f32x2& a, f32x2& g) {
    f32x2 az, q, e;
    gelu_tail2(z, az, q, e);
    float zl, zh, hl, hh;
    upk2(z, zl, zh);
    upk2(fma2(q, splat2(-1.0f), splat2(0.5f)), hl, hh);
    const f32x2 cdf = add2(splat2(0.5f), pk2(copysignf(hl, zl), copysignf(hh, zh)));
    a = mul2(z, cdf);
    g = fma2(mul2(z, splat2(0.3989422804014327f)), e, cdf);
}

// 2-byte element (row r, column c) of a K-major canonical tile with kt columns
__device__ __forceinline__ uint32_t canon_elem_offset(int r, int c, int kt) {
    return static_cast<uint32_t>((r >> 3) * (kt >> 3) * 128 + (c >> 3) * 128 + (r & 7) * 16 + (c & 7) * 2);
}

}  // namespace mlp
}  // namespace de
