// mlp_head_bwd.cu -- backward of the 64-wide projection head on tcgen05 / TMEM (sm_100a), one kernel.
//
//   forward:  h = x W1^T + b1 ;  z = LN(h) (optional) ;  a = act(z) ;  y = a W2^T + b2
//   given gy = dL/dy, this kernel produces dL/dx (the gradient that feeds the hash-grid backward) and
//   accumulates dW1, db1, dgamma, dbeta, dW2, db2.
//
// Per 128-point tile (one persistent 128-thread CTA per SM, all 512 TMEM columns):
//   GEMM 1  D_h  [128 x 64]  = X   [128 x K1] * W1^T          (recompute of the hidden pre-activation)
//   GEMM 2  D_da [128 x 64]  = GY  [128 x 64] * W2            (gradient w.r.t. the activation)
//   epilogue A: every thread owns one point: LayerNorm statistics, activation and its derivative, LayerNorm
//               backward, all in registers; dgamma / dbeta column sums by a 31-shuffle transposing warp reduction
//   GEMM 3  D_dx [128 x K1]  = DH  [128 x 64] * W1            (gradient w.r.t. the features)
//   GEMM 4  D_dW [128 x N4] += [GY^T ; DH^T] [128 x 128 pts] * [A^T ; X^T ; 1] ^T
//           rows 0..63  x cols 0..63        = GY^T A  = dW2        rows 0..63  x col 64+K1 = sum GY = db2
//           rows 64..127 x cols 64..64+K1   = DH^T X  = dW1        rows 64..127 x col 64+K1 = sum DH = db1
//           (the two off-diagonal blocks are computed and ignored: tensor flops are free on this HBM-bound path).
//           D_dW stays in TMEM for the CTA's whole life (accumulate flag) and is reduced into global memory once.
// All operands are bf16 in the UMMA canonical K-major no-swizzle layout; the point-major tiles (GY, DH, A, X)
// are also written transposed (2-byte shared stores) so that GEMM 4 contracts over points with K-major operands.
#include <cstdlib>

#include "launch.h"
#include "mlp_common.cuh"

namespace de {

using namespace mlp;

namespace {

constexpr int kBwdTmemCols = 512;
constexpr int kColDa = 64, kColDx = 128, kColDw = 256;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// lane l returns sum over the warp's 32 lanes of vals[l] (31 shuffles; vals is destroyed)
__device__ __forceinline__ float warp_colsum32(float (&vals)[32], int lane) {
#pragma unroll
    for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < n / 2; ++i) {
            const float send = upper ? vals[i] : vals[i + n / 2];
            const float keep = upper ? vals[i + n / 2] : vals[i];
            vals[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return vals[0];
}

__device__ __forceinline__ void st_bf16(unsigned char* base, uint32_t off, float v) {
    *reinterpret_cast<__nv_bfloat16*>(base + off) = __float2bfloat16_rn(v);
}

__device__ __forceinline__ void red_f32(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// fp32 rows of a (n, KT) matrix -> bf16 tile, K-major at `dst` and transposed (feature-major, point columns) at
// `dst_t` rows [t_row0, t_row0 + KT).  Rows >= n are zero.  Warp w fills tile rows 32w .. 32w+31.
template <int KT>
struct TileRegs {
    static constexpr int kIt = (4 * KT) / 32;
    float4 a[kIt], b[kIt];
};

// Issue every global load of a 128 x KT fp32 tile (64 B x KT/8 per thread in flight); rows >= n read as zero.
template <int KT>
__device__ __forceinline__ void tile_issue(TileRegs<KT>& t, const float* src, long long stride, long long p0, long long n,
                                           int warp, int lane) {
#pragma unroll
    for (int it = 0; it < TileRegs<KT>::kIt; ++it) {
        const int c = lane + 32 * it;
        const int r = warp * 32 + (c / KT) * 8 + ((c % KT) & 7);
        const int k8 = (c % KT) >> 3;
        t.a[it] = t.b[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p0 + r < n) {
            const float4* s4 = reinterpret_cast<const float4*>(src + (p0 + r) * stride + k8 * 8);
            t.a[it] = __ldcs(s4);
            t.b[it] = __ldcs(s4 + 1);
        }
    }
}

// bf16 tile, K-major at `dst` and transposed (feature-major, point columns) at `dst_t` rows [t_row0, t_row0 + KT)
template <int KT>
__device__ __forceinline__ void tile_store(const TileRegs<KT>& t, unsigned char* dst, unsigned char* dst_t, int t_row0,
                                           int warp, int lane) {
#pragma unroll
    for (int it = 0; it < TileRegs<KT>::kIt; ++it) {
        const int c = lane + 32 * it;
        const int r = warp * 32 + (c / KT) * 8 + ((c % KT) & 7);
        const int k8 = (c % KT) >> 3;
        const float4 a = t.a[it], b = t.b[it];
        *reinterpret_cast<uint4*>(dst + canon_chunk_offset(r, k8, KT)) =
            make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
        const float e[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int q = 0; q < 8; ++q) st_bf16(dst_t, canon_elem_offset(t_row0 + k8 * 8 + q, r, kTileM), e[q]);
    }
}

}  // namespace

// dynamic shared memory (bf16, canonical K-major):
//   X [128 x K1] | W1 [64 x K1] | GY [128 x 64] | W2T [64 x 64] | DH [128 x 64] | W1T [K1 x 64] |
//   AT [128 x 128 pts] = [GY^T ; DH^T] | BT [N4 x 128 pts] = [A^T ; X^T ; ones(16)]
template <int K1>
__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_head_bwd_kernel(const __grid_constant__ MlpHeadBwdParams P) {
    constexpr int N4 = kH + K1 + 16;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float s_vec[3][kH];            // b1, ln gamma, ln beta
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_tmem;

    unsigned char* sX = smem;
    unsigned char* sW1 = sX + kTileM * K1 * 2;
    unsigned char* sGY = sW1 + kH * K1 * 2;
    unsigned char* sW2T = sGY + kTileM * kH * 2;
    unsigned char* sDH = sW2T + kH * kH * 2;
    unsigned char* sW1T = sDH + kTileM * kH * 2;
    unsigned char* sAT = sW1T + K1 * kH * 2;
    unsigned char* sBT = sAT + kTileM * kTileM * 2;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // ---- one-time setup ------------------------------------------------------------------------
    const __nv_bfloat16* w1 = static_cast<const __nv_bfloat16*>(P.w1);
    const __nv_bfloat16* w1t = static_cast<const __nv_bfloat16*>(P.w1t);
    const __nv_bfloat16* w2t = static_cast<const __nv_bfloat16*>(P.w2t);
    for (int c = tid; c < kH * (K1 / 8); c += kMlpThreads) {        // W1 (64, K1): B of GEMM 1
        const int n = c / (K1 / 8), k8 = c % (K1 / 8);
        *reinterpret_cast<uint4*>(sW1 + canon_chunk_offset(n, k8, K1)) = *reinterpret_cast<const uint4*>(w1 + static_cast<size_t>(n) * K1 + k8 * 8);
    }
    for (int c = tid; c < K1 * (kH / 8); c += kMlpThreads) {        // W1^T (K1, 64): B of GEMM 3
        const int n = c / (kH / 8), k8 = c % (kH / 8);
        *reinterpret_cast<uint4*>(sW1T + canon_chunk_offset(n, k8, kH)) = *reinterpret_cast<const uint4*>(w1t + static_cast<size_t>(n) * kH + k8 * 8);
    }
    for (int c = tid; c < kH * (kH / 8); c += kMlpThreads) {        // W2^T (64 hidden, 64 out): B of GEMM 2
        const int n = c / (kH / 8), k8 = c % (kH / 8);
        *reinterpret_cast<uint4*>(sW2T + canon_chunk_offset(n, k8, kH)) = *reinterpret_cast<const uint4*>(w2t + static_cast<size_t>(n) * kH + k8 * 8);
    }
    for (int c = tid; c < 16 * (kTileM / 8); c += kMlpThreads) {    // 16 rows of ones: bias-gradient columns of GEMM 4
        const int n = kH + K1 + c / (kTileM / 8), k8 = c % (kTileM / 8);
        *reinterpret_cast<uint4*>(sBT + canon_chunk_offset(n, k8, kTileM)) = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
    }
    if (tid < kH) {
        s_vec[0][tid] = P.b1 ? P.b1[tid] : 0.0f;
        s_vec[1][tid] = P.ln_gamma ? P.ln_gamma[tid] : 1.0f;
        s_vec[2][tid] = P.ln_beta ? P.ln_beta[tid] : 0.0f;
    }
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kBwdTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    const uint32_t tmem_row = tmem + (static_cast<uint32_t>(warp * 32) << 16);
    uint32_t phase = 0;
    const int row = tid;
    float acc_beta[2] = {0.f, 0.f}, acc_gamma[2] = {0.f, 0.f};      // lane l: columns l and 32 + l, this warp's rows
    bool first = true;

    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        const long long p0 = static_cast<long long>(tile) * kTileM;

        // ---- 1. stage X, X^T, GY, GY^T ---------------------------------------------------------
        {
            TileRegs<K1> tx;
            TileRegs<kH> tg;
            tile_issue<K1>(tx, P.x, P.x_stride, p0, P.n, warp, lane);
            tile_issue<kH>(tg, P.gy, P.gy_stride, p0, P.n, warp, lane);
            tile_store<K1>(tx, sX, sBT, kH, warp, lane);
            tile_store<kH>(tg, sGY, sAT, 0, warp, lane);
        }
        fence_async_smem();
        tc_fence_before();
        __syncthreads();

        // ---- 2. GEMM 1 (hidden recompute) and GEMM 2 (d activation) ----------------------------
        if (tid == 0) {
            tc_fence_after();
            const uint32_t xa = smem_u32(sX), w1a = smem_u32(sW1), ga = smem_u32(sGY), w2a = smem_u32(sW2T);
#pragma unroll
            for (int k = 0; k < K1 / 16; ++k)
                umma_bf16(tmem, umma_desc(xa + k * 256, 128, (K1 / 8) * 128), umma_desc(w1a + k * 256, 128, (K1 / 8) * 128),
                          umma_idesc(kH), k > 0 ? 1u : 0u);
#pragma unroll
            for (int k = 0; k < kH / 16; ++k)
                umma_bf16(tmem + kColDa, umma_desc(ga + k * 256, 128, (kH / 8) * 128), umma_desc(w2a + k * 256, 128, (kH / 8) * 128),
                          umma_idesc(kH), k > 0 ? 1u : 0u);
            umma_commit(&s_bar);
        }
        mbar_wait(&s_bar, phase);
        phase ^= 1u;
        tc_fence_after();

        // ---- 3. epilogue A: one point per thread -----------------------------------------------
        {
            float v[kH];      // h, then xhat (LayerNorm) or h (no LayerNorm)
            float d[kH];      // dz = da * act'(z), then dh
            {
                float lo[32], hi[32];
                tmem_ld32(tmem_row, lo);
                tmem_ld32(tmem_row + 32, hi);
#pragma unroll
                for (int j = 0; j < 32; ++j) { v[j] = lo[j] + s_vec[0][j]; v[32 + j] = hi[j] + s_vec[0][32 + j]; }
            }
            float rstd = 1.0f;
            if (P.layernorm) {
                float mean = 0.0f;
#pragma unroll
                for (int j = 0; j < kH; ++j) mean += v[j];
                mean *= (1.0f / kH);
                float var = 0.0f;
#pragma unroll
                for (int j = 0; j < kH; ++j) { const float t = v[j] - mean; var = fmaf(t, t, var); }
                rstd = rsqrtf(var * (1.0f / kH) + P.ln_eps);
#pragma unroll
                for (int j = 0; j < kH; ++j) v[j] = (v[j] - mean) * rstd;
            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float da[32];
                tmem_ld32(tmem_row + kColDa + 32 * half, da);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int j = 32 * half + i;
                    const float z = P.layernorm ? fmaf(v[j], s_vec[1][j], s_vec[2][j]) : v[j];
                    float a = z, g = 1.0f;
                    if (P.activation == 1) {
                        a = fmaxf(z, 0.0f);
                        g = z > 0.0f ? 1.0f : 0.0f;
                    } else if (P.activation == 2) {
                        gelu_fwd_bwd(z, a, g);
                    }
                    st_bf16(sBT, canon_elem_offset(j, row, kTileM), a);                  // A^T: row = hidden unit, column = point
                    d[j] = da[i] * g;
                }
            }
            if (P.layernorm) {
                // dbeta_j += sum_p dz ; dgamma_j += sum_p dz * xhat  (this warp's 32 points, lane l <- column l)
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float t[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) t[i] = d[32 * half + i];
                    acc_beta[half] += warp_colsum32(t, lane);
#pragma unroll
                    for (int i = 0; i < 32; ++i) t[i] = d[32 * half + i] * v[32 * half + i];
                    acc_gamma[half] += warp_colsum32(t, lane);
                }
                float m1 = 0.0f, m2 = 0.0f;
#pragma unroll
                for (int j = 0; j < kH; ++j) {
                    d[j] *= s_vec[1][j];                    // dxhat
                    m1 += d[j];
                    m2 = fmaf(d[j], v[j], m2);
                }
                m1 *= (1.0f / kH);
                m2 *= (1.0f / kH);
#pragma unroll
                for (int j = 0; j < kH; ++j) d[j] = rstd * (d[j] - m1 - v[j] * m2);
            }
#pragma unroll
            for (int k8 = 0; k8 < kH / 8; ++k8) {
                *reinterpret_cast<uint4*>(sDH + canon_chunk_offset(row, k8, kH)) =
                    make_uint4(pack_bf16(d[8 * k8], d[8 * k8 + 1]), pack_bf16(d[8 * k8 + 2], d[8 * k8 + 3]),
                               pack_bf16(d[8 * k8 + 4], d[8 * k8 + 5]), pack_bf16(d[8 * k8 + 6], d[8 * k8 + 7]));
            }
#pragma unroll
            for (int j = 0; j < kH; ++j) st_bf16(sAT, canon_elem_offset(kH + j, row, kTileM), d[j]);   // DH^T
        }
        fence_async_smem();
        tc_fence_before();
        __syncthreads();

        // ---- 4. GEMM 3 (d features) and GEMM 4 (weight / bias gradients, accumulated in TMEM) --
        if (tid == 0) {
            tc_fence_after();
            const uint32_t dha = smem_u32(sDH), w1ta = smem_u32(sW1T), ata = smem_u32(sAT), bta = smem_u32(sBT);
#pragma unroll
            for (int k = 0; k < kH / 16; ++k)
                umma_bf16(tmem + kColDx, umma_desc(dha + k * 256, 128, (kH / 8) * 128), umma_desc(w1ta + k * 256, 128, (kH / 8) * 128),
                          umma_idesc(K1), k > 0 ? 1u : 0u);
#pragma unroll
            for (int k = 0; k < kTileM / 16; ++k)
                umma_bf16(tmem + kColDw, umma_desc(ata + k * 256, 128, (kTileM / 8) * 128), umma_desc(bta + k * 256, 128, (kTileM / 8) * 128),
                          umma_idesc(N4), (!first || k > 0) ? 1u : 0u);
            umma_commit(&s_bar);
        }
        first = false;
        mbar_wait(&s_bar, phase);
        phase ^= 1u;
        tc_fence_after();

        // ---- 5. epilogue B: stream the feature-gradient row ------------------------------------
#pragma unroll
        for (int c0 = 0; c0 < K1; c0 += 16) {
            float q[16];
            tmem_ld16(tmem_row + kColDx + c0, q);
            if (p0 + row < P.n) {
                float4* dst = reinterpret_cast<float4*>(P.dx + (p0 + row) * P.dx_stride + c0);
#pragma unroll
                for (int u = 0; u < 4; ++u) __stcs(dst + u, make_float4(q[4 * u], q[4 * u + 1], q[4 * u + 2], q[4 * u + 3]));
            }
        }
    }

    // ---- weight / bias gradients of this CTA's tiles -> global (fp32 reductions) -------------------
    if (!first) {
        if (row < kH) {                    // rows 0..63: output unit o = row -> dW2[o][:], db2[o]
#pragma unroll
            for (int c0 = 0; c0 < kH; c0 += 16) {
                float q[16];
                tmem_ld16(tmem_row + kColDw + c0, q);
#pragma unroll
                for (int u = 0; u < 16; ++u) red_f32(P.dw2 + row * kH + c0 + u, q[u]);
            }
            float q[16];
            tmem_ld16(tmem_row + kColDw + kH + K1, q);
            if (P.db2) red_f32(P.db2 + row, q[0]);
        } else {                           // rows 64..127: hidden unit j = row - 64 -> dW1[j][:], db1[j]
            const int j = row - kH;
#pragma unroll
            for (int c0 = 0; c0 < K1; c0 += 16) {
                float q[16];
                tmem_ld16(tmem_row + kColDw + kH + c0, q);
#pragma unroll
                for (int u = 0; u < 16; ++u) red_f32(P.dw1 + j * K1 + c0 + u, q[u]);
            }
            float q[16];
            tmem_ld16(tmem_row + kColDw + kH + K1, q);
            if (P.db1) red_f32(P.db1 + j, q[0]);
        }
        if (P.layernorm) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                if (P.dgamma) red_f32(P.dgamma + 32 * half + lane, acc_gamma[half]);
                if (P.dbeta) red_f32(P.dbeta + 32 * half + lane, acc_beta[half]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kBwdTmemCols) : "memory");
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
template <int K1>
static cudaError_t launch_bwd_k(const MlpHeadBwdParams& P, int n_sms, cudaStream_t s) {
    auto kernel = mlp_head_bwd_kernel<K1>;
    constexpr int N4 = kH + K1 + 16;
    const size_t smem = static_cast<size_t>(kTileM) * K1 * 2 + kH * K1 * 2 + kTileM * kH * 2 + kH * kH * 2 + kTileM * kH * 2 +
                        K1 * kH * 2 + kTileM * kTileM * 2 + static_cast<size_t>(N4) * kTileM * 2;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int grid = n_sms;                                  // 512 TMEM columns: one CTA per SM
    if (grid > P.n_tiles) grid = P.n_tiles;
    kernel<<<grid, kMlpThreads, smem, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_mlp_head_bwd(MlpHeadBwdParams P, int in_dim, int n_sms, cudaStream_t s) {
    // default: the ping-pong kernel of mlp_head_bwd2.cu; DE_HEAD_BWD_VARIANT=1 selects this file's serial kernel (A/B)
    const char* v = getenv("DE_HEAD_BWD_VARIANT");
    P.variant = v ? atoi(v) : 0;
    if (!(P.variant & 1) || P.dx_group_w) return launch_mlp_head_bwd_pp(P, in_dim, n_sms, s);
    P.n_tiles = static_cast<int>((P.n + kTileM - 1) / kTileM);
    if (P.n_tiles == 0) return cudaSuccess;
    switch (in_dim) {
        case 32: return launch_bwd_k<32>(P, n_sms, s);
        case 48: return launch_bwd_k<48>(P, n_sms, s);
        case 64: return launch_bwd_k<64>(P, n_sms, s);
        case 96: return launch_bwd_k<96>(P, n_sms, s);
        case 128: return launch_bwd_k<128>(P, n_sms, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
