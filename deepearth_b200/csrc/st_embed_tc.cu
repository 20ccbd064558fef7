// st_embed_tc.cu -- the temporal MLP of SpatialTemporalEmbedding (models/cross_modal_fusion.py:155-158) on the tensor cores.
//
//   emb_t = W2 . GELU(t w1 + b1) + b2      (Linear(1, D/2) -> GELU -> Linear(D/2, D/2), fp32 in the reference)
//
// The second Linear is 2 (D/2)^2 flop per token: 17 TFLOP/s through the fp32 SIMT tile GEMM of st_embed.cu -- at D = 2048 half the
// speed of the reference's own cuBLAS SGEMM on the same B200.  Here it runs on gemm_bf16_tma_kernel (proj_kernels.cu: TMA +
// tcgen05, CTA pairs) at fp32 accuracy: every fp32 operand x is split into three bf16 parts,
//       x0 = bf16(x),  x1 = bf16(x - x0),  x2 = bf16(x - x0 - x1)        (8 + 8 + 8 mantissa bits: x = x0 + x1 + x2 to 2^-24 |x|)
// and the product is the six terms  a2 b0 + a1 b1 + a0 b2 + a1 b0 + a0 b1 + a0 b0  (small first; the dropped ones are below 2^-24 |a||b|),
// accumulated in fp32 in tensor memory as ONE GEMM whose contraction runs over six operand segments (GemmParams::n_seg): the parts lie
// side by side in the operands' columns and the TMA producer picks (part of A, part of B) per segment -- nothing is replicated.
// Tokens are processed in slabs so that the split operands stay a bounded workspace.
//
//   forward:   H'' = split3(GELU(t w1 + b1))  (slab, 3 D/2);  out[:, D/2:] = tokens + (H'' . W2''^T + (b2 + mod))   [residual epilogue]
//   backward:  dpre = (split3(g_t) . split3(W2^T)^T) * GELU'(pre)  [GELU' in the GEMM epilogue];  d w1, d b1 = column sums of dpre (* t);
//              d W2 += split3(g_t)^T . H''
//              (contraction over tokens: row-contracted operands, split over the CTAs, fp32 red.global.add)
#include <cuda_bf16.h>

#include "launch.h"
#include "st_embed_common.cuh"

namespace de {

namespace {

__device__ __forceinline__ void split3(float x, __nv_bfloat16& p0, __nv_bfloat16& p1, __nv_bfloat16& p2) {
    p0 = __float2bfloat16_rn(x);
    const float r1 = __fsub_rn(x, __bfloat162float(p0));          // exact
    p1 = __float2bfloat16_rn(r1);
    p2 = __float2bfloat16_rn(__fsub_rn(r1, __bfloat162float(p1)));
}

// eight values -> their three parts, each as one 16-byte store
__device__ __forceinline__ void split3_store8(const float (&x)[8], __nv_bfloat16* d, long long part_stride) {
    __align__(16) __nv_bfloat16 p[3][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split3(x[i], p[0][i], p[1][i], p[2][i]);
#pragma unroll
    for (int q = 0; q < 3; ++q) *reinterpret_cast<uint4*>(d + q * part_stride) = *reinterpret_cast<const uint4*>(p[q]);
}

// dst (rows, 3 cols) bf16 = the three parts of src(r, c) = src[r * ld_r + c * ld_c], side by side
__global__ void __launch_bounds__(256) split3_kernel(const float* __restrict__ src, long long ld_r, long long ld_c, long long rows,
                                                     int cols, __nv_bfloat16* __restrict__ dst) {
    const long long total = rows * cols;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        const long long r = e / cols;
        const int c = static_cast<int>(e - r * cols);
        __nv_bfloat16 p0, p1, p2;
        split3(__ldg(src + r * ld_r + c * ld_c), p0, p1, p2);
        __nv_bfloat16* d = dst + r * 3 * cols + c;
        d[0] = p0; d[cols] = p1; d[2 * cols] = p2;
    }
}

// same for rows that are contiguous in memory (ld_c = 1), eight columns per thread: cols % 8 == 0, 16-byte aligned rows
__global__ void __launch_bounds__(256) split3_rows8_kernel(const float* __restrict__ src, long long ld_r, long long rows, int cols,
                                                           __nv_bfloat16* __restrict__ dst) {
    const int c8n = cols / 8;
    const long long total = rows * c8n;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        const long long r = e / c8n;
        const int c = static_cast<int>(e - r * c8n) * 8;
        const float4 a = __ldcs(reinterpret_cast<const float4*>(src + r * ld_r + c)), b = __ldcs(reinterpret_cast<const float4*>(src + r * ld_r + c) + 1);
        const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        split3_store8(x, dst + r * 3 * cols + c, cols);
    }
}

// dst (rows, 3 Dh) = split3(GELU(t[r] * w1[k] + b1[k])), eight hidden units per thread (Dh % 8 == 0)
__global__ void __launch_bounds__(256) split3_hidden_kernel(const float* __restrict__ t, const float* __restrict__ w1,
                                                            const float* __restrict__ b1, long long rows, int Dh,
                                                            __nv_bfloat16* __restrict__ dst) {
    const int k8n = Dh / 8;
    const long long total = rows * k8n;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        const long long r = e / k8n;
        const int k = static_cast<int>(e - r * k8n) * 8;
        const float tr = __ldg(t + r);
        const float4 wa = __ldg(reinterpret_cast<const float4*>(w1 + k)), wb = __ldg(reinterpret_cast<const float4*>(w1 + k) + 1);
        const float4 ba = __ldg(reinterpret_cast<const float4*>(b1 + k)), bb = __ldg(reinterpret_cast<const float4*>(b1 + k) + 1);
        const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w}, b[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = gelu_erf(__fmaf_rn(tr, w[i], b[i]));
        split3_store8(x, dst + r * 3 * Dh + k, Dh);
    }
}

// bias[j] = b2[j] (+ mod[Dh + j])
__global__ void bias_sum_kernel(const float* __restrict__ b2, const float* __restrict__ mod_t, int Dh, float* __restrict__ out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < Dh) out[j] = mod_t ? __fadd_rn(__ldg(b2 + j), __ldg(mod_t + j)) : __ldg(b2 + j);
}

// d w1[k] += sum_m dpre[m, k] * t_m ;  d b1[k] += sum_m dpre[m, k], dpre = c1 * GELU'(t_m w1_k + b1_k).  APPLY = false: c1 already
// carries the GELU' factor (the GEMM epilogue applied it under its mainloop).
template <bool APPLY>
__global__ void __launch_bounds__(256) dpre_colsum_kernel(const float* __restrict__ c1, const float* __restrict__ t,
                                                          const float* __restrict__ w1, const float* __restrict__ b1, long long n,
                                                          int Dh, float* d_w1, float* d_b1, long long rows_per_cta) {
    __shared__ float s1[8][33], s2[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int r = threadIdx.x >> 5;
    const long long m_begin = static_cast<long long>(blockIdx.y) * rows_per_cta;
    const long long m_end = (m_begin + rows_per_cta < n) ? m_begin + rows_per_cta : n;
    float a1 = 0.0f, a2 = 0.0f;
    if (c < Dh) {
        const float w = __ldg(w1 + c), b = __ldg(b1 + c);
        for (long long m = m_begin + r; m < m_end; m += 8) {
            const float tm = __ldg(t + m);
            float v = __ldcs(c1 + m * Dh + c);
            if constexpr (APPLY) v *= gelu_erf_grad(__fmaf_rn(tm, w, b));
            a2 += v;
            a1 = fmaf(v, tm, a1);
        }
    }
    s1[r][threadIdx.x & 31] = a1;
    s2[r][threadIdx.x & 31] = a2;
    __syncthreads();
    if (r == 0 && c < Dh) {
#pragma unroll
        for (int i = 1; i < 8; ++i) { a1 += s1[i][threadIdx.x]; a2 += s2[i][threadIdx.x]; }
        if (d_w1) atomicAdd(d_w1 + c, a1);
        if (d_b1) atomicAdd(d_b1 + c, a2);
    }
}

long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }
size_t align256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

unsigned grid_for(long long elems) {
    long long b = ceil_div_ll(elems, 256 * 4);
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return static_cast<unsigned>(b);
}

// tokens per slab: the split operands of one slab are 6 Dh bytes per token and operand
long long slab_tokens(long long n, int Dh) {
    const long long slab = Dh >= 512 ? 65536 : 262144;
    return n < slab ? (n + 63) / 64 * 64 : slab;
}

// the six (part of A, part of B) products, SMALL terms first: the tensor core truncates its fp32 accumulator once per MMA
// instruction, by up to an ulp of the running sum whatever the size of the addend -- with the a0 b0 segment first, the five
// segments of small terms cost five sixths of those truncations at full magnitude and contribute almost nothing (measured at
// D = 2048: d w1 1.2e-5 of max against fp64 with large terms first; the order below is what the tests print now)
void set_segments(GemmParams& G, int seg_blocks, long long stride) {
    static const unsigned char sa[6] = {2, 1, 0, 1, 0, 0}, sb[6] = {0, 1, 2, 0, 1, 0};
    G.n_seg = 6;
    G.seg_blocks = seg_blocks;
    G.seg_stride_a = G.seg_stride_b = stride;
    for (int i = 0; i < 6; ++i) { G.seg_a[i] = sa[i]; G.seg_b[i] = sb[i]; }
}

}  // namespace

bool st_embed_tc_supported(int D) { return D >= 128 && D % 128 == 0; }      // D / 2 a multiple of the 64-element k-block

size_t st_embed_tc_workspace_bytes(long long n, int D, bool backward) {
    if (n <= 0 || !st_embed_tc_supported(D)) return 0;
    const size_t Dh = static_cast<size_t>(D / 2), slab = static_cast<size_t>(slab_tokens(n, D / 2));
    const size_t w = align256(Dh * 3 * Dh * 2), op = align256(slab * 3 * Dh * 2);
    return backward ? w + 2 * op + align256(slab * Dh * 4) : w + align256(Dh * 4) + op;
}

cudaError_t launch_st_embed_forward_tc(const StEmbedParams& P, void* ws, int n_sms, cudaStream_t s) {
    const int Dh = P.D / 2;
    cudaError_t e = launch_st_embed_spatial(P, P.t ? Dh : P.D, s);
    if (e != cudaSuccess || !P.t) return e;
    const long long slab = slab_tokens(P.n, Dh);
    char* w = static_cast<char*>(ws);
    __nv_bfloat16* w2s = reinterpret_cast<__nv_bfloat16*>(w);
    float* bias = reinterpret_cast<float*>(w + align256(static_cast<size_t>(Dh) * 3 * Dh * 2));
    __nv_bfloat16* hs = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(bias) + align256(static_cast<size_t>(Dh) * 4));
    split3_kernel<<<grid_for(static_cast<long long>(Dh) * Dh), 256, 0, s>>>(P.w2, Dh, 1, Dh, Dh, w2s);     // B(n, k) = W2[n][k]
    bias_sum_kernel<<<(Dh + 255) / 256, 256, 0, s>>>(P.b2, P.mod ? P.mod + Dh : nullptr, Dh, bias);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    for (long long m0 = 0; m0 < P.n; m0 += slab) {
        const long long ns = P.n - m0 < slab ? P.n - m0 : slab;
        split3_hidden_kernel<<<grid_for(ns * Dh / 8), 256, 0, s>>>(P.t + m0, P.w1, P.b1, ns, Dh, hs);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        GemmParams G{};
        G.M = ns; G.N = Dh; G.K = 6LL * Dh;
        G.c = P.out + m0 * P.D + Dh; G.ldc = P.D; G.bias = bias;
        G.residual = P.tokens + m0 * P.D + Dh; G.ldr = P.D;
        set_segments(G, Dh / 64, Dh);
        G.a_rows = ns; G.a_cols = 3LL * Dh; G.b_rows = Dh; G.b_cols = 3LL * Dh;
        if ((e = launch_gemm_bf16(hs, 3LL * Dh, w2s, 3LL * Dh, 0, G, 0, n_sms, s)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_st_embed_backward_tc(const StEmbedBwdParams& P, void* ws, int n_sms, cudaStream_t s) {
    const int Dh = P.D / 2;
    cudaError_t e = launch_st_embed_backward_common(P, s);
    if (e != cudaSuccess || !P.t) return e;
    const long long slab = slab_tokens(P.n, Dh);
    const size_t op = align256(static_cast<size_t>(slab) * 3 * Dh * 2);
    char* w = static_cast<char*>(ws);
    __nv_bfloat16* w2ts = reinterpret_cast<__nv_bfloat16*>(w);
    __nv_bfloat16* hs = reinterpret_cast<__nv_bfloat16*>(w + align256(static_cast<size_t>(Dh) * 3 * Dh * 2));
    __nv_bfloat16* gs = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(hs) + op);
    float* c1 = reinterpret_cast<float*>(reinterpret_cast<char*>(gs) + op);
    // B of the first product: B(n = k, kk = j) = W2[j][k]
    split3_kernel<<<grid_for(static_cast<long long>(Dh) * Dh), 256, 0, s>>>(P.w2, 1, Dh, Dh, Dh, w2ts);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    for (long long m0 = 0; m0 < P.n; m0 += slab) {
        const long long ns = P.n - m0 < slab ? P.n - m0 : slab;
        if ((reinterpret_cast<uintptr_t>(P.g) & 15) == 0) split3_rows8_kernel<<<grid_for(ns * Dh / 8), 256, 0, s>>>(P.g + m0 * P.D + Dh, P.D, ns, Dh, gs);
        else split3_kernel<<<grid_for(ns * Dh), 256, 0, s>>>(P.g + m0 * P.D + Dh, P.D, 1, ns, Dh, gs);
        split3_hidden_kernel<<<grid_for(ns * Dh / 8), 256, 0, s>>>(P.t + m0, P.w1, P.b1, ns, Dh, hs);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        {       // C1 (ns, Dh) = g_t . W2
            GemmParams G{};
            G.M = ns; G.N = Dh; G.K = 6LL * Dh; G.c = c1; G.ldc = Dh;
            G.gg_row_t = P.t + m0; G.gg_col_w = P.w1; G.gg_col_b = P.b1;         // C1 leaves the GEMM as dpre
            set_segments(G, Dh / 64, Dh);
            G.a_rows = ns; G.a_cols = 3LL * Dh; G.b_rows = Dh; G.b_cols = 3LL * Dh;
            if ((e = launch_gemm_bf16(gs, 3LL * Dh, w2ts, 3LL * Dh, 0, G, 0, n_sms, s)) != cudaSuccess) return e;
        }
        {       // d w1, d b1
            const long long rows = 4096;
            dim3 grid(static_cast<unsigned>((Dh + 31) / 32), static_cast<unsigned>(ceil_div_ll(ns, rows)));
            dpre_colsum_kernel<false><<<grid, 256, 0, s>>>(c1, P.t + m0, P.w1, P.b1, ns, Dh, P.d_w1, P.d_b1, rows);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        {       // d W2[j][k] += sum_m g_t[m, j] h[m, k]: the tokens are the contraction (rows of both operands)
            const int seg_blocks = static_cast<int>(ceil_div_ll(ns, 64));
            GemmParams G{};
            G.M = Dh; G.N = Dh; G.K = 6LL * seg_blocks * 64; G.c = P.d_w2; G.ldc = Dh;
            set_segments(G, seg_blocks, Dh);
            G.a_rows = ns; G.a_cols = 3LL * Dh; G.b_rows = ns; G.b_cols = 3LL * Dh;
            if ((e = launch_gemm_bf16(gs, 3LL * Dh, hs, 3LL * Dh, 1, G, 2, n_sms, s)) != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

}  // namespace de
