// st_embed.cu -- SpatialTemporalEmbedding on sm_100a (the other consumer of xyzt in the integrated model).
//
// Restates /root/reference/models/cross_modal_fusion.py:130-164:
//   x_idx = (pos[..., 0] * R).long();  y_idx = (pos[..., 1] * R).long()          (:145-146, R = 64 table rows)
//   emb[:, :D/2]  = cat(embed_x[x_idx], embed_y[y_idx])                          (:148-152)
//   emb[:, D/2:]  = Linear(D/2, D/2)(GELU(Linear(1, D/2)(t)))                    (:155-158)
//   emb          += modality                                                     (:161-162)
//   out           = tokens + emb                                                 (:164)
// The additions are performed in the reference's order (0 + e is exact, then + modality, then tokens + emb), so the
// spatial half is bit-identical to the reference; the temporal half differs only by the fp32 summation order of the
// second Linear.  `.long()` truncates toward zero and a negative index addresses from the end (PyTorch advanced
// indexing); anything outside [-R, R) -- where the reference raises IndexError -- sets *err and writes NaN.
//
// The temporal MLP is a (n x D/2) x (D/2 x D/2) contraction whose left operand GELU(t * w1 + b1) is generated on the fly;
// it runs as a 64 x 64-tile fp32 SIMT GEMM (token counts on this path are thousands, not millions; fp32 keeps the
// reference's numerics).  The same tile kernel, with other operand / epilogue functors, serves the backward.
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "launch.h"
#include "st_embed_common.cuh"

namespace de {

namespace {

constexpr int kTile = 64;   // tokens x outputs per CTA
constexpr int kTk = 16;     // contraction chunk


// (pos * R).long() with PyTorch's negative-index wrap; -1 = outside [-R, R) or not finite
__device__ __forceinline__ int table_row(float p, int R) {
    const float s = __fmul_rn(p, static_cast<float>(R));
    if (!(fabsf(s) < 3.0e9f)) return -1;                 // NaN / inf / far out of range
    long long i = __float2ll_rz(s);
    if (i < 0) i += R;
    return (i < 0 || i >= R) ? -1 : static_cast<int>(i);
}

// ---- spatial half (+ pass-through of the temporal half when no temporal positions are given) -----------------
__global__ void __launch_bounds__(256) st_embed_spatial_kernel(const StEmbedParams P, int c_end) {
    const int Dq = P.D / 4;
    const long long total = P.n * c_end;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        const long long m = e / c_end;
        const int c = static_cast<int>(e - m * c_end);
        float emb = 0.0f;
        bool has = false;
        if (P.pos && c < 2 * Dq) {
            const int which = c >= Dq;
            const int row = table_row(__ldg(P.pos + m * P.pos_stride + which), P.R);
            if (row < 0) {
                if (P.err && (c == 0 || c == Dq)) atomicOr(P.err, 1);
                emb = CUDART_NAN_F;
            } else {
                emb = __ldg((which ? P.ey : P.ex) + static_cast<long long>(row) * Dq + (c - which * Dq));
            }
            has = true;
        }
        if (P.mod) emb = has ? __fadd_rn(emb, __ldg(P.mod + c)) : __ldg(P.mod + c);
        const float tok = __ldcs(P.tokens + m * P.D + c);
        __stcs(P.out + m * P.D + c, __fadd_rn(tok, emb));
    }
}

// Same, four columns per thread (16-byte loads and stores): D a multiple of 16 -- a float4 never straddles the x / y / temporal
// column ranges -- and 16-byte aligned buffers.  The additions per element are the scalar kernel's, in its order.
__global__ void __launch_bounds__(256) st_embed_spatial_vec4_kernel(const StEmbedParams P, int c_end) {
    const int Dq = P.D / 4, c4_end = c_end / 4;
    const long long total = P.n * c4_end;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        const long long m = e / c4_end;
        const int c = static_cast<int>(e - m * c4_end) * 4;
        float4 emb = make_float4(0.f, 0.f, 0.f, 0.f);
        bool has = false;
        if (P.pos && c < 2 * Dq) {
            const int which = c >= Dq;
            const int row = table_row(__ldg(P.pos + m * P.pos_stride + which), P.R);
            if (row < 0) {
                if (P.err && (c == 0 || c == Dq)) atomicOr(P.err, 1);
                emb = make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F);
            } else {
                emb = __ldg(reinterpret_cast<const float4*>((which ? P.ey : P.ex) + static_cast<long long>(row) * Dq + (c - which * Dq)));
            }
            has = true;
        }
        if (P.mod) {
            const float4 md = __ldg(reinterpret_cast<const float4*>(P.mod + c));
            emb = has ? make_float4(__fadd_rn(emb.x, md.x), __fadd_rn(emb.y, md.y), __fadd_rn(emb.z, md.z), __fadd_rn(emb.w, md.w)) : md;
        }
        const float4 tok = __ldcs(reinterpret_cast<const float4*>(P.tokens + m * P.D + c));
        __stcs(reinterpret_cast<float4*>(P.out + m * P.D + c),
               make_float4(__fadd_rn(tok.x, emb.x), __fadd_rn(tok.y, emb.y), __fadd_rn(tok.z, emb.z), __fadd_rn(tok.w, emb.w)));
    }
}

// ---- 64 x 64 fp32 tile GEMM with functor operands:  C(m, n) = sum_k A(m, k) * B(k, n)  over k in this CTA's split ------
template <class ALoad, class BLoad, class Epi>
__global__ void __launch_bounds__(256) st_tile_gemm_kernel(long long M, int N, long long K, long long k_per_split,
                                                           ALoad a, BLoad b, Epi epi) {
    __shared__ float As[kTk][kTile + 4];
    __shared__ float Bs[kTk][kTile + 4];
    const long long m0 = static_cast<long long>(blockIdx.x) * kTile;
    const int n0 = blockIdx.y * kTile;
    const long long k_begin = static_cast<long long>(blockIdx.z) * k_per_split;
    const long long k_end = (k_begin + k_per_split < K) ? k_begin + k_per_split : K;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (long long k0 = k_begin; k0 < k_end; k0 += kTk) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int kk = threadIdx.x & 15, mm = (threadIdx.x >> 4) + 16 * i;
            As[kk][mm] = (m0 + mm < M && k0 + kk < k_end) ? a(m0 + mm, k0 + kk) : 0.0f;
            const int nn = threadIdx.x & 63, kb = (threadIdx.x >> 6) + 4 * i;
            Bs[kb][nn] = (n0 + nn < N && k0 + kb < k_end) ? b(k0 + kb, n0 + nn) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kTk; ++k) {
            const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long m = m0 + ty * 4 + i;
            const int n = n0 + tx * 4 + j;
            if (m < M && n < N) epi(m, n, acc[i][j]);
        }
}

// forward: A(m, k) = GELU(t_m * w1_k + b1_k);  B(k, n) = W2[n][k];  out[m, D/2 + n] = tokens + ((acc + b2_n) [+ mod])
struct FwdA {
    const float *t, *w1, *b1;
    __device__ float operator()(long long m, long long k) const {
        return gelu_erf(__fmaf_rn(__ldg(t + m), __ldg(w1 + k), __ldg(b1 + k)));
    }
};
struct FwdB {
    const float* w2;
    int Dh;
    __device__ float operator()(long long k, int n) const { return __ldg(w2 + static_cast<long long>(n) * Dh + k); }
};
struct FwdEpi {
    const float *tokens, *b2, *mod;
    float* out;
    int D, Dh;
    __device__ void operator()(long long m, int n, float acc) const {
        float emb = __fadd_rn(acc, __ldg(b2 + n));
        if (mod) emb = __fadd_rn(emb, __ldg(mod + Dh + n));
        const long long o = m * D + Dh + n;
        out[o] = __fadd_rn(__ldg(tokens + o), emb);
    }
};

// backward 1: dpre[m, k] = (sum_j g[m, D/2 + j] * W2[j][k]) * GELU'(t_m * w1_k + b1_k)
struct DaA {
    const float* g;
    int D, Dh;
    __device__ float operator()(long long m, long long j) const { return __ldg(g + m * D + Dh + j); }
};
struct DaB {
    const float* w2;
    int Dh;
    __device__ float operator()(long long j, int k) const { return __ldg(w2 + j * Dh + k); }
};
struct DaEpi {
    const float *t, *w1, *b1;
    float* dpre;
    int Dh;
    __device__ void operator()(long long m, int k, float acc) const {
        dpre[m * Dh + k] = acc * gelu_erf_grad(__fmaf_rn(__ldg(t + m), __ldg(w1 + k), __ldg(b1 + k)));
    }
};

// backward 2: dW2[j][k] += sum_m g[m, D/2 + j] * GELU(t_m * w1_k + b1_k)     (contraction over tokens, split over CTAs)
struct Dw2A {
    const float* g;
    int D, Dh;
    __device__ float operator()(long long j, long long m) const { return __ldg(g + m * D + Dh + j); }
};
struct Dw2B {
    const float *t, *w1, *b1;
    __device__ float operator()(long long m, int k) const {
        return gelu_erf(__fmaf_rn(__ldg(t + m), __ldg(w1 + k), __ldg(b1 + k)));
    }
};
struct Dw2Epi {
    float* dw2;
    int Dh;
    __device__ void operator()(long long j, int k, float acc) const { atomicAdd(dw2 + j * Dh + k, acc); }
};

// ---- column sums over tokens:  o1[c] += sum_m f[m, c] (* s[m]),  o2[c] += sum_m f[m, c] ------------------------------
__global__ void __launch_bounds__(256) st_colsum_kernel(const float* __restrict__ f, long long n, int n_cols, long long stride,
                                                        int col0, const float* __restrict__ scale, float* o_scaled,
                                                        float* o_plain, long long rows_per_cta, float* o_plain_hi = nullptr,
                                                        int hi_col0 = 0) {
    __shared__ float s1[8][33], s2[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int r = threadIdx.x >> 5;
    const long long m_begin = static_cast<long long>(blockIdx.y) * rows_per_cta;
    const long long m_end = (m_begin + rows_per_cta < n) ? m_begin + rows_per_cta : n;
    float a1 = 0.0f, a2 = 0.0f;
    if (c < n_cols)
        for (long long m = m_begin + r; m < m_end; m += 8) {
            const float v = __ldg(f + m * stride + col0 + c);
            a2 += v;
            if (scale) a1 = fmaf(v, __ldg(scale + m), a1);
        }
    s1[r][threadIdx.x & 31] = a1;
    s2[r][threadIdx.x & 31] = a2;
    __syncthreads();
    if (r == 0 && c < n_cols) {
#pragma unroll
        for (int i = 1; i < 8; ++i) { a1 += s1[i][threadIdx.x]; a2 += s2[i][threadIdx.x]; }
        if (o_scaled) atomicAdd(o_scaled + c, a1);
        if (o_plain) atomicAdd(o_plain + c, a2);
        if (o_plain_hi && c >= hi_col0) atomicAdd(o_plain_hi + (c - hi_col0), a2);      // a second consumer of the upper columns
    }
}

// ---- spatial tables: d_embed[row(m)][c] += g[m, c0 + c]; 64 rows take every token, so a CTA accumulates a
// (R x 32 column) slab in shared memory over its token range and flushes it once ------------------------------------
__global__ void __launch_bounds__(256) st_embed_table_grad_kernel(const float* __restrict__ g, const float* __restrict__ pos,
                                                                  long long pos_stride, long long n, int D, int R,
                                                                  float* d_ex, float* d_ey, long long rows_per_cta) {
    extern __shared__ float slab[];                    // R x 32
    const int Dq = D / 4;
    const int blocks_per_table = (Dq + 31) / 32;
    const int which = blockIdx.x / blocks_per_table;
    const int cb = (blockIdx.x - which * blocks_per_table) * 32;
    const int lane = threadIdx.x & 31, r = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < R * 32; i += 256) slab[i] = 0.0f;
    __syncthreads();
    const long long m_begin = static_cast<long long>(blockIdx.y) * rows_per_cta;
    const long long m_end = (m_begin + rows_per_cta < n) ? m_begin + rows_per_cta : n;
    if (cb + lane < Dq) {
        long long m = m_begin + r;
        for (; m + 24 < m_end; m += 32) {                        // four tokens in flight per thread
            float pv[4], gv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                pv[u] = __ldg(pos + (m + 8 * u) * pos_stride + which);
                gv[u] = __ldcs(g + (m + 8 * u) * D + which * Dq + cb + lane);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int row = table_row(pv[u], R);
                if (row >= 0) atomicAdd(&slab[row * 32 + lane], gv[u]);
            }
        }
        for (; m < m_end; m += 8) {
            const int row = table_row(__ldg(pos + m * pos_stride + which), R);
            if (row >= 0) atomicAdd(&slab[row * 32 + lane], __ldg(g + m * D + which * Dq + cb + lane));
        }
    }
    __syncthreads();
    float* dst = which ? d_ey : d_ex;
    for (int i = threadIdx.x; i < R * 32; i += 256) {
        const int row = i >> 5, c = cb + (i & 31);
        if (c < Dq && slab[i] != 0.0f) atomicAdd(dst + static_cast<long long>(row) * Dq + c, slab[i]);
    }
}

long long ceil_div(long long a, long long b) { return (a + b - 1) / b; }

}  // namespace

// the streaming kernel alone over columns [0, c_end) (st_embed_tc.cu runs the temporal half on the tensor cores)
cudaError_t launch_st_embed_spatial(const StEmbedParams& P, int c_end, cudaStream_t s) {
    const auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = P.D % 16 == 0 && c_end % 4 == 0 && al16(P.tokens) && al16(P.out) && al16(P.ex) && al16(P.ey) && al16(P.mod);
    long long blocks = ceil_div(P.n * (vec ? c_end / 4 : c_end), 256 * 4);
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    if (vec) st_embed_spatial_vec4_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(P, c_end);
    else st_embed_spatial_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(P, c_end);
    return cudaGetLastError();
}

cudaError_t launch_st_embed_forward(const StEmbedParams& P, cudaStream_t s) {
    const int Dh = P.D / 2;
    const int c_end = P.t ? Dh : P.D;                 // columns the streaming kernel owns
    cudaError_t e = launch_st_embed_spatial(P, c_end, s);
    if (e != cudaSuccess || !P.t) return e;
    dim3 grid(static_cast<unsigned>(ceil_div(P.n, kTile)), static_cast<unsigned>(ceil_div(Dh, kTile)), 1);
    st_tile_gemm_kernel<<<grid, 256, 0, s>>>(P.n, Dh, static_cast<long long>(Dh), static_cast<long long>(Dh),
                                              FwdA{P.t, P.w1, P.b1}, FwdB{P.w2, Dh},
                                              FwdEpi{P.tokens, P.b2, P.mod, P.out, P.D, Dh});
    return cudaGetLastError();
}

// everything but the temporal MLP: d(modality), d(spatial tables), d(b2)
cudaError_t launch_st_embed_backward_common(const StEmbedBwdParams& P, cudaStream_t s) {
    const int Dh = P.D / 2, Dq = P.D / 4;
    cudaError_t e = cudaSuccess;
    const bool b2_with_mod = P.d_mod && P.t && P.d_b2;      // d(b2)[j] = sum_m g[m, D/2 + j]: the upper half of the same column sums
    if (P.d_mod) {          // d(modality)[c] = sum_m g[m, c]
        const long long rows = 4096;
        dim3 grid(static_cast<unsigned>(ceil_div(P.D, 32)), static_cast<unsigned>(ceil_div(P.n, rows)));
        st_colsum_kernel<<<grid, 256, 0, s>>>(P.g, P.n, P.D, P.D, 0, nullptr, nullptr, P.d_mod, rows, b2_with_mod ? P.d_b2 : nullptr, Dh);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    if (P.pos && P.d_ex && P.d_ey) {
        const long long rows = 2048;
        dim3 grid(static_cast<unsigned>(2 * ceil_div(Dq, 32)), static_cast<unsigned>(ceil_div(P.n, rows)));
        st_embed_table_grad_kernel<<<grid, 256, static_cast<size_t>(P.R) * 32 * sizeof(float), s>>>(
            P.g, P.pos, P.pos_stride, P.n, P.D, P.R, P.d_ex, P.d_ey, rows);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    if (P.t && P.d_b2 && !b2_with_mod) {
        const long long rows = 4096;
        dim3 grid(static_cast<unsigned>(ceil_div(Dh, 32)), static_cast<unsigned>(ceil_div(P.n, rows)));
        st_colsum_kernel<<<grid, 256, 0, s>>>(P.g, P.n, Dh, P.D, Dh, nullptr, nullptr, P.d_b2, rows);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_st_embed_backward(const StEmbedBwdParams& P, cudaStream_t s) {
    const int Dh = P.D / 2;
    cudaError_t e = launch_st_embed_backward_common(P, s);
    if (e != cudaSuccess) return e;
    if (P.t) {
        const long long rows = 4096;
        {                   // dpre = (g_t . W2) * GELU'(pre)
            dim3 grid(static_cast<unsigned>(ceil_div(P.n, kTile)), static_cast<unsigned>(ceil_div(Dh, kTile)), 1);
            st_tile_gemm_kernel<<<grid, 256, 0, s>>>(P.n, Dh, static_cast<long long>(Dh), static_cast<long long>(Dh),
                                                      DaA{P.g, P.D, Dh}, DaB{P.w2, Dh}, DaEpi{P.t, P.w1, P.b1, P.dpre, Dh});
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        {                   // d(w1)[k] = sum_m dpre[m, k] * t_m ; d(b1)[k] = sum_m dpre[m, k]
            dim3 grid(static_cast<unsigned>(ceil_div(Dh, 32)), static_cast<unsigned>(ceil_div(P.n, rows)));
            st_colsum_kernel<<<grid, 256, 0, s>>>(P.dpre, P.n, Dh, Dh, 0, P.t, P.d_w1, P.d_b1, rows);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        {                   // d(W2) += g_t^T . GELU(pre): tokens are the contraction, split over blockIdx.z
            long long splits = ceil_div(P.n, 1024);
            const long long tiles = ceil_div(Dh, kTile) * ceil_div(Dh, kTile);
            if (splits * tiles > 148 * 8) splits = ceil_div(148 * 8, tiles);
            if (splits < 1) splits = 1;
            if (splits > 65535) splits = 65535;
            long long per = ceil_div(P.n, splits);
            per = ceil_div(per, kTk) * kTk;
            splits = ceil_div(P.n, per);
            dim3 grid(static_cast<unsigned>(ceil_div(Dh, kTile)), static_cast<unsigned>(ceil_div(Dh, kTile)),
                      static_cast<unsigned>(splits));
            st_tile_gemm_kernel<<<grid, 256, 0, s>>>(static_cast<long long>(Dh), Dh, P.n, per, Dw2A{P.g, P.D, Dh},
                                                      Dw2B{P.t, P.w1, P.b1}, Dw2Epi{P.d_w2, Dh});
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

}  // namespace de
