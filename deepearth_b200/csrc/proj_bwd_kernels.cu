// proj_bwd_kernels.cu -- backward of the wide projection (proj_kernels.cu), sm_100a.
//
//   forward:   h = W1 x + b1;  xh = (h - mean) rstd;  v = gamma xh + beta;  a = GELU(v);  y = W2 a + b2
//   backward:  da   = gy . W2                       TMA / tcgen05 GEMM (gemm_bf16_tma_kernel), bf16 out
//              dv   = da * GELU'(v);  t = gamma dv
//              dpre = rstd (t - mean_j t - xh mean_j (t xh))              <- proj_dpre_kernel (this file)
//              dgamma = sum_p dv xh;  dbeta = sum_p dv;  db1 = sum_p dpre  <- same kernel (warp transpose sums)
//              dW2 = gy^T a;  dW1 = dpre^T x                              GEMMs contracting over points (split-K, fp32 red)
//              dx  = dpre . W1                                            GEMM, fp32 out: the gradient the hash grids consume
//              db2 = sum_p gy                                             cast_colsum_kernel (also makes the bf16 copy of gy)
//
// proj_dpre_kernel mirrors proj_hidden_kernel: the first Linear is recomputed in 128-column TMEM chunks, twice per tile --
// pass 1 for the two row sums the LayerNorm backward needs (and dgamma / dbeta), pass 2 for dpre itself (and db1).  The
// pre-activation is never stored by the forward; the saved row statistics (mean, rstd) make the recompute exact.
#include "launch.h"
#include "mlp_common.cuh"
#include "proj_common.cuh"

namespace de {

using namespace mlp;

// fire-and-forget reduction (the pointer arrives through a parameter struct, so atomicAdd would compile to a generic ATOM)
__device__ __forceinline__ void red_f32(float* p, float v) { asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

// shared-memory accumulation through an explicit shared address (s_acc is carved out of the aligned dynamic buffer, which
// hides its address space from the compiler: atomicAdd would become a generic ATOM)
__device__ __forceinline__ void red_shared_f32(float* p, float v) {
    asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(smem_u32(p)), "f"(v) : "memory");
}

// ---- gy fp32 -> bf16 copy + column sums (db2) -----------------------------------------------------------------------
__global__ void __launch_bounds__(256) cast_colsum_kernel(const float* __restrict__ src, long long ld_src, __nv_bfloat16* __restrict__ dst,
                                                          long long ld_dst, long long n_rows, int n_cols, float* colsum,
                                                          long long rows_per_cta) {
    __shared__ float s_part[8][128];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int c0 = blockIdx.x * 128 + cx * 4;
    const long long r_begin = static_cast<long long>(blockIdx.y) * rows_per_cta;
    const long long r_end = (r_begin + rows_per_cta < n_rows) ? r_begin + rows_per_cta : n_rows;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c0 < n_cols) {                                            // n_cols is a multiple of 4
        long long r = r_begin + ry;
        for (; r + 24 < r_end; r += 32) {                         // four rows in flight per thread (the loop was one load deep)
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldcs(reinterpret_cast<const float4*>(src + (r + 8 * u) * ld_src + c0));
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
                if (dst) *reinterpret_cast<uint2*>(dst + (r + 8 * u) * ld_dst + c0) = make_uint2(pack_bf16(v[u].x, v[u].y), pack_bf16(v[u].z, v[u].w));
            }
        }
        for (; r < r_end; r += 8) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(src + r * ld_src + c0));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            if (dst) *reinterpret_cast<uint2*>(dst + r * ld_dst + c0) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
        }
    }
    if (!colsum) return;
    *reinterpret_cast<float4*>(&s_part[ry][cx * 4]) = acc;
    __syncthreads();
    if (threadIdx.x < 128) {
        float a = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) a += s_part[i][threadIdx.x];
        const int c = blockIdx.x * 128 + threadIdx.x;
        if (c < n_cols) red_f32(colsum + c, a);
    }
}

cudaError_t launch_cast_colsum(const float* src, long long ld_src, void* dst_bf16, long long ld_dst, long long n_rows, int n_cols,
                               float* colsum, int n_sms, cudaStream_t s) {
    if (n_rows == 0) return cudaSuccess;
    const int bx = (n_cols + 127) / 128;
    long long by = (static_cast<long long>(n_sms) * 8 + bx - 1) / bx;
    long long rows = (n_rows + by - 1) / by;
    if (rows < 64) rows = 64;
    rows = (rows + 7) / 8 * 8;
    by = (n_rows + rows - 1) / rows;
    cast_colsum_kernel<<<dim3(bx, static_cast<unsigned>(by)), 256, 0, s>>>(src, ld_src, static_cast<__nv_bfloat16*>(dst_bf16), ld_dst,
                                                                         n_rows, n_cols, colsum, rows);
    return cudaGetLastError();
}

// ---- column sums of a bf16 matrix (db2 when grad_y arrives in bf16: autocast) --------------------------------------------
// 32 lanes x 8 columns (one 16-byte load each) = 256 columns per CTA column, 8 row phases per CTA; n_cols is a multiple of 8.
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src, long long n_rows,
                                                          int n_cols, float* colsum, long long rows_per_cta) {
    __shared__ float s_part[8][256];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int c0 = blockIdx.x * 256 + cx * 8;
    const long long r_begin = static_cast<long long>(blockIdx.y) * rows_per_cta;
    const long long r_end = (r_begin + rows_per_cta < n_rows) ? r_begin + rows_per_cta : n_rows;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (c0 < n_cols) {
        for (long long r = r_begin + ry; r < r_end; r += 8) {
            const uint4 v = __ldcs(reinterpret_cast<const uint4*>(src + r * ld_src + c0));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                acc[2 * i] += __uint_as_float(w[i] << 16);
                acc[2 * i + 1] += __uint_as_float(w[i] & 0xffff0000u);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) s_part[ry][cx * 8 + i] = acc[i];
    __syncthreads();
    float a = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) a += s_part[i][threadIdx.x];
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < n_cols) red_f32(colsum + c, a);
}

cudaError_t launch_colsum_bf16(const void* src_bf16, long long ld_src, long long n_rows, int n_cols, float* colsum, int n_sms,
                               cudaStream_t s) {
    if (n_rows == 0 || !colsum) return cudaSuccess;
    const int bx = (n_cols + 255) / 256;
    long long by = (static_cast<long long>(n_sms) * 8 + bx - 1) / bx;
    long long rows = (n_rows + by - 1) / by;
    if (rows < 64) rows = 64;
    rows = (rows + 7) / 8 * 8;
    by = (n_rows + rows - 1) / rows;
    colsum_bf16_kernel<<<dim3(bx, static_cast<unsigned>(by)), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(src_bf16), ld_src, n_rows,
                                                                         n_cols, colsum, rows);
    return cudaGetLastError();
}

// ---- bf16 transpose (fallback operand path of the weight-gradient GEMMs, DE_PROJ_DW_TRANSPOSE=1) ----------------------
// src (rows, cols) row-major -> dst (cols, ld_dst) row-major; 64 x 64 tiles through shared memory
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src,
                                                             __nv_bfloat16* __restrict__ dst, long long ld_dst, long long rows, int cols) {
    __shared__ __nv_bfloat16 tile[64][66];
    const long long r0 = static_cast<long long>(blockIdx.x) * 64;
    const int c0 = blockIdx.y * 64;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    for (int i = ty; i < 64; i += 4) {
        const long long r = r0 + i;
        const int c = c0 + tx;
        tile[i][tx] = (r < rows && c < cols) ? src[r * ld_src + c] : __float2bfloat16_rn(0.0f);
    }
    __syncthreads();
    for (int i = ty; i < 64; i += 4) {
        const int c = c0 + i;
        const long long r = r0 + tx;
        if (c < cols && r < rows) dst[static_cast<long long>(c) * ld_dst + r] = tile[tx][i];
    }
}

cudaError_t launch_transpose_bf16(const void* src, long long ld_src, void* dst, long long ld_dst, long long rows, int cols,
                                  cudaStream_t s) {
    if (rows == 0 || cols == 0) return cudaSuccess;
    const dim3 grid(static_cast<unsigned>((rows + 63) / 64), static_cast<unsigned>((cols + 63) / 64));
    transpose_bf16_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(src), ld_src, static_cast<__nv_bfloat16*>(dst), ld_dst,
                                                rows, cols);
    return cudaGetLastError();
}

// ---- dpre ----------------------------------------------------------------------------------------------------------------
constexpr int kBwdChunkN = 128;

// 16 bf16 columns of a row: one 32-byte request (not allocated in L1), or two 16-byte ones when the rows are only 16-byte aligned
__device__ __forceinline__ void load_piece(const void* p, uint32_t (&r)[8], bool v32) {
    if (v32) {
        ld_nc_32B(p, r);
    } else {
        const uint4 a = *reinterpret_cast<const uint4*>(p), b = *(reinterpret_cast<const uint4*>(p) + 1);
        r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
    }
}
__device__ __forceinline__ void store_piece(void* p, const uint32_t (&r)[8], bool v32) {
    if (v32) {
        st_32B(p, r);
    } else {
        reinterpret_cast<uint4*>(p)[0] = make_uint4(r[0], r[1], r[2], r[3]);
        reinterpret_cast<uint4*>(p)[1] = make_uint4(r[4], r[5], r[6], r[7]);
    }
}

// W1 chunks arrive by TMA exactly as in proj_hidden_kernel (a continuous chunk stream, requested one epilogue ahead).
// dynamic shared memory (1024 B aligned): B[2] [Kp / 64 boxes x 16 KB] | A [128 x Kp] bf16 canonical | acc[3][H] floats
// Two threads per row (256 per CTA, each takes half of a chunk's columns; the two row sums are merged once per tile): the
// epilogue is instruction- and latency-bound, and 2 x 4 warps per SM issued 27 % of the time (ncu r02f).
constexpr int kDpreThreads = 256;
template <int K1>
__global__ void __launch_bounds__(kDpreThreads, 2)
proj_dpre_kernel(const __grid_constant__ CUtensorMap map_w1, const __grid_constant__ ProjDpreParams P) {
    constexpr int NT = kDpreThreads;
    __shared__ float s_part[2][kTileM][2];
    constexpr int Kp = (K1 + 63) / 64 * 64;
    constexpr uint32_t kChunkBytes = (Kp / 64) * 16384;
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_mma[2], s_tma[2];
    __shared__ uint32_t s_tmem;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    unsigned char* sB[2] = {smem, smem + kChunkBytes};
    unsigned char* sA = smem + 2 * kChunkBytes;
    const int H = P.hidden, NC = H / kBwdChunkN;
    float* s_acc = reinterpret_cast<float*>(sA + kTileM * Kp * 2);     // [3][H]: dgamma, dbeta, db1
    const int tid = threadIdx.x, warp = (tid >> 5) & 3, lane = tid & 31, half = tid >> 7;
    // the bf16 rows move as one 32-byte request per piece and lane when everything is 32-byte aligned
    const bool v32 = ((reinterpret_cast<uintptr_t>(P.dact) | reinterpret_cast<uintptr_t>(P.dpre_out)) & 31) == 0 && P.dact_stride % 16 == 0;

    for (int i = tid; i < 3 * H; i += NT) s_acc[i] = 0.0f;
    if (tid == 0) {
        mbar_init(&s_mma[0], 1); mbar_init(&s_mma[1], 1);
        mbar_init(&s_tma[0], 1); mbar_init(&s_tma[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(2 * kBwdChunkN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (Kp != K1)
        for (int i = tid; i < kTileM * Kp * 2 / 16; i += NT) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    constexpr uint32_t kIdesc = idesc_bf16(kTileM, kBwdChunkN, false, false);
    uint32_t mma_phase0 = 0u, mma_phase1 = 0u, tma_phase0 = 0u, tma_phase1 = 0u;
    const int row = tid & 127;
    const float inv_h = 1.0f / static_cast<float>(H);
    const int my_tiles = (P.n_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const long long total_chunks = static_cast<long long>(my_tiles) * 2 * NC;

    auto request_chunk = [&](long long q) {
        const int c = static_cast<int>(q % (2 * NC)) % NC;
        uint64_t* bar = &s_tma[q & 1];
        mbar_expect_tx(bar, kChunkBytes);
#pragma unroll
        for (int b = 0; b < Kp / 64; ++b) tma_load_2d(sB[q & 1] + b * 16384, &map_w1, b * 64, c * kBwdChunkN, bar);
    };
    auto issue_chunk = [&](long long q) {
        if (q & 1) { mbar_wait(&s_tma[1], tma_phase1); tma_phase1 ^= 1u; }
        else { mbar_wait(&s_tma[0], tma_phase0); tma_phase0 ^= 1u; }
        tc_fence_after();
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB[q & 1]);
#pragma unroll
        for (int k = 0; k < Kp / 16; ++k)
            umma_bf16(tmem + static_cast<uint32_t>(q & 1) * kBwdChunkN, umma_desc(a0 + k * 256, 128, (Kp / 8) * 128),
                      umma_desc_sw128(b0 + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024), kIdesc, k > 0 ? 1u : 0u);
        umma_commit(&s_mma[q & 1]);
    };
    if (tid == 0 && total_chunks > 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_w1)) : "memory");
        request_chunk(0);
        if (total_chunks > 1) request_chunk(1);
    }

    long long q0 = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, q0 += 2 * NC) {
        const long long p0 = static_cast<long long>(tile) * kTileM;
        const int n_valid = static_cast<int>(P.n - p0 < kTileM ? P.n - p0 : kTileM);
        const float* tile_base = P.x + p0 * P.x_stride;
        const int xs = static_cast<int>(P.x_stride);
        const bool a32 = ((reinterpret_cast<uintptr_t>(tile_base) | (static_cast<uintptr_t>(xs) << 2)) & 31) == 0;
        constexpr int kIt = (4 * K1) / 32;
        constexpr int kRound = kIt >= 2 ? kIt / 2 : kIt;
#pragma unroll
        for (int it0 = 0; it0 < (half == 0 ? kIt : 0); it0 += kRound) {       // the first 128 threads fill the A tile
            float4 va[kRound], vb[kRound];
#pragma unroll
            for (int u = 0; u < kRound; ++u) {
                const int c = lane + 32 * (it0 + u);
                const int r = min(warp * 32 + (c / K1) * 8 + ((c % K1) & 7), n_valid - 1);
                const int k8 = (c % K1) >> 3;
                ld_stream_32B(tile_base + (static_cast<long long>(r) * xs + k8 * 8), va[u], vb[u], a32);
            }
#pragma unroll
            for (int u = 0; u < kRound; ++u) {
                const int c = lane + 32 * (it0 + u);
                const int r = warp * 32 + (c / K1) * 8 + ((c % K1) & 7);
                const int k8 = (c % K1) >> 3;
                const uint4 pk = make_uint4(pack_bf16(va[u].x, va[u].y), pack_bf16(va[u].z, va[u].w), pack_bf16(vb[u].x, vb[u].y), pack_bf16(vb[u].z, vb[u].w));
                *reinterpret_cast<uint4*>(sA + canon_chunk_offset(r, k8, Kp)) = pk;
                // bf16 copy of the features for the dW1 GEMM (the contraction over points reads it as a (points, K1) matrix)
                if (P.x_bf16_out && r < n_valid)
                    *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(P.x_bf16_out) + (p0 + r) * K1 + k8 * 8) = pk;
            }
        }
        fence_async_smem();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) issue_chunk(q0);

        const bool live = row < n_valid;
        const float mean = live ? __ldg(P.stats + 2 * (p0 + row)) : 0.0f;
        const float rstd = live ? __ldg(P.stats + 2 * (p0 + row) + 1) : 0.0f;
        const __nv_bfloat16* da_row = static_cast<const __nv_bfloat16*>(P.dact) + (p0 + row) * P.dact_stride;
        __nv_bfloat16* out_row = static_cast<__nv_bfloat16*>(P.dpre_out) + (p0 + row) * static_cast<long long>(H);
        float s1 = 0.0f, s2 = 0.0f;                       // sum_j t_j, sum_j t_j xh_j
        f32x2 s1p = 0ull, s2p = 0ull;                     // their running pairs (even / odd columns) during pass 1

        for (int g = 0; g < 2 * NC; ++g) {
            const long long qs = q0 + g;
            const int c = g < NC ? g : g - NC;
            if (g + 1 < 2 * NC) {
                tc_fence_before();
                __syncthreads();
                if (tid == 0) issue_chunk(qs + 1);
            }
            if (qs & 1) { mbar_wait(&s_mma[1], mma_phase1); mma_phase1 ^= 1u; }
            else { mbar_wait(&s_mma[0], mma_phase0); mma_phase0 ^= 1u; }
            tc_fence_after();
            if (tid == 0 && qs + 2 < total_chunks) request_chunk(qs + 2);
            const uint32_t t_base = tmem + lane_off + static_cast<uint32_t>(qs & 1) * kBwdChunkN;
            // The thread's 64 columns of the chunk come as four 16-column pieces (ncu r02n: with 32-column pieces the kernel sat
            // at the 128-register cap, and the reloads of its spilled loop state were ~12 % of the stall samples); piece i + 1
            // is requested from tensor memory as soon as piece i has arrived, and a piece's global loads go out before its wait.
            constexpr int NP = kBwdChunkN / (NT / 128) / 16;
            const uint32_t t_mine = t_base + half * (kBwdChunkN / 2);
            const int col_mine = c * kBwdChunkN + half * (kBwdChunkN / 2);
            uint32_t rb[2][16];
            tmem_ld16_issue(t_mine, rb[0]);
            const f32x2 nmean2 = splat2(-mean), rstd2 = splat2(rstd);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int col0 = col_mine + i * 16;
                const float4* b1v = reinterpret_cast<const float4*>(P.b1 + col0);           // 16-byte aligned (C ABI)
                uint4* piece = reinterpret_cast<uint4*>(out_row + col0);                    // this row's 16 columns of dpre_out
                float4 bb[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) bb[u] = __ldg(b1v + u);
                if (g < NC) {
                    // ---- pass 1: dv = da GELU'(v), t = gamma dv; row sums of t and t xh; dgamma, dbeta.  t is parked (bf16)
                    //      in dpre_out so that pass 2 needs neither da nor the GELU derivative again
                    uint32_t dw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                    if (live) load_piece(da_row + col0, dw, v32);
                    const float4* gav = reinterpret_cast<const float4*>(P.ln_gamma + col0);
                    const float4* bev = reinterpret_cast<const float4*>(P.ln_beta + col0);
                    float4 gg[4], ee[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) { gg[u] = __ldg(gav + u); ee[u] = __ldg(bev + u); }
                    tmem_ld16_wait(rb[i & 1]);
                    if (i + 1 < NP) tmem_ld16_issue(t_mine + (i + 1) * 16, rb[(i + 1) & 1]);
                    float dvv[16], dvx[16];
                    uint32_t tp[8];
                    // packed pairs (mlp_common.cuh): columns (2u, 2u + 1) of the piece share one FFMA2 / FMUL2 / FADD2
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int q4 = u >> 1;
                        const f32x2 b2 = (u & 1) ? pk2(bb[q4].z, bb[q4].w) : pk2(bb[q4].x, bb[q4].y);
                        const f32x2 g2 = (u & 1) ? pk2(gg[q4].z, gg[q4].w) : pk2(gg[q4].x, gg[q4].y);
                        const f32x2 e2 = (u & 1) ? pk2(ee[q4].z, ee[q4].w) : pk2(ee[q4].x, ee[q4].y);
                        const f32x2 xh = mul2(add2(add2(pk2u(rb[i & 1][2 * u], rb[i & 1][2 * u + 1]), b2), nmean2), rstd2);
                        f32x2 act, gp;
                        gelu_fwd_bwd2(fma2(g2, xh, e2), act, gp);
                        const uint32_t pr = dw[u];
                        const f32x2 dv = mul2(pk2u(pr << 16, pr & 0xffff0000u), gp);               // bf16 -> fp32, times GELU'
                        const f32x2 tt = mul2(g2, dv);
                        s1p = add2(s1p, tt);
                        s2p = fma2(tt, xh, s2p);
                        upk2(dv, dvv[2 * u], dvv[2 * u + 1]);                        // column sums: dbeta
                        upk2(mul2(dv, xh), dvx[2 * u], dvx[2 * u + 1]);              //              dgamma
                        float t0, t1;
                        upk2(tt, t0, t1);
                        tp[u] = pack_bf16(t0, t1);
                    }
                    if (live) store_piece(piece, tp, v32);
                    const float sb = warp_transpose_sum16(dvv, lane);
                    const float sg = warp_transpose_sum16(dvx, lane);
                    if (lane < 16) {
                        red_shared_f32(&s_acc[H + col0 + lane], sb);
                        red_shared_f32(&s_acc[col0 + lane], sg);
                    }
                } else {
                    // ---- pass 2: dpre = rstd (t - mean t - xh mean(t xh)); db1 ---------------------------------------------
                    uint32_t tw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                    if (live) load_piece(piece, tw, v32);                                             // written by this thread in pass 1
                    tmem_ld16_wait(rb[i & 1]);
                    if (i + 1 < NP) tmem_ld16_issue(t_mine + (i + 1) * 16, rb[(i + 1) & 1]);
                    const f32x2 nm1 = splat2(-(s1 * inv_h)), nm2 = splat2(-(s2 * inv_h));
                    float dd[16];
                    uint32_t pk[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int q4 = u >> 1;
                        const f32x2 b2 = (u & 1) ? pk2(bb[q4].z, bb[q4].w) : pk2(bb[q4].x, bb[q4].y);
                        const f32x2 xh = mul2(add2(add2(pk2u(rb[i & 1][2 * u], rb[i & 1][2 * u + 1]), b2), nmean2), rstd2);
                        const uint32_t pr = tw[u];
                        const f32x2 t = pk2u(pr << 16, pr & 0xffff0000u);
                        float d0, d1;
                        upk2(mul2(rstd2, fma2(xh, nm2, add2(t, nm1))), d0, d1);
                        pk[u] = pack_bf16(d0, d1);
                        // db1 sums the bf16-ROUNDED dpre, the values the dW1 / dx GEMMs consume
                        dd[2 * u] = live ? __uint_as_float(pk[u] << 16) : 0.0f;
                        dd[2 * u + 1] = live ? __uint_as_float(pk[u] & 0xffff0000u) : 0.0f;
                    }
                    if (live) store_piece(piece, pk, v32);
                    const float sd = warp_transpose_sum16(dd, lane);
                    if (lane < 16) red_shared_f32(&s_acc[2 * H + col0 + lane], sd);
                }
            }
            if (g == NC - 1) {                                    // both halves of a row need the row's full sums for pass 2
                {
                    float a0, a1, c0, c1;
                    upk2(s1p, a0, a1);
                    upk2(s2p, c0, c1);
                    s1 = a0 + a1;
                    s2 = c0 + c1;
                }
                s_part[half][row][0] = s1;
                s_part[half][row][1] = s2;
                __syncthreads();
                s1 = s_part[0][row][0] + s_part[1][row][0];
                s2 = s_part[0][row][1] + s_part[1][row][1];
            }
        }
        tc_fence_before();
        __syncthreads();
    }
    // one flush per CTA
    for (int i = tid; i < H; i += NT) {
        if (P.dgamma) red_f32(P.dgamma + i, s_acc[i]);
        if (P.dbeta) red_f32(P.dbeta + i, s_acc[H + i]);
        if (P.db1) red_f32(P.db1 + i, s_acc[2 * H + i]);
    }
    tc_fence_before();
    __syncthreads();
    if (tid < 32)                     // ONE warp: with 256 threads "warp" (the TMEM lane quarter) is 0 for warps 0 and 4, and a second
                                      // dealloc of the same columns traps (intermittently: only when it sees the first one's update)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * kBwdChunkN) : "memory");
}

template <int K1>
static cudaError_t launch_dpre_k(const ProjDpreParams& P, int n_sms, cudaStream_t s) {
    constexpr int Kp = (K1 + 63) / 64 * 64;
    const size_t smem = static_cast<size_t>(2) * (Kp / 64) * 16384 + static_cast<size_t>(kTileM) * Kp * 2 +
                        static_cast<size_t>(3) * P.hidden * sizeof(float) + 1024;
    CUtensorMap map_w1;
    if (!make_map_bf16(&map_w1, P.w1, P.hidden, K1, K1, kBwdChunkN)) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(proj_dpre_kernel<K1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = static_cast<int>(232448u / (smem + 2048));       // 227 KB of shared memory per SM, ~1 KB reserved per CTA
    if (per_sm > 2) per_sm = 2;
    if (per_sm < 1) per_sm = 1;
    int grid = n_sms * per_sm;
    if (grid > P.n_tiles) grid = P.n_tiles;
    proj_dpre_kernel<K1><<<grid, kDpreThreads, smem, s>>>(map_w1, P);
    return cudaGetLastError();
}

cudaError_t launch_proj_dpre(ProjDpreParams P, int in_dim, int n_sms, cudaStream_t s) {
    P.n_tiles = static_cast<int>((P.n + kTileM - 1) / kTileM);
    switch (in_dim) {
        case 32: return launch_dpre_k<32>(P, n_sms, s);
        case 48: return launch_dpre_k<48>(P, n_sms, s);
        case 64: return launch_dpre_k<64>(P, n_sms, s);
        case 96: return launch_dpre_k<96>(P, n_sms, s);
        case 128: return launch_dpre_k<128>(P, n_sms, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
