// proj_kernels.cu -- Grid4DEncoder.projection at its real width (hidden_dim = 768 by default, 2048 in the integrated
// model) on the 5th-generation tensor cores (sm_100a: tcgen05.mma + TMEM + TMA).
//
//   y = W2 . GELU( LayerNorm( W1 . x + b1 ) ) + b2        /root/reference/models/encoders.py:47-52, models/configs.py:44
//
// Two kernels, because a 128-row tile's hidden row (H >= 768 fp32 accumulators) does not fit tensor memory (512 columns)
// next to the output accumulators, and the LayerNorm statistics need the whole row before any of it can be normalised:
//
//   proj_hidden_kernel<K1, NT>   x (n, K1) fp32  ->  a = GELU(LN(W1 x + b1)) (n, H) bf16.   The first Linear is small
//       (K1 = 48 or 128), so it is simply run TWICE per tile in 128-column chunks: pass 1 reads the chunks back from TMEM only
//       to accumulate the row statistics (chunk-wise mean / M2, merged with Chan's formula -- no E[h^2] - mean^2
//       cancellation), pass 2 recomputes each chunk, normalises, applies GELU and writes bf16.  The pre-activation never
//       exists in memory; the tile's features are read once, the hidden row is written once.  W1 chunks arrive by TMA as a
//       continuous stream through a double buffer, one epilogue ahead; two threads per row; the epilogue runs on packed fp32
//       pairs (FFMA2 / FADD2 / FMUL2), software-pipelined 16-column TMEM pieces and one 32-byte store per piece.
//
//   gemm_bf16_tma_kernel<A_MN, B_MN, EPI, MC>   C (M, N) = A (M, K) . B (N, K)^T [+ bias] [+ residual]: the second Linear
//       (and, with other operands, the GEMMs of the backward pass and of st_embed_tc.cu).  Persistent, warp specialised: warp
//       0 issues TMA loads (cp.async.bulk.tensor, 128 B swizzle) into the operand ring, warp 1 issues tcgen05.mma into one of
//       two 256-column TMEM accumulators, warps 4..11 drain the other one, so the epilogue of tile i overlaps the mainloop of
//       tile i + 1.  MC = 2 (default): CTA pairs -- cta_group::2 MMAs with M = 256, each CTA stages its 128 rows of A and
//       half of B (6-stage ring of 32 KB); MC = 1: single CTAs, M = 128 (4 stages of 48 KB).  fp32 output leaves through
//       swizzled shared memory and TMA tensor stores; split-K partial sums by red.global.add.v4.f32; the contraction may run
//       over operand segments (GemmParams::n_seg: split-precision products).
//
// Operand layouts follow cute::UMMA (CUTLASS 4.x, cute/atom/mma_traits_sm100.hpp): K-major SWIZZLE_128B tiles are 8-row x
// 128-byte atoms (stride byte offset 1024), exactly what a TMA box {64 bf16, rows} with CU_TENSOR_MAP_SWIZZLE_128B writes.
#include <cuda.h>

#include <cstdlib>

#include "launch.h"
#include "mlp_common.cuh"
#include "proj_common.cuh"
#include "st_embed_common.cuh"

namespace de {

using namespace mlp;

// ---------------------------------------------------------------------------------------------------------------------
// kernel 1: hidden activations
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kChunkN = 128;                                    // hidden columns per MMA chunk / TMEM buffer
constexpr int kHiddenThreadsDefault = 256;            // B200 (r02e): forward 4.13 -> 3.42 ms per 2^20 rows against 128 threads

// W1 chunks arrive by TMA: {64 k, 128 rows} boxes with the 128 B swizzle, Kp / 64 boxes per chunk (Kp = K1 rounded up to 64;
// columns past K1 are zero-filled by the copy engine, and the A tile carries the same zero columns).  The chunk stream is
// continuous across the CTA's tiles -- chunk q + 2 is requested as soon as the MMAs of chunk q have released its buffer,
// i.e. one whole epilogue ahead of its use -- so no thread ever waits for weights.
// dynamic shared memory (1024 B aligned): B[2] [Kp / 64 boxes x 16 KB] | A [128 x Kp] bf16, canonical K-major no-swizzle
// NT = 128: one thread per row does the whole chunk; NT = 256: two threads per row, each takes half of a chunk's columns
// (the LayerNorm / GELU epilogue is instruction-bound at 8 warps per SM) and the row statistics are merged once per tile.
template <int K1, int NT>
__global__ void __launch_bounds__(NT)
proj_hidden_kernel(const __grid_constant__ CUtensorMap map_w1, const __grid_constant__ ProjHiddenParams P) {
    constexpr int NH = NT / 128;                                // threads per row
    __shared__ float s_stat[NH][kTileM][3];
    constexpr int Kp = (K1 + 63) / 64 * 64;
    constexpr uint32_t kChunkBytes = (Kp / 64) * 16384;
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_mma[2], s_tma[2];
    __shared__ uint32_t s_tmem;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    unsigned char* sB[2] = {smem, smem + kChunkBytes};
    unsigned char* sA = smem + 2 * kChunkBytes;
    const int tid = threadIdx.x, warp = (tid >> 5) & 3, lane = tid & 31, half = tid >> 7;
    const int H = P.hidden, NC = H / kChunkN;
    const bool st32 = (reinterpret_cast<uintptr_t>(P.hidden_out) & 31) == 0 && P.hidden_stride % 16 == 0;

    if (tid == 0) {
        mbar_init(&s_mma[0], 1); mbar_init(&s_mma[1], 1);
        mbar_init(&s_tma[0], 1); mbar_init(&s_tma[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(2 * kChunkN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (Kp != K1)                                               // the pad columns of the A tile stay zero for the CTA's life
        for (int i = tid; i < kTileM * Kp * 2 / 16; i += NT) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    constexpr uint32_t kIdesc = idesc_bf16(kTileM, kChunkN, false, false);
    uint32_t mma_phase0 = 0u, mma_phase1 = 0u;                  // every thread waits on the MMA barriers
    uint32_t tma_phase0 = 0u, tma_phase1 = 0u;                  // thread 0 only
    const int row = tid & 127;
    const int my_tiles = (P.n_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const long long total_chunks = static_cast<long long>(my_tiles) * 2 * NC;

    auto request_chunk = [&](long long q) {                     // thread 0: W1 rows of stream position q -> buffer q & 1
        const int c = static_cast<int>(q % (2 * NC)) % NC;
        uint64_t* bar = &s_tma[q & 1];
        mbar_expect_tx(bar, kChunkBytes);
#pragma unroll
        for (int b = 0; b < Kp / 64; ++b) tma_load_2d(sB[q & 1] + b * 16384, &map_w1, b * 64, c * kChunkN, bar);
    };
    auto issue_chunk = [&](long long q) {                       // thread 0: wait for the weights, then the chunk's MMAs
        if (q & 1) { mbar_wait(&s_tma[1], tma_phase1); tma_phase1 ^= 1u; }
        else { mbar_wait(&s_tma[0], tma_phase0); tma_phase0 ^= 1u; }
        tc_fence_after();
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB[q & 1]);
#pragma unroll
        for (int k = 0; k < Kp / 16; ++k)
            umma_bf16(tmem + static_cast<uint32_t>(q & 1) * kChunkN, umma_desc(a0 + k * 256, 128, (Kp / 8) * 128),
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
        // ---- features -> bf16 A tile (lane mapping of mlp_head.cu: conflict-free 16-byte shared stores) ----------------
        constexpr int kIt = (4 * K1) / 32;
        constexpr int kRound = kIt >= 2 ? kIt / 2 : kIt;
#pragma unroll
        for (int it0 = 0; it0 < (half == 0 ? kIt : 0); it0 += kRound) {
            float4 va[kRound], vb[kRound];
#pragma unroll
            for (int u = 0; u < kRound; ++u) {
                const int c = lane + 32 * (it0 + u);
                const int r = min(warp * 32 + (c / K1) * 8 + ((c % K1) & 7), n_valid - 1);     // tail rows re-read the last valid row
                const int k8 = (c % K1) >> 3;
                ld_stream_32B(tile_base + (static_cast<long long>(r) * xs + k8 * 8), va[u], vb[u], a32);
            }
#pragma unroll
            for (int u = 0; u < kRound; ++u) {
                const int c = lane + 32 * (it0 + u);
                const int r = warp * 32 + (c / K1) * 8 + ((c % K1) & 7);
                const int k8 = (c % K1) >> 3;
                *reinterpret_cast<uint4*>(sA + canon_chunk_offset(r, k8, Kp)) =
                    make_uint4(pack_bf16(va[u].x, va[u].y), pack_bf16(va[u].z, va[u].w), pack_bf16(vb[u].x, vb[u].y), pack_bf16(vb[u].z, vb[u].w));
            }
        }
        fence_async_smem();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) issue_chunk(q0);

        float cnt = 0.0f, mean = 0.0f, m2 = 0.0f, rstd = 0.0f;
        __nv_bfloat16* out_row = static_cast<__nv_bfloat16*>(P.hidden_out) + (p0 + row) * static_cast<long long>(P.hidden_stride);
        for (int g = 0; g < 2 * NC; ++g) {
            const long long q = q0 + g;
            const int c = g < NC ? g : g - NC;
            // the next chunk's MMAs run while this chunk's epilogue executes
            if (g + 1 < 2 * NC) {
                tc_fence_before();
                __syncthreads();                      // everyone is done reading TMEM buffer (q + 1) & 1 (chunk q - 1)
                if (tid == 0) issue_chunk(q + 1);
            }
            if (q & 1) { mbar_wait(&s_mma[1], mma_phase1); mma_phase1 ^= 1u; }
            else { mbar_wait(&s_mma[0], mma_phase0); mma_phase0 ^= 1u; }
            tc_fence_after();
            if (tid == 0 && q + 2 < total_chunks) request_chunk(q + 2);      // buffer q & 1 is free again
            const uint32_t t_base = tmem + lane_off + static_cast<uint32_t>(q & 1) * kChunkN;
            // The thread's columns of the chunk come in 16-column pieces; piece i + 1 is requested from tensor memory as soon
            // as piece i has arrived, and the constants of a piece are requested before its wait (ncu r02g: a third of the
            // stall samples sat on the first FADD after each 32-column tcgen05.ld).
            constexpr int NP = kChunkN / NH / 16;
            const uint32_t t_mine = t_base + half * (kChunkN / NH);
            const int col_mine = c * kChunkN + half * (kChunkN / NH);
            uint32_t rb[2][16];
            tmem_ld16_issue(t_mine, rb[0]);
            if (g < NC) {
                // ---- pass 1: row statistics ---------------------------------------------------------------------------
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const float4* b1 = reinterpret_cast<const float4*>(P.b1 + col_mine + i * 16);      // 16-byte aligned (C ABI)
                    float4 bb[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) bb[u] = __ldg(b1 + u);
                    tmem_ld16_wait(rb[i & 1]);
                    if (i + 1 < NP) tmem_ld16_issue(t_mine + (i + 1) * 16, rb[(i + 1) & 1]);
                    // packed pairs (FADD2 / FFMA2): half the fp32-pipe instructions of the scalar form
                    f32x2 v[8];
                    f32x2 sa = 0ull, sb = 0ull;                     // (+0, +0)
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        v[2 * u] = add2(pk2u(rb[i & 1][4 * u], rb[i & 1][4 * u + 1]), pk2(bb[u].x, bb[u].y));
                        v[2 * u + 1] = add2(pk2u(rb[i & 1][4 * u + 2], rb[i & 1][4 * u + 3]), pk2(bb[u].z, bb[u].w));
                        sa = add2(sa, v[2 * u]);
                        sb = add2(sb, v[2 * u + 1]);
                    }
                    float s0, s1, s2, s3;
                    upk2(sa, s0, s1);
                    upk2(sb, s2, s3);
                    const float mb = ((s0 + s1) + (s2 + s3)) * (1.0f / 16.0f);
                    const f32x2 nmb = splat2(-mb);
                    sa = sb = 0ull;
#pragma unroll
                    for (int j = 0; j < 8; j += 2) {
                        const f32x2 d0 = add2(v[j], nmb), d1 = add2(v[j + 1], nmb);
                        sa = fma2(d0, d0, sa);
                        sb = fma2(d1, d1, sb);
                    }
                    upk2(sa, s0, s1);
                    upk2(sb, s2, s3);
                    chan_merge(cnt, mean, m2, 16.0f, mb, (s0 + s1) + (s2 + s3));
                }
                if (g == NC - 1) {
                    if constexpr (NH > 1) {                      // merge the two halves' statistics, in a fixed order
                        s_stat[half][row][0] = cnt; s_stat[half][row][1] = mean; s_stat[half][row][2] = m2;
                        __syncthreads();
                        cnt = s_stat[0][row][0]; mean = s_stat[0][row][1]; m2 = s_stat[0][row][2];
                        chan_merge(cnt, mean, m2, s_stat[1][row][0], s_stat[1][row][1], s_stat[1][row][2]);
                    }
                    rstd = rsqrtf(m2 / static_cast<float>(H) + P.ln_eps);
                    if (P.stats_out && row < n_valid && half == 0) {
                        P.stats_out[2 * (p0 + row)] = mean;
                        P.stats_out[2 * (p0 + row) + 1] = rstd;
                    }
                }
            } else {
                // ---- pass 2: normalise, GELU, bf16, store ---------------------------------------------------------------
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int col0 = col_mine + i * 16;
                    const float4* b1 = reinterpret_cast<const float4*>(P.b1 + col0);
                    const float4* ga = reinterpret_cast<const float4*>(P.ln_gamma + col0);
                    const float4* be = reinterpret_cast<const float4*>(P.ln_beta + col0);
                    float4 bb[4], gg[4], ee[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) { bb[u] = __ldg(b1 + u); gg[u] = __ldg(ga + u); ee[u] = __ldg(be + u); }
                    tmem_ld16_wait(rb[i & 1]);
                    if (i + 1 < NP) tmem_ld16_issue(t_mine + (i + 1) * 16, rb[(i + 1) & 1]);
                    uint32_t packed[8];
                    const f32x2 nmean = splat2(-mean), rstd2 = splat2(rstd);
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int q4 = u >> 1;
                        const f32x2 b = (u & 1) ? pk2(bb[q4].z, bb[q4].w) : pk2(bb[q4].x, bb[q4].y);
                        const f32x2 ga2 = (u & 1) ? pk2(gg[q4].z, gg[q4].w) : pk2(gg[q4].x, gg[q4].y);
                        const f32x2 be2 = (u & 1) ? pk2(ee[q4].z, ee[q4].w) : pk2(ee[q4].x, ee[q4].y);
                        const f32x2 pre = add2(add2(pk2u(rb[i & 1][2 * u], rb[i & 1][2 * u + 1]), b), nmean);
                        const f32x2 act = gelu_fwd2(fma2(mul2(pre, rstd2), ga2, be2));
                        float lo, hi;
                        upk2(act, lo, hi);
                        packed[u] = pack_bf16(lo, hi);
                    }
                    if (row < n_valid) {
                        if (st32) {                                 // rows 32-byte aligned: one whole-sector store per piece
                            st_32B(out_row + col0, packed);
                        } else {
                            uint4* dst = reinterpret_cast<uint4*>(out_row + col0);
                            dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                            dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
                        }
                    }
                }
            }
        }
        tc_fence_before();
        __syncthreads();          // the next tile overwrites the A tile and TMEM buffer q0 & 1
    }
    tc_fence_before();
    __syncthreads();
    if (tid < 32)                     // ONE warp: with 256 threads "warp" (the TMEM lane quarter) is 0 for warps 0 and 4, and a second
                                      // dealloc of the same columns traps (intermittently: only when it sees the first one's update)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * kChunkN) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------------
// kernel 2: C = A . B^T (+ bias), bf16 operands by TMA, fp32 accumulation in TMEM
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kGemmBM = 128, kGemmBN = 256, kGemmBK = 64;
constexpr int kGemmMaxStages = 6;
constexpr int kGemmThreads = 384;                      // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warps 4..11 epilogue
constexpr uint32_t kGemmRingBytes = 192 * 1024;
constexpr uint32_t kGemmStoreBytes = 8 * 4096;               // one 32 x 32 fp32 staging tile per epilogue warp (TMA stores)
constexpr uint32_t kGemmSmemBytes = kGemmRingBytes + kGemmStoreBytes + 1024;      // + alignment slack
constexpr int kGemmPrefetch = 12;                    // k-blocks the L2 prefetch cursor runs ahead of the loads
constexpr int kGemmPairDefault = 1;                   // B200 r02p: fwd+bwd 11.06 -> 10.58 ms at 2^20 x 768 against single CTAs

// A_MN / B_MN = 1: the operand is stored with the contraction index as its ROW index ((K, M) resp. (K, N) row-major: a
// "transposed" operand, MN-major in UMMA terms) and arrives as 64 x 64 boxes; 0: (M, K) resp. (N, K) row-major, K-major.
// EPI 0: C fp32 = acc + bias; 1: C bf16 = acc + bias; 2: C fp32 += acc (red.global.add: split-K partial sums).
// MC = 2: CTA PAIRS (tcgen05 cta_group::2).  The two CTAs of a cluster (the two SMs of a TPC) run one M = 256 MMA on a
// 256 x 256 output tile: each CTA loads its own 128 rows of A and HALF of the B tile (128 of the 256 N rows) into its own
// shared memory and holds the accumulator rows of its A half in its own TMEM; the leader (rank 0) issues every MMA and its
// commits are multicast to both CTAs' barriers.  Per CTA and k-block that is 32 KB of L2 reads and 8 KB of shared-memory
// operand reads per MMA instruction instead of 48 KB and 12 KB: the single-CTA kernel keeps the shared-memory pipe busy
// 3/4 of the time with operand reads alone (ncu r02m: l1tex the busiest unit at 68 %, tensor pipe 55 %), which is why the
// vendor's kernels for these shapes are all 2-CTA ones.  Ring: 6 stages of 32 KB (MC = 1: 4 of 48 KB).
// live MMA width of a tile with n_left columns to its right edge: a multiple of 16 per CTA's share
template <int MC>
__device__ __forceinline__ int gemm_n_eff(int n_left) {
    const int n = n_left < kGemmBN ? n_left : kGemmBN;
    return (n + 16 * MC - 1) / (16 * MC) * (16 * MC);
}

template <int A_MN, int B_MN, int EPI, int MC>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tma_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_c, const __grid_constant__ GemmParams P) {
    constexpr int kGemmStages = MC == 2 ? 6 : 4;
    constexpr uint32_t kGemmStageBytes = (kGemmBM + kGemmBN / MC) * kGemmBK * 2;         // this CTA's share of a k-block
    static_assert(kGemmStages <= kGemmMaxStages && kGemmStages * kGemmStageBytes <= kGemmRingBytes, "operand ring");
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t s_full[kGemmMaxStages], s_empty[kGemmMaxStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (P.N + kGemmBN - 1) / kGemmBN;
    const long long tiles_m = (P.M + kGemmBM - 1) / kGemmBM;
    // With operand segments a split is a range of k-blocks INSIDE a segment, walked once per segment (small-term segments
    // first): every work item then ends with its own share of the large a0 b0 products instead of some items holding nothing
    // else -- the tensor core truncates the accumulator once per MMA, so what counts is how many MMAs run at full magnitude.
    const int k_blocks_all = P.n_seg ? P.seg_blocks : static_cast<int>((P.K + kGemmBK - 1) / kGemmBK);
    const int kb_per = (k_blocks_all + P.k_splits - 1) / P.k_splits;          // host guarantees every split is non-empty
    const int seg_rep = P.n_seg ? P.n_seg : 1;
    const long long tiles_mc = (tiles_m + MC - 1) / MC;                        // MC vertically adjacent tiles per cluster
    const long long n_tiles = tiles_mc * tiles_n * P.k_splits;                 // work items: (output tile [pair], K split)
    const uint32_t rank = MC > 1 ? cluster_rank() : 0u;
    const long long unit = blockIdx.x / MC, n_units = gridDim.x / MC;          // the cluster is the scheduling unit

    if (threadIdx.x == 0) {
        for (int i = 0; i < kGemmStages; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 8 * MC); }   // MC = 2: the leader's counts both CTAs' epilogue warps
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                                   // MC = 2: the same warp of both CTAs, collectively
        if constexpr (MC == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    if constexpr (MC > 1) cluster_sync_all();          // the peer's barriers exist before anything can arrive on them
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp == 0 && lane == 0) {
        // ===== TMA producer =====
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        int stage = 0;
        uint32_t phase = 0;
        // The ring holds 4 k-blocks (192 KB): ~2000 clocks of tensor work, about one HBM round trip -- ncu r02e showed the
        // tensor pipe 42 % active, waiting for operands that stream from HBM.  A second cursor therefore runs kGemmPrefetch
        // k-blocks ahead of the loads (across tile boundaries) and asks L2 for the streamed operand boxes, so the loads
        // themselves find their data in L2.
        long long pf_work = unit;
        int pf_kb = 0, pf_end = 0;
        auto pf_range = [&]() {
            const int sp = static_cast<int>(pf_work % P.k_splits);
            pf_kb = sp * kb_per;
            pf_end = min(k_blocks_all, (sp + 1) * kb_per);
        };
        auto pf_step = [&]() {
            if (pf_work >= n_tiles || !P.prefetch) return;
            const long long t = pf_work / P.k_splits;
            const int pm0 = static_cast<int>((t / tiles_n) * MC + rank) * kGemmBM, pn0 = static_cast<int>(t % tiles_n) * kGemmBN;
            if constexpr (A_MN) {
#pragma unroll
                for (int i = 0; i < kGemmBM / 64; ++i) tma_prefetch_2d(&map_a, pm0 + 64 * i, pf_kb * kGemmBK);
            } else {
                tma_prefetch_2d(&map_a, pf_kb * kGemmBK, pm0);
            }
            if constexpr (B_MN) {                                  // weight-gradient form: B streams from HBM as well
#pragma unroll
                for (int i = 0; i < kGemmBN / 64; ++i)
                    if (MC == 1 || (i >> 1) == static_cast<int>(rank)) tma_prefetch_2d(&map_b, pn0 + 64 * i, pf_kb * kGemmBK);
            }
            if (++pf_kb == pf_end) { pf_work += n_units; if (pf_work < n_tiles) pf_range(); }
        };
        if (pf_work < n_tiles) pf_range();
        for (int i = 0; i < kGemmPrefetch; ++i) pf_step();
        for (long long work = unit; work < n_tiles; work += n_units) {
            const long long tile = work / P.k_splits;
            const int split = static_cast<int>(work - tile * P.k_splits);
            const int m0 = static_cast<int>((tile / tiles_n) * MC + rank) * kGemmBM, n0 = static_cast<int>(tile % tiles_n) * kGemmBN;
            const int kb_begin = split * kb_per, kb_len = min(k_blocks_all, (split + 1) * kb_per) - kb_begin;
            for (int sg = 0, kb = kb_begin, it = 0; it < seg_rep * kb_len; ++it, ++kb) {
                if (kb == kb_begin + kb_len) { kb = kb_begin; ++sg; }          // next segment over the same k-block range
                pf_step();
                mbar_wait(&s_empty[stage], phase ^ 1u);
                unsigned char* sa = smem + stage * kGemmStageBytes;
                unsigned char* sb = sa + kGemmBM * kGemmBK * 2;
                // coordinates of this k-block in the operands (operand segments: see GemmParams)
                int ka = kb * kGemmBK, kbb = kb * kGemmBK, am0 = m0, bn0 = n0;
                if (P.n_seg) {
                    const int within = kb * kGemmBK;
                    if constexpr (A_MN) { ka = within; am0 = m0 + static_cast<int>(P.seg_a[sg] * P.seg_stride_a); }
                    else ka = static_cast<int>(P.seg_a[sg] * P.seg_stride_a) + within;
                    if constexpr (B_MN) { kbb = within; bn0 = n0 + static_cast<int>(P.seg_b[sg] * P.seg_stride_b); }
                    else kbb = static_cast<int>(P.seg_b[sg] * P.seg_stride_b) + within;
                }
                if constexpr (MC == 2) {
                    // both CTAs' boxes report to the LEADER's full barrier, which the leader arms for the pair's bytes
                    if (rank == 0) mbar_expect_tx(&s_full[stage], 2 * kGemmStageBytes);
                    const uint32_t full = mapa_u32(smem_u32(&s_full[stage]), 0);
                    // this CTA's half of the B tile: the MMA reads the first n_eff / 2 rows of each CTA's B buffer
                    const int nb = bn0 + static_cast<int>(rank) * (gemm_n_eff<MC>(P.full_n ? kGemmBN : P.N - n0) / 2);
                    if constexpr (A_MN) {
#pragma unroll
                        for (int i = 0; i < kGemmBM / 64; ++i) tma_load_2d_pair(sa + i * 8192, &map_a, am0 + 64 * i, ka, full);
                    } else {
                        tma_load_2d_pair(sa, &map_a, ka, m0, full);
                    }
                    if constexpr (B_MN) {
#pragma unroll
                        for (int i = 0; i < kGemmBN / 128; ++i) tma_load_2d_pair(sb + i * 8192, &map_b, nb + 64 * i, kbb, full);
                    } else {
                        tma_load_2d_pair(sb, &map_b, kbb, nb, full);
                    }
                } else {
                    mbar_expect_tx(&s_full[stage], kGemmStageBytes);
                    if constexpr (A_MN) {
#pragma unroll
                        for (int i = 0; i < kGemmBM / 64; ++i) tma_load_2d(sa + i * 8192, &map_a, am0 + 64 * i, ka, &s_full[stage]);
                    } else {
                        tma_load_2d(sa, &map_a, ka, m0, &s_full[stage]);
                    }
                    if constexpr (B_MN) {
#pragma unroll
                        for (int i = 0; i < kGemmBN / 64; ++i) tma_load_2d(sb + i * 8192, &map_b, bn0 + 64 * i, kbb, &s_full[stage]);
                    } else {
                        tma_load_2d(sb, &map_b, kbb, n0, &s_full[stage]);
                    }
                }
                if (++stage == kGemmStages) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1 && lane == 0 && rank == 0) {
        // ===== MMA issuer (MC = 2: the pair's leader) =====
        int stage = 0;
        uint32_t phase = 0;
        int as = 0;
        uint32_t aphase = 0;
        for (long long work = unit; work < n_tiles; work += n_units) {
            const int split = static_cast<int>(work % P.k_splits);
            const int kb_begin = split * kb_per, kb_end = min(k_blocks_all, (split + 1) * kb_per);
            // the MMA's N is the tile's live width (a multiple of 16): a 128-column matrix (dx, dW1 at in_dim 128) costs half
            // the tensor time of a full 256-column tile instead of multiplying zero-filled rows
            const int n_left = P.N - static_cast<int>((work / P.k_splits) % tiles_n) * kGemmBN;
            const uint32_t kIdesc = idesc_bf16(kGemmBM * MC, gemm_n_eff<MC>(P.full_n ? kGemmBN : n_left), A_MN != 0, B_MN != 0);
            mbar_wait(&s_tempty[as], aphase ^ 1u);                  // epilogue has drained this accumulator
            tc_fence_after();
            const int n_it = seg_rep * (kb_end - kb_begin);
            for (int it = 0; it < n_it; ++it) {
                mbar_wait(&s_full[stage], phase);
                tc_fence_after();
                const uint32_t a0 = smem_u32(smem + stage * kGemmStageBytes);
                const uint32_t b0 = a0 + kGemmBM * kGemmBK * 2;
#pragma unroll
                for (int k = 0; k < kGemmBK / 16; ++k) {
                    // K-major: 16 elements = 32 B further along the 128 B swizzled row; MN-major: 16 rows of 128 B = two
                    // 1024 B atoms further (LBO = 8192 B between 64-wide MN blocks, SBO = 1024 B between 8-row K groups)
                    const uint64_t da = A_MN ? umma_desc_sw128(a0 + k * 2048, 8192, 1024) : umma_desc_sw128(a0 + k * 32, 16, 1024);
                    const uint64_t db = B_MN ? umma_desc_sw128(b0 + k * 2048, 8192, 1024) : umma_desc_sw128(b0 + k * 32, 16, 1024);
                    if constexpr (MC == 2) umma_bf16_pair(tmem + as * kGemmBN, da, db, kIdesc, (it > 0 || k) ? 1u : 0u);
                    else umma_bf16(tmem + as * kGemmBN, da, db, kIdesc, (it > 0 || k) ? 1u : 0u);
                }
                if constexpr (MC == 2) {
                    umma_commit_pair(&s_empty[stage], 3);            // the slot is refilled by BOTH producers: tell both
                    if (it == n_it - 1) umma_commit_pair(&s_tfull[as], 3);         // each CTA drains its own half
                } else {
                    umma_commit(&s_empty[stage]);                    // frees the slot when these MMAs have read it
                    if (it == n_it - 1) umma_commit(&s_tfull[as]);   // accumulator complete
                }
                if (++stage == kGemmStages) { stage = 0; phase ^= 1u; }
            }
            if (++as == 2) { as = 0; aphase ^= 1u; }
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> global =====
        // eight warps: two per TMEM lane quarter, each draining half of the accumulator's columns.  With four, the drain of a
        // 128 x 256 fp32 tile (8 dependent tcgen05.ld -> bias -> store rounds per thread) took about as long as the next
        // tile's mainloop and the MMA warp waited for a free accumulator (tensor pipe 42 % active with fp32 output, 57 % with
        // bf16, ncu r02e).
        const int quarter = warp & 3;                                // TMEM lanes this warp may read
        const int col_half = (warp - 4) >> 2;
        int as = 0;
        uint32_t aphase = 0;
        for (long long work = unit; work < n_tiles; work += n_units) {
            const long long tile = work / P.k_splits;
            const long long m0 = ((tile / tiles_n) * MC + rank) * kGemmBM;
            const int n0 = static_cast<int>(tile % tiles_n) * kGemmBN;
            mbar_wait(&s_tfull[as], aphase);
            tc_fence_after();
            const long long r = m0 + quarter * 32 + lane;
            const uint32_t t_base = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + as * kGemmBN;
#pragma unroll 1
            for (int q = col_half * (kGemmBN / 64); q < (col_half + 1) * (kGemmBN / 64); ++q) {
                const int col0 = n0 + q * 32;
                if (col0 >= P.N) break;                              // warp-uniform
                float v[32];
                tmem_ld32(t_base + q * 32, v);
                if (P.bias) {
                    if (col0 + 32 <= P.N && P.bias_vec) {
                        const float4* b4 = reinterpret_cast<const float4*>(P.bias + col0);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 b = __ldg(b4 + j);
                            v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (col0 + j < P.N) ? __ldg(P.bias + col0 + j) : 0.0f;
                    }
                }
                if (EPI == 0 && P.residual && r < P.M) {
                    const float* rp = P.residual + r * P.ldr + col0;
                    if (col0 + 32 <= P.N && ((reinterpret_cast<uintptr_t>(rp) & 15) == 0)) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 q4 = __ldcs(reinterpret_cast<const float4*>(rp) + j);
                            v[4 * j] += q4.x; v[4 * j + 1] += q4.y; v[4 * j + 2] += q4.z; v[4 * j + 3] += q4.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (col0 + j < P.N) ? __ldcs(rp + j) : 0.0f;
                    }
                }
                if (EPI == 0 && P.gg_row_t) {
                    const float tr = r < P.M ? __ldg(P.gg_row_t + r) : 0.0f;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (col0 + j < P.N) v[j] *= gelu_erf_grad(__fmaf_rn(tr, __ldg(P.gg_col_w + col0 + j), __ldg(P.gg_col_b + col0 + j)));
                }
                if (EPI == 0 && P.tma_store) {
                    // fp32 rows through shared memory and a tensor store: the 32 lanes' 32 x 128 B leave as one box of whole
                    // lines (ncu r02p: with per-lane 32 B stores the fp32 epilogue, not the tensor pipe (58 % active) or HBM
                    // (45 %), set the pace).  128 B swizzle: lane r keeps 16-byte chunk c at chunk c ^ (r & 7) -- conflict-free
                    // stores, and the layout the tensor map (SWIZZLE_128B, box 32 x 32) expects.  The copy engine clips the box
                    // at the matrix edge.
                    unsigned char* stg = smem + kGemmRingBytes + (warp - 4) * 4096;
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");      // the previous box has left
                    __syncwarp();
                    const uint32_t row_addr = smem_u32(stg) + lane * 128;
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + ((c ^ (lane & 7)) << 4)), "f"(v[4 * c]),
                                     "f"(v[4 * c + 1]), "f"(v[4 * c + 2]), "f"(v[4 * c + 3]) : "memory");
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];"
                                     ::"l"(reinterpret_cast<uint64_t>(&map_c)), "r"(col0), "r"(static_cast<int>(m0) + quarter * 32),
                                       "r"(smem_u32(stg)) : "memory");
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                } else if (r < P.M) {
                    if constexpr (EPI == 2) {
                        float* dst = static_cast<float*>(P.c) + r * P.ldc + col0;
                        if (col0 + 32 <= P.N && P.vec_ok) {
#pragma unroll
                            for (int u = 0; u < 8; ++u)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * u), "f"(v[4 * u]), "f"(v[4 * u + 1]),
                                             "f"(v[4 * u + 2]), "f"(v[4 * u + 3]) : "memory");
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < P.N) atomicAdd(dst + j, v[j]);
                        }
                    } else if constexpr (EPI == 1) {
                        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(P.c) + r * P.ldc + col0;
                        if (col0 + 32 <= P.N && P.vec_ok) {
                            // two 32-byte stores: whole sectors (four 16-byte stores left every sector half written per instruction)
#pragma unroll
                            for (int u = 0; u < 2; ++u)
                                asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * u),
                                             "r"(pack_bf16(v[16 * u], v[16 * u + 1])), "r"(pack_bf16(v[16 * u + 2], v[16 * u + 3])),
                                             "r"(pack_bf16(v[16 * u + 4], v[16 * u + 5])), "r"(pack_bf16(v[16 * u + 6], v[16 * u + 7])),
                                             "r"(pack_bf16(v[16 * u + 8], v[16 * u + 9])), "r"(pack_bf16(v[16 * u + 10], v[16 * u + 11])),
                                             "r"(pack_bf16(v[16 * u + 12], v[16 * u + 13])), "r"(pack_bf16(v[16 * u + 14], v[16 * u + 15])) : "memory");
                        } else {
                            #pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < P.N) dst[j] = __float2bfloat16_rn(v[j]);
                        }
                    } else {
                        float* dst = static_cast<float*>(P.c) + r * P.ldc + col0;
                        if (col0 + 32 <= P.N && P.vec_ok) {
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                st_stream_32B(dst + 8 * u, make_float4(v[8 * u], v[8 * u + 1], v[8 * u + 2], v[8 * u + 3]),
                                              make_float4(v[8 * u + 4], v[8 * u + 5], v[8 * u + 6], v[8 * u + 7]), true);
                        } else {
                            #pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < P.N) dst[j] = v[j];
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (MC == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&s_tempty[as]), 0));     // the leader issues the MMAs
                else mbar_arrive(&s_tempty[as]);
            }
            if (++as == 2) { as = 0; aphase ^= 1u; }
        }
    }
    if (EPI == 0 && warp >= 4 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // tensor stores complete
    tc_fence_before();
    __syncthreads();
    if constexpr (MC > 1) cluster_sync_all();          // no CTA leaves while its peer may still write into it or signal it
    if (warp == 2) {
        if constexpr (MC == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

}  // namespace

// row-major bf16 matrix (rows, cols), leading dimension ld elements; box = {64 columns, box_rows}, 128 B swizzle.
// K-major operand (M or N rows, K columns): box_rows = the tile's MN extent; MN-major operand (K rows, MN columns): 64.
bool make_map_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
    const cuuint32_t box[2] = {64u, static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1u, 1u};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool proj_hidden_supported(int in_dim, int hidden) {
    const bool k_ok = in_dim == 32 || in_dim == 48 || in_dim == 64 || in_dim == 96 || in_dim == 128;
    return k_ok && hidden >= kChunkN && hidden % kChunkN == 0 && hidden <= 4096;
}

template <int K1, int NT>
static cudaError_t launch_hidden_k(const ProjHiddenParams& P, int n_sms, cudaStream_t s) {
    constexpr int Kp = (K1 + 63) / 64 * 64;
    const size_t smem = static_cast<size_t>(2) * (Kp / 64) * 16384 + static_cast<size_t>(kTileM) * Kp * 2 + 1024;
    CUtensorMap map_w1;
    if (!make_map_bf16(&map_w1, P.w1, P.hidden, K1, K1, kChunkN)) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(proj_hidden_kernel<K1, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = static_cast<int>(232448u / (smem + 2048));       // 227 KB of shared memory per SM, ~1 KB reserved per CTA
    if (per_sm > 2) per_sm = 2;                                   // 2 x 256 TMEM columns
    if (per_sm < 1) per_sm = 1;
    int grid = n_sms * per_sm;
    if (grid > P.n_tiles) grid = P.n_tiles;
    proj_hidden_kernel<K1, NT><<<grid, NT, smem, s>>>(map_w1, P);
    return cudaGetLastError();
}

cudaError_t launch_proj_hidden(ProjHiddenParams P, int in_dim, int n_sms, cudaStream_t s) {
    P.n_tiles = static_cast<int>((P.n + kTileM - 1) / kTileM);
    static const int threads = []() {                    // DE_PROJ_HIDDEN_THREADS (read once): 128 or 256 threads per CTA
        const char* e = getenv("DE_PROJ_HIDDEN_THREADS");
        return (e && atoi(e) == 128) ? 128 : ((e && atoi(e) == 256) ? 256 : kHiddenThreadsDefault);
    }();
#define DE_HID(K) (threads == 256 ? launch_hidden_k<K, 256>(P, n_sms, s) : launch_hidden_k<K, 128>(P, n_sms, s))
    switch (in_dim) {
        case 32: return DE_HID(32);
        case 48: return DE_HID(48);
        case 64: return DE_HID(64);
        case 96: return DE_HID(96);
        case 128: return DE_HID(128);
        default: return cudaErrorInvalidValue;
    }
#undef DE_HID
}

// fp32 matrix (rows, cols), rows ld elements apart: box {32 columns, 32 rows}, 128 B swizzle -- the epilogue's tensor stores
static bool make_map_f32_store(CUtensorMap* map, void* base, long long rows, long long cols, long long ld) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
    const cuuint32_t box[2] = {32u, 32u};
    const cuuint32_t estr[2] = {1u, 1u};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int A_MN, int B_MN, int EPI, int MC>
static cudaError_t launch_gemm_t(const CUtensorMap& map_a, const CUtensorMap& map_b, const CUtensorMap& map_c, const GemmParams& P, int grid,
                                 cudaStream_t s) {
    auto kernel = gemm_bf16_tma_kernel<A_MN, B_MN, EPI, MC>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kGemmSmemBytes));
    if (e != cudaSuccess) return e;
    if (MC == 1) {
        kernel<<<grid, kGemmThreads, kGemmSmemBytes, s>>>(map_a, map_b, map_c, P);
        return cudaGetLastError();
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(static_cast<unsigned>(grid));
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = kGemmSmemBytes;
    cfg.stream = s;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = MC;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, map_a, map_b, map_c, P);
}

template <int A_MN, int B_MN, int MC>
static cudaError_t launch_gemm_e(const CUtensorMap& map_a, const CUtensorMap& map_b, const CUtensorMap& map_c, const GemmParams& P, int epi,
                                 int grid, cudaStream_t s) {
    switch (epi) {
        case 0: return launch_gemm_t<A_MN, B_MN, 0, MC>(map_a, map_b, map_c, P, grid, s);
        case 1: return launch_gemm_t<A_MN, B_MN, 1, MC>(map_a, map_b, map_c, P, grid, s);
        default: return launch_gemm_t<A_MN, B_MN, 2, MC>(map_a, map_b, map_c, P, grid, s);
    }
}

// DE_GEMM_PAIR (read once per process): 1 = CTA pairs (cta_group::2, 256 x 256 tiles), 0 = single CTAs (128 x 256)
static int gemm_pair() {
    static const int v = []() {
        const char* e = getenv("DE_GEMM_PAIR");
        return (e && *e) ? (atoi(e) != 0 ? 1 : 0) : kGemmPairDefault;
    }();
    return v;
}

// C (M, N) = A . B^T (+ bias | accumulated).  transposed = 0: A is (M, K), B is (N, K) row-major (K contiguous);
// transposed = 1: A is (K, M), B is (K, N) row-major -- the contraction runs over ROWS (weight gradients: K = points).
// epi: 0 fp32 store, 1 bf16 store, 2 fp32 atomic accumulation with the K range cut into `k_splits` work items per tile.
cudaError_t launch_gemm_bf16(const void* a, long long lda, const void* b, long long ldb, int transposed, GemmParams P, int epi,
                             int n_sms, cudaStream_t s) {
    const long long tiles_m = (P.M + kGemmBM - 1) / kGemmBM, tiles_n = (P.N + kGemmBN - 1) / kGemmBN;
    const int mc = (gemm_pair() && tiles_m >= 2) ? 2 : 1;
    CUtensorMap map_a, map_b;
    bool ok;
    // (rows, cols) of the operand matrices as stored; with operand segments the caller states them (K is the segmented total)
    const long long ar = P.n_seg ? P.a_rows : (transposed ? P.K : P.M), ac = P.n_seg ? P.a_cols : (transposed ? P.M : P.K);
    const long long br = P.n_seg ? P.b_rows : (transposed ? P.K : static_cast<long long>(P.N));
    const long long bc = P.n_seg ? P.b_cols : (transposed ? static_cast<long long>(P.N) : P.K);
    if (transposed) ok = make_map_bf16(&map_a, a, ar, ac, lda, 64) && make_map_bf16(&map_b, b, br, bc, ldb, 64);
    else ok = make_map_bf16(&map_a, a, ar, ac, lda, kGemmBM) && make_map_bf16(&map_b, b, br, bc, ldb, kGemmBN / mc);
    if (!ok) return cudaErrorInvalidValue;
    const long long units = ((tiles_m + mc - 1) / mc) * tiles_n;           // scheduling units: tiles or vertical tile pairs
    const long long max_units = n_sms / mc;
    const long long k_blocks = P.n_seg ? P.seg_blocks : (P.K + kGemmBK - 1) / kGemmBK;      // what the splits cut (see the kernel)
    long long splits = 1;
    if (epi == 2) {
        // fill the machine ONCE: units * splits <= the resident CTAs (pairs).  Rounding up instead (9 splits of 9 tile pairs
        // on 74 CTA pairs = 81 work items) left 7 pairs with two items and 67 with one: 1.46 ms for 1.24 TFLOP (ncu r02p)
        splits = max_units / units;
        if (splits > k_blocks) splits = k_blocks;
        if (splits < 1) splits = 1;
        const long long per = (k_blocks + splits - 1) / splits;
        splits = (k_blocks + per - 1) / per;                        // every split non-empty
    }
    P.k_splits = static_cast<int>(splits);
    static const int prefetch = []() { const char* e = getenv("DE_GEMM_PREFETCH"); return (e && *e) ? (atoi(e) != 0) : 0; }();   // B200 r02g: 13.7 ms with, 11.9 ms without
    P.prefetch = prefetch && !P.n_seg;
    static const int full_n = []() { const char* e = getenv("DE_GEMM_FULL_N"); return (e && *e) ? (atoi(e) != 0) : 0; }();
    P.full_n = full_n;
    const long long work = units * splits;
    const int grid = static_cast<int>((work < max_units ? work : max_units) * mc);
    const size_t elem = epi == 1 ? 2 : 4;
    P.vec_ok = (reinterpret_cast<uintptr_t>(P.c) % 32 == 0 && (P.ldc * elem) % 32 == 0) ? 1 : 0;
    P.bias_vec = (P.bias && reinterpret_cast<uintptr_t>(P.bias) % 16 == 0) ? 1 : 0;
    // fp32 output through tensor stores when the matrix allows a tensor map (16-byte aligned base and row pitch)
    static const int tma_store = []() { const char* e = getenv("DE_GEMM_TMA_STORE"); return (e && *e) ? (atoi(e) != 0) : 1; }();
    CUtensorMap map_c = map_a;                             // placeholder when unused
    P.tma_store = 0;
    if (epi == 0 && tma_store && !P.residual && reinterpret_cast<uintptr_t>(P.c) % 16 == 0 && (P.ldc * 4) % 16 == 0 &&
        make_map_f32_store(&map_c, P.c, P.M, P.N, P.ldc))
        P.tma_store = 1;
    if (mc == 2)
        return transposed ? launch_gemm_e<1, 1, 2>(map_a, map_b, map_c, P, epi, grid, s) : launch_gemm_e<0, 0, 2>(map_a, map_b, map_c, P, epi, grid, s);
    return transposed ? launch_gemm_e<1, 1, 1>(map_a, map_b, map_c, P, epi, grid, s) : launch_gemm_e<0, 0, 1>(map_a, map_b, map_c, P, epi, grid, s);
}

}  // namespace de
