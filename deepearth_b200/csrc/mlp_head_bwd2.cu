// mlp_head_bwd2.cu -- backward of the 64-wide projection head, warp-specialised ping-pong version (sm_100a).
//
//   forward:  h = x W1^T + b1 ;  z = LN(h) (optional) ;  a = act(z) ;  y = a W2^T + b2
//   (structure of Grid4DEncoder.projection, /root/reference/models/encoders.py:47-52, at the 64-wide size of
//   BASELINE.json configs[4]); given gy = dL/dy this kernel produces dL/dx and accumulates dW1, db1, dgamma,
//   dbeta, dW2, db2.  Four GEMMs per 128-point tile:
//
//   GEMM 1  D_h  [128 x 64]  = X  [128 x K1] * W1^T           GEMM 2  D_da [128 x 64] = GY [128 x 64] * W2
//   GEMM 3  D_dx [128 x K1]  = DH [128 x 64] * W1
//   GEMM 4  D_dW [128 x N4] += [GY | DH]^T [128 cols x 128 pts] * [A | X | 1] [128 pts x N4]   (dW2, dW1, db2, db1)
//
// What changed against the first version (16.8 ms per 2^24 points, one 4-warp CTA per SM, every phase serial; removed):
//   * GEMM 4 contracts over POINTS.  Its operands are the point-major tiles the other GEMMs already use, read
//     as MN-major operands (instruction-descriptor bits 15/16): no transposed copies, which were ~400 two-byte
//     shared stores per thread per tile and 84 KB of shared memory.
//   * The freed shared memory holds TWO tiles in flight: two 128-thread consumer groups ping-pong (one stages /
//     runs its register epilogue while the other waits on the tensor core), a ninth warp issues every
//     tcgen05.mma in arrival order (so the shared dW accumulator in TMEM sees one in-order issuer), and each
//     group prefetches its next tile's global loads into registers before it waits for GEMM 3/4.
//   * TMEM: group g owns columns [128 g, 128 g + 128): D_h | D_da, then D_dx on top of them; dW at 256..256+N4.
#include <cstdlib>

#include "launch.h"
#include "mlp_common.cuh"

namespace de {

using namespace mlp;

namespace {

constexpr int kPPThreads = 384;          // 2 consumer warpgroups + 1 warpgroup whose first warp issues the MMAs
                                         // (registers are allocated per warpgroup; setmaxnreg moves them to the consumers)
constexpr int kColDw2 = 256;

__device__ __forceinline__ void tmem_ld16b(uint32_t taddr, float (&v)[16]) {
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
__device__ __forceinline__ float colsum32(float (&vals)[32], int lane) {
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

__device__ __forceinline__ void red_add_f32(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// A 128 x KT fp32 tile held in registers between its global loads and its shared-memory stores.
template <int KT>
struct TileR {
    static constexpr int kIt = (4 * KT) / 32;
    float4 a[kIt], b[kIt];
};

// Issue every global load of rows [p0, p0+128) (64 B x KT/8 per thread in flight); rows >= n read as zero.
// Warp w (0..3 inside its group) covers tile rows 32w .. 32w+31; 8 consecutive lanes take 8 consecutive rows of
// one 16-byte bf16 chunk column (conflict-free shared stores), 4 chunk columns per request (128 B segments).
template <int KT>
__device__ __forceinline__ void tile_issue2(TileR<KT>& t, const float* tile_base, int stride, int n_valid, int warp, int lane) {
    const bool a32 = ((reinterpret_cast<uintptr_t>(tile_base) | (static_cast<uintptr_t>(stride) << 2)) & 31) == 0;
#pragma unroll
    for (int it = 0; it < TileR<KT>::kIt; ++it) {
        const int c = lane + 32 * it;
        const int r = warp * 32 + (c / KT) * 8 + ((c % KT) & 7);
        const int k8 = (c % KT) >> 3;
        t.a[it] = t.b[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < n_valid) ld_stream_32B(tile_base + (r * stride + k8 * 8), t.a[it], t.b[it], a32);
    }
}

// Same, without predicates or zero fill: rows past the end of the batch re-read the last valid row.  Only for the
// FEATURE tile: a padded row's gradient row is zero, so whatever features it holds drop out of every product.
template <int KT>
__device__ __forceinline__ void tile_issue2_clamped(TileR<KT>& t, const float* tile_base, int stride, int n_valid, int warp, int lane) {
    const bool a32 = ((reinterpret_cast<uintptr_t>(tile_base) | (static_cast<uintptr_t>(stride) << 2)) & 31) == 0;
#pragma unroll
    for (int it = 0; it < TileR<KT>::kIt; ++it) {
        const int c = lane + 32 * it;
        const int r = min(warp * 32 + (c / KT) * 8 + ((c % KT) & 7), n_valid - 1);
        const int k8 = (c % KT) >> 3;
        ld_stream_32B(tile_base + (r * stride + k8 * 8), t.a[it], t.b[it], a32);
    }
}

// registers -> bf16 chunks [chunk0, chunk0 + KT/8) of a canonical K-major tile that is `cols` columns wide
template <int KT>
__device__ __forceinline__ void tile_store2(const TileR<KT>& t, unsigned char* dst, int chunk0, int cols, int warp, int lane) {
#pragma unroll
    for (int it = 0; it < TileR<KT>::kIt; ++it) {
        const int c = lane + 32 * it;
        const int r = warp * 32 + (c / KT) * 8 + ((c % KT) & 7);
        const int k8 = (c % KT) >> 3;
        const float4 a = t.a[it], b = t.b[it];
        *reinterpret_cast<uint4*>(dst + canon_chunk_offset(r, chunk0 + k8, cols)) =
            make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
    }
}

// Epilogue B: one row of D_dx (TMEM) -> dL/dx.  Row-major rows, or the group-major planes the hash-grid backward walks
// (dx[c / W][p][c % W]: consecutive threads = consecutive points write consecutive 16 B pieces of one plane).
template <int K1>
__device__ __forceinline__ void store_dx_row(const MlpHeadBwdParams& P, uint32_t tmem_row, long long pr) {
    const bool ok = pr < P.n;
    const int W = P.dx_group_w ? P.dx_group_w : (1 << 30);          // row-major = one "plane" as wide as the row
    const long long plane = P.dx_group_w ? P.n * P.dx_group_w : 0;  // floats between planes
    float* pl = P.dx + (P.dx_group_w ? pr * P.dx_group_w : pr * P.dx_stride);
    const bool a32 = ((reinterpret_cast<uintptr_t>(P.dx) | (static_cast<uintptr_t>(P.dx_group_w ? P.dx_group_w : P.dx_stride) << 2)) & 31) == 0;
    int within = 0;
#pragma unroll
    for (int c0 = 0; c0 < K1; c0 += 16) {
        float q[16];
        tmem_ld16b(tmem_row + c0, q);                               // warp-collective: outside the row predicate
        if (W >= 8) {                                               // 32 B pieces stay inside one plane (or one row)
#pragma unroll
            for (int u = 0; u < 4; u += 2) {
                if (ok) st_stream_32B(pl + within, make_float4(q[4 * u], q[4 * u + 1], q[4 * u + 2], q[4 * u + 3]),
                                      make_float4(q[4 * u + 4], q[4 * u + 5], q[4 * u + 6], q[4 * u + 7]), a32);
                within += 8;
                if (within == W) { within = 0; pl += plane; }
            }
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (ok) __stcs(reinterpret_cast<float4*>(pl + within), make_float4(q[4 * u], q[4 * u + 1], q[4 * u + 2], q[4 * u + 3]));
                within += 4;
                if (within == W) { within = 0; pl += plane; }
            }
        }
    }
}

}  // namespace

// dynamic shared memory (bf16, canonical K-major, no swizzle):
//   W1 [64 x K1] | W1T [K1 x 64] | W2T [64 x 64] | per group g: XA_g [128 x NB] = [A | X | ones(16)] , GD_g [128 x 128] = [GY | DH]
template <int K1, int SPLIT>
__global__ void __launch_bounds__(kPPThreads, 1)
mlp_head_bwd_pp_kernel(const __grid_constant__ MlpHeadBwdParams P) {
    constexpr int NB = kH + K1 + 16;                  // columns of XA = N of GEMM 4
    constexpr uint32_t kXaBytes = kTileM * NB * 2, kGdBytes = kTileM * 128 * 2;
    constexpr uint32_t kXaGroup = (NB / 8) * 128;     // bytes between 8-point groups of XA
    constexpr uint32_t kGdGroup = (128 / 8) * 128;    // ... of GD
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(16) float s_vec[3][kH];      // b1, ln gamma, ln beta (read as packed pairs)
    __shared__ __align__(8) uint64_t s_in_full[2], s_d12[2], s_mid_full[2], s_d34[2];
    __shared__ uint32_t s_tmem;

    unsigned char* sW1 = smem;
    unsigned char* sW1T = sW1 + kH * K1 * 2;
    unsigned char* sW2T = sW1T + K1 * kH * 2;
    unsigned char* sTiles = sW2T + kH * kH * 2;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = warp >> 2;                        // 0, 1: consumer groups; 2: the MMA warp

    // ---- one-time setup ------------------------------------------------------------------------
    const __nv_bfloat16* w1 = static_cast<const __nv_bfloat16*>(P.w1);
    const __nv_bfloat16* w1t = static_cast<const __nv_bfloat16*>(P.w1t);
    const __nv_bfloat16* w2t = static_cast<const __nv_bfloat16*>(P.w2t);
    for (int c = tid; c < kH * (K1 / 8); c += kPPThreads) {         // W1 (64, K1): B of GEMM 1
        const int n = c / (K1 / 8), k8 = c % (K1 / 8);
        *reinterpret_cast<uint4*>(sW1 + canon_chunk_offset(n, k8, K1)) = *reinterpret_cast<const uint4*>(w1 + static_cast<size_t>(n) * K1 + k8 * 8);
    }
    for (int c = tid; c < K1 * (kH / 8); c += kPPThreads) {         // W1^T (K1, 64): B of GEMM 3
        const int n = c / (kH / 8), k8 = c % (kH / 8);
        *reinterpret_cast<uint4*>(sW1T + canon_chunk_offset(n, k8, kH)) = *reinterpret_cast<const uint4*>(w1t + static_cast<size_t>(n) * kH + k8 * 8);
    }
    for (int c = tid; c < kH * (kH / 8); c += kPPThreads) {         // W2^T (64 hidden, 64 out): B of GEMM 2
        const int n = c / (kH / 8), k8 = c % (kH / 8);
        *reinterpret_cast<uint4*>(sW2T + canon_chunk_offset(n, k8, kH)) = *reinterpret_cast<const uint4*>(w2t + static_cast<size_t>(n) * kH + k8 * 8);
    }
    for (int c = tid; c < 2 * kTileM * 2; c += kPPThreads) {        // the 16 columns of ones of both XA tiles (bias gradients)
        const int g = c / (kTileM * 2), r = (c / 2) % kTileM, k8 = (kH + K1) / 8 + (c & 1);
        *reinterpret_cast<uint4*>(sTiles + g * (kXaBytes + kGdBytes) + canon_chunk_offset(r, k8, NB)) =
            make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
    }
    if (tid < kH) {
        s_vec[0][tid] = P.b1 ? P.b1[tid] : 0.0f;
        s_vec[1][tid] = P.ln_gamma ? P.ln_gamma[tid] : 1.0f;
        s_vec[2][tid] = P.ln_beta ? P.ln_beta[tid] : 0.0f;
    }
    if (tid == 0) {
        for (int g = 0; g < 2; ++g) {
            mbar_init(&s_in_full[g], 128);
            mbar_init(&s_mid_full[g], 128);
            mbar_init(&s_d12[g], 1);
            mbar_init(&s_d34[g], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    // tiles of this CTA: blockIdx.x + m * gridDim.x, m = 0, 1, ...; group g takes m = g, g + 2, ...
    const int n_mine = P.n_tiles > static_cast<int>(blockIdx.x) ? (P.n_tiles - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    const int cnt0 = (n_mine + 1) / 2, cnt1 = n_mine / 2;

    if (grp == 2) {
        // =========================== MMA warp: one thread issues every tcgen05.mma ===================
        // register budget: 384 threads x 168 at launch; this warpgroup gives back (168 - 56) x 128 = 14336 registers, exactly
        // what the two consumer warpgroups take (2 x (224 - 168) x 128) -- a larger request would wait forever.
        // SPLIT (DE_HEAD_BWD_SPLIT, default 2): 0 = 56 / 224, 1 = 88 / 208, 2 = 104 / 200, 3 = 120 / 192.  With 56 registers
        // the MMA thread spilled its descriptors and every GEMM 1 / 2 issue took ~1 us (13 % of the consumers' time was the
        // wait for it); measured per 2^24 points: 10.6 / 9.2 / 8.4 / 8.7 ms.
        if constexpr (SPLIT == 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        else if constexpr (SPLIT == 1) asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
        else if constexpr (SPLIT == 2 || SPLIT == 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 104;");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 120;");
        if (warp == 8 && lane == 0) {
            int issued12[2] = {0, 0}, issued34[2] = {0, 0};
            const int cnt[2] = {cnt0, cnt1};
            bool dw_started = false;
            constexpr uint32_t kIdesc64 = umma_idesc(kH), kIdescK1 = umma_idesc(K1);
            constexpr uint32_t kIdescDw = umma_idesc(NB) | (1u << 15) | (1u << 16);     // A and B MN-major
            while (issued34[0] < cnt[0] || issued34[1] < cnt[1]) {
                bool any = false;
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    const uint32_t xa = smem_u32(sTiles + g * (kXaBytes + kGdBytes)), gd = xa + kXaBytes;
                    const uint32_t tcol = tmem + g * 128;
                    if (issued12[g] == issued34[g]) {
                        if (issued12[g] < cnt[g] && mbar_test(&s_in_full[g], issued12[g] & 1)) {
                            tc_fence_after();
                            const uint32_t w1a = smem_u32(sW1), w2a = smem_u32(sW2T);
#pragma unroll
                            for (int k = 0; k < K1 / 16; ++k)       // GEMM 1: X (XA chunks 8..) * W1^T -> D_h
                                umma_bf16(tcol, umma_desc(xa + (kH / 8) * 128 + k * 256, 128, kXaGroup), umma_desc(w1a + k * 256, 128, (K1 / 8) * 128),
                                          kIdesc64, k > 0 ? 1u : 0u);
#pragma unroll
                            for (int k = 0; k < kH / 16; ++k)       // GEMM 2: GY (GD chunks 0..7) * W2 -> D_da
                                umma_bf16(tcol + 64, umma_desc(gd + k * 256, 128, kGdGroup), umma_desc(w2a + k * 256, 128, (kH / 8) * 128),
                                          kIdesc64, k > 0 ? 1u : 0u);
                            umma_commit(&s_d12[g]);
                            ++issued12[g];
                            any = true;
                        }
                    } else if (mbar_test(&s_mid_full[g], issued34[g] & 1)) {
                        tc_fence_after();
                        const uint32_t w1ta = smem_u32(sW1T);
#pragma unroll
                        for (int k = 0; k < kH / 16; ++k)           // GEMM 3: DH (GD chunks 8..15) * W1 -> D_dx (over D_h | D_da)
                            umma_bf16(tcol, umma_desc(gd + (kH / 8) * 128 + k * 256, 128, kGdGroup), umma_desc(w1ta + k * 256, 128, (kH / 8) * 128),
                                      kIdescK1, k > 0 ? 1u : 0u);
#pragma unroll
                        for (int k = 0; k < kTileM / 16; ++k) {     // GEMM 4: [GY | DH]^T * [A | X | 1], K = 16 points per step
                            const uint32_t aa = gd + k * 2 * kGdGroup, ba = xa + k * 2 * kXaGroup;
                            // MN-major, no swizzle: SBO = bytes between 8-column core matrices (128), LBO = bytes between
                            // 8-point groups (checked on B200 against the first, serial version of this kernel, since removed)
                            umma_bf16(tmem + kColDw2, umma_desc(aa, kGdGroup, 128), umma_desc(ba, kXaGroup, 128), kIdescDw,
                                      (dw_started || k > 0) ? 1u : 0u);
                        }
                        dw_started = true;
                        umma_commit(&s_d34[g]);
                        ++issued34[g];
                        any = true;
                    }
                }
                if constexpr (SPLIT != 4) { if (!any) __nanosleep(40); }      // SPLIT 4 (A/B): spin without backing off
            }
        }
        __syncwarp();
    } else {
        // =========================== consumer groups ================================================
        if constexpr (SPLIT == 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
        else if constexpr (SPLIT == 1) asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
        else if constexpr (SPLIT == 2 || SPLIT == 4) asm volatile("setmaxnreg.inc.sync.aligned.u32 200;");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 192;");
        const int cnt = grp == 0 ? cnt0 : cnt1;
        const int wq = warp & 3;                                    // warp inside the group = TMEM lane quarter
        const int row = wq * 32 + lane;                             // tile row (point) owned in the epilogues
        unsigned char* sXA = sTiles + grp * (kXaBytes + kGdBytes);
        unsigned char* sGD = sXA + kXaBytes;
        const uint32_t tmem_row = tmem + (static_cast<uint32_t>(wq * 32) << 16) + grp * 128;
        float acc_beta[2] = {0.f, 0.f}, acc_gamma[2] = {0.f, 0.f};  // lane l: columns l and 32 + l, this warp's rows

        long long p0 = (static_cast<long long>(blockIdx.x) + static_cast<long long>(grp) * gridDim.x) * kTileM;
        const long long p_step = 2LL * gridDim.x * kTileM;
        const int xs = static_cast<int>(P.x_stride), gs = static_cast<int>(P.gy_stride);
        for (int i = 0; i < cnt; ++i, p0 += p_step) {
            const int n_valid = static_cast<int>(P.n - p0 < kTileM ? P.n - p0 : kTileM);
            // ---- this tile's feature rows: loads issued now, in flight while the tensor core finishes GEMM 3 / 4
            //      of the previous tile and its gradient rows are streamed out ---------------------------
            TileR<K1> tx;
            TileR<kH> tg;
            tile_issue2_clamped<K1>(tx, P.x + p0 * P.x_stride, xs, n_valid, wq, lane);
            tile_issue2<kH>(tg, P.gy + p0 * P.gy_stride, gs, n_valid, wq, lane);
            if (i > 0) {                                            // epilogue B of the previous tile
                mbar_wait(&s_d34[grp], (i - 1) & 1);
                tc_fence_after();
                store_dx_row<K1>(P, tmem_row, p0 - p_step + row);
            }
            // ---- stage X and GY (both already in flight) ---------------------------------------------
            tile_store2<K1>(tx, sXA, kH / 8, NB, wq, lane);
            tile_store2<kH>(tg, sGD, 0, 128, wq, lane);
            fence_async_smem();
            tc_fence_before();
            mbar_arrive(&s_in_full[grp]);

            mbar_wait(&s_d12[grp], i & 1);
            tc_fence_after();

            // ---- epilogue A: one point per thread (LayerNorm statistics, activation and its derivative,
            //      LayerNorm backward, all in registers) -----------------------------------------------
            //      The arithmetic runs on packed fp32 pairs (mlp_common.cuh: FFMA2 / FADD2 / FMUL2, columns 2k and 2k + 1 in
            //      one issue slot; each half an ordinary IEEE operation, sums in the scalar version's order): the consumers
            //      are latency-bound at two warps per scheduler, and this epilogue was ~60 % of the kernel's instructions.
            {
                f32x2 v[kH / 2];      // h, then xhat (LayerNorm) or h (no LayerNorm)
                f32x2 d[kH / 2];      // dz = da * act'(z), then dh
                const f32x2* vb1 = reinterpret_cast<const f32x2*>(s_vec[0]);
                const f32x2* vga = reinterpret_cast<const f32x2*>(s_vec[1]);
                const f32x2* vbe = reinterpret_cast<const f32x2*>(s_vec[2]);
                {
                    float lo[32], hi[32];
                    tmem_ld32(tmem_row, lo);
                    tmem_ld32(tmem_row + 32, hi);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        v[j] = add2(pk2(lo[2 * j], lo[2 * j + 1]), vb1[j]);
                        v[16 + j] = add2(pk2(hi[2 * j], hi[2 * j + 1]), vb1[16 + j]);
                    }
                }
                float rstd = 1.0f;
                if (P.layernorm) {
                    f32x2 sa = 0ull, sb = 0ull;                         // (+0, +0); the four chains of sum64
#pragma unroll
                    for (int j = 0; j < kH / 2; j += 2) { sa = add2(sa, v[j]); sb = add2(sb, v[j + 1]); }
                    float a0, a1, a2, a3;
                    upk2(sa, a0, a1);
                    upk2(sb, a2, a3);
                    const float mean = ((a0 + a1) + (a2 + a3)) * (1.0f / kH);
                    const f32x2 nmean = splat2(-mean);
                    sa = sb = 0ull;
#pragma unroll
                    for (int j = 0; j < kH / 2; j += 2) {
                        v[j] = add2(v[j], nmean);
                        v[j + 1] = add2(v[j + 1], nmean);
                        sa = fma2(v[j], v[j], sa);
                        sb = fma2(v[j + 1], v[j + 1], sb);
                    }
                    upk2(sa, a0, a1);
                    upk2(sb, a2, a3);
                    rstd = rsqrtf(((a0 + a1) + (a2 + a3)) * (1.0f / kH) + P.ln_eps);
                    const f32x2 r2 = splat2(rstd);
#pragma unroll
                    for (int j = 0; j < kH / 2; ++j) v[j] = mul2(v[j], r2);
                }
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float da[32];
                    tmem_ld32(tmem_row + 64 + 32 * half, da);
#pragma unroll
                    for (int c8 = 0; c8 < 4; ++c8) {
                        uint32_t a8[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int i = 8 * c8 + 2 * u, jp = 16 * half + 4 * c8 + u;       // column in this half, pair index
                            const f32x2 z = P.layernorm ? fma2(v[jp], vga[jp], vbe[jp]) : v[jp];
                            f32x2 a = z, g = splat2(1.0f);
                            if (P.activation == 2) {
                                gelu_fwd_bwd2(z, a, g);
                            } else if (P.activation == 1) {
                                float z0, z1;
                                upk2(z, z0, z1);
                                a = pk2(fmaxf(z0, 0.0f), fmaxf(z1, 0.0f));
                                g = pk2(z0 > 0.0f ? 1.0f : 0.0f, z1 > 0.0f ? 1.0f : 0.0f);
                            }
                            float a0, a1;
                            upk2(a, a0, a1);
                            a8[u] = pack_bf16(a0, a1);
                            d[jp] = mul2(pk2(da[i], da[i + 1]), g);
                        }
                        // A: columns 0..63 of XA (B operand of GEMM 4)
                        *reinterpret_cast<uint4*>(sXA + canon_chunk_offset(row, 4 * half + c8, NB)) = make_uint4(a8[0], a8[1], a8[2], a8[3]);
                    }
                }
                if (P.layernorm) {
                    // dbeta_j += sum_p dz ; dgamma_j += sum_p dz * xhat  (this warp's 32 points, lane l <- column l)
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        float t[32];
#pragma unroll
                        for (int i = 0; i < 16; ++i) upk2(d[16 * half + i], t[2 * i], t[2 * i + 1]);
                        acc_beta[half] += colsum32(t, lane);
#pragma unroll
                        for (int i = 0; i < 16; ++i) upk2(mul2(d[16 * half + i], v[16 * half + i]), t[2 * i], t[2 * i + 1]);
                        acc_gamma[half] += colsum32(t, lane);
                    }
                    f32x2 sa = 0ull, sb = 0ull, da2 = 0ull, db2 = 0ull;            // sum64(d), dot64(d, v): four chains each
#pragma unroll
                    for (int j = 0; j < kH / 2; j += 2) {
                        d[j] = mul2(d[j], vga[j]);                                  // dxhat
                        d[j + 1] = mul2(d[j + 1], vga[j + 1]);
                        sa = add2(sa, d[j]);
                        sb = add2(sb, d[j + 1]);
                        da2 = fma2(d[j], v[j], da2);
                        db2 = fma2(d[j + 1], v[j + 1], db2);
                    }
                    float a0, a1, a2, a3;
                    upk2(sa, a0, a1);
                    upk2(sb, a2, a3);
                    const f32x2 nm1 = splat2(-(((a0 + a1) + (a2 + a3)) * (1.0f / kH)));
                    upk2(da2, a0, a1);
                    upk2(db2, a2, a3);
                    const f32x2 nm2 = splat2(-(((a0 + a1) + (a2 + a3)) * (1.0f / kH)));
                    const f32x2 r2 = splat2(rstd);
#pragma unroll
                    for (int j = 0; j < kH / 2; ++j) d[j] = mul2(r2, fma2(v[j], nm2, add2(d[j], nm1)));
                }
                // DH: columns 64..127 of GD (A operand of GEMM 3, rows 64..127 of GEMM 4's A)
#pragma unroll
                for (int k8 = 0; k8 < kH / 8; ++k8) {
                    uint32_t q[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        float d0, d1;
                        upk2(d[4 * k8 + u], d0, d1);
                        q[u] = pack_bf16(d0, d1);
                    }
                    *reinterpret_cast<uint4*>(sGD + canon_chunk_offset(row, kH / 8 + k8, 128)) = make_uint4(q[0], q[1], q[2], q[3]);
                }
            }
            fence_async_smem();
            tc_fence_before();
            mbar_arrive(&s_mid_full[grp]);

        }
        if (cnt > 0) {                                              // epilogue B of the last tile
            mbar_wait(&s_d34[grp], (cnt - 1) & 1);
            tc_fence_after();
            store_dx_row<K1>(P, tmem_row, p0 - p_step + row);
            if (P.layernorm) {
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    if (P.dgamma) red_add_f32(P.dgamma + 32 * half + lane, acc_gamma[half]);
                    if (P.dbeta) red_add_f32(P.dbeta + 32 * half + lane, acc_beta[half]);
                }
            }
        }
    }

    // ---- weight / bias gradients of this CTA's tiles: TMEM -> global (fp32 reductions) ---------------
    // every group has waited for its last GEMM 4 (s_d34), so after this barrier the dW accumulator is final
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (grp == 0 && n_mine > 0) {
        const int row = tid;                                        // TMEM lane = row of D_dW
        const uint32_t tmem_row = tmem + (static_cast<uint32_t>(warp * 32) << 16);
        if (row < kH) {                    // rows 0..63: output unit o = row -> dW2[o][:], db2[o]
#pragma unroll
            for (int c0 = 0; c0 < kH; c0 += 16) {
                float q[16];
                tmem_ld16b(tmem_row + kColDw2 + c0, q);
#pragma unroll
                for (int u = 0; u < 16; ++u) red_add_f32(P.dw2 + row * kH + c0 + u, q[u]);
            }
            float q[16];
            tmem_ld16b(tmem_row + kColDw2 + kH + K1, q);
            if (P.db2) red_add_f32(P.db2 + row, q[0]);
        } else {                           // rows 64..127: hidden unit j = row - 64 -> dW1[j][:], db1[j]
            const int j = row - kH;
#pragma unroll
            for (int c0 = 0; c0 < K1; c0 += 16) {
                float q[16];
                tmem_ld16b(tmem_row + kColDw2 + kH + c0, q);
#pragma unroll
                for (int u = 0; u < 16; ++u) red_add_f32(P.dw1 + j * K1 + c0 + u, q[u]);
            }
            float q[16];
            tmem_ld16b(tmem_row + kColDw2 + kH + K1, q);
            if (P.db1) red_add_f32(P.db1 + j, q[0]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
template <int K1>
static cudaError_t launch_pp_k(const MlpHeadBwdParams& P, int n_sms, cudaStream_t s) {
    // register split between the MMA warp and the consumer warpgroups (A/B knob, read once per process; 2 = measured best)
    static const int split = []() { const char* sp = getenv("DE_HEAD_BWD_SPLIT"); return sp ? atoi(sp) : 2; }();
    auto kernel = split == 0 ? mlp_head_bwd_pp_kernel<K1, 0> : split == 1 ? mlp_head_bwd_pp_kernel<K1, 1>
                : split == 3 ? mlp_head_bwd_pp_kernel<K1, 3> : split == 4 ? mlp_head_bwd_pp_kernel<K1, 4> : mlp_head_bwd_pp_kernel<K1, 2>;
    constexpr int NB = kH + K1 + 16;
    const size_t smem = static_cast<size_t>(kH) * K1 * 2 + K1 * kH * 2 + kH * kH * 2 + 2 * (static_cast<size_t>(kTileM) * NB * 2 + kTileM * 128 * 2);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int grid = n_sms;                                  // 512 TMEM columns: one CTA per SM
    if (grid > P.n_tiles) grid = P.n_tiles;
    kernel<<<grid, kPPThreads, smem, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_mlp_head_bwd(MlpHeadBwdParams P, int in_dim, int n_sms, cudaStream_t s) {
    P.n_tiles = static_cast<int>((P.n + kTileM - 1) / kTileM);
    if (P.n_tiles == 0) return cudaSuccess;
    switch (in_dim) {
        case 32: return launch_pp_k<32>(P, n_sms, s);
        case 48: return launch_pp_k<48>(P, n_sms, s);
        case 64: return launch_pp_k<64>(P, n_sms, s);
        case 96: return launch_pp_k<96>(P, n_sms, s);
        case 128: return launch_pp_k<128>(P, n_sms, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace de
