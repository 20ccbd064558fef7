// grid4d_field_kernels.cu -- dense-field evaluation (sm_100a): the encoder over a REGULAR lat/lon grid at fixed altitude
// and time, inference only.  Replaces the query loop + preprocessing + encoder forward of
// /root/reference/SPECIFICATIONS.MD:1963-2088 (`visualize_embedding_field`: np.linspace lats x np.linspace lons,
// lat-major point order) and core/plant_data_loader.py:543-549 (species distribution maps).
//
// What is different from the general forward (grid4d_pair_kernels.cu):
//   * no coordinate tensor: a CTA generates the geodetic points of its 16 x 16 patch of the grid, runs the prologue
//     arithmetic (geo_math.cuh, the same single-rounding fp64 sequence as geo2xyz.cu) and keeps the 256 normalised
//     (x, y, z, t) rows in shared memory;
//   * a warp holds 16 NEIGHBOURING points of ONE level (the general kernel: 16 levels of one point), so at the coarse
//     and middle levels -- where neighbouring grid points fall into the same or adjacent cells -- the lanes of a gather
//     instruction land in the same 128 B lines and the L1 pipeline is charged one wavefront instead of sixteen
//     (ncu r02e: 233 M load wavefronts per 2^22 points against 1208 M for the general kernel);
//   * feature rows are assembled in shared memory and leave the CTA as whole G*F*4-byte pieces.
// Per-corner index, weight and summation order are those of the pair kernel, so the features are bit-identical to
// de_grid4d_forward over the same points (tests/test_field.py).
// Compiled with -fmad=false (build.py), like geo2xyz.cu; the interpolation arithmetic uses explicit intrinsics.
#include "geo_math.cuh"
#include "grid4d_common.cuh"
#include "grid4d_pair.cuh"
#include "launch.h"

namespace de {

namespace {

constexpr int kFieldTile = 16;                        // patch edge: 16 x 16 points per CTA
constexpr int kFieldPoints = kFieldTile * kFieldTile;

// np.linspace(start, stop, n)[i]: start + i * step with step = (stop - start) / (n - 1) -- a multiplication and an
// addition, each rounded (numpy evaluates `arange * step + start` elementwise) -- and the last sample set to `stop`.
__device__ __forceinline__ double linspace_at(double start, double stop, double step, int n, int i) {
    if (n > 1 && i == n - 1) return stop;
    return __dadd_rn(__dmul_rn(static_cast<double>(i), step), start);
}

__device__ __forceinline__ float4 field_point(const FieldParams& Fp, int gi, int gj) {
    const double lat = linspace_at(Fp.lat_start, Fp.lat_stop, Fp.lat_step, Fp.n_lat, gi);
    const double lon = linspace_at(Fp.lon_start, Fp.lon_stop, Fp.lon_step, Fp.n_lon, gj);
    double xyz[3];
    ecef_from_geodetic(lat, lon, Fp.altitude, xyz);
    float4 o;
    o.x = norm_coord(xyz[0], Fp.bbox_min[0], Fp.bbox_span[0]);
    o.y = norm_coord(xyz[1], Fp.bbox_min[1], Fp.bbox_span[1]);
    o.z = norm_coord(xyz[2], Fp.bbox_min[2], Fp.bbox_span[2]);
    o.w = Fp.has_time ? norm_coord(Fp.time, Fp.t_min, Fp.t_span) : 0.0f;
    return o;
}

}  // namespace

// coordinates only: xyzt[(gi * n_lon + gj)] for the whole grid (Mode R callers, tests)
__global__ void __launch_bounds__(256) field_xyzt_kernel(const FieldParams Fp, float* __restrict__ xyzt) {
    const long long n = static_cast<long long>(Fp.n_lat) * Fp.n_lon;
    const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int gi = static_cast<int>(p / Fp.n_lon), gj = static_cast<int>(p - static_cast<long long>(gi) * Fp.n_lon);
    reinterpret_cast<float4*>(xyzt)[p] = field_point(Fp, gi, gj);
}

// one (point, level) of the lane pair: cell, the lane's four corner indices, and -- after the gathers -- the feature sum
template <int F, bool POW2>
struct FieldLevel {
    float w[3];
    uint32_t idx[4];
    // c: the point's (x, y, z, t) row; the level's three columns are picked with selects on loop-invariant predicates
    // (pick4's nested conditionals compiled to branches: ~50 of 153 instructions per lane and level, ncu r02e; picking by
    // shared-memory address instead put the load latency on the index arithmetic's critical path, r02g)
    __device__ __forceinline__ void prepare(const LevelParams& lv, const float4& c, int d0, int d1, int d2, int ox) {
        const int dd[3] = {d0, d1, d2};
        float x[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            float r = c.x;
            r = dd[d] == 1 ? c.y : r;
            r = dd[d] == 2 ? c.z : r;
            r = dd[d] == 3 ? c.w : r;
            x[d] = r;
        }
        long long g[3];
        cell_of<3>(lv, x, g, w);
        CellTerms<3, POW2> ct;
        ct.set(lv, g, ox);
#pragma unroll
        for (int j = 0; j < 4; ++j) idx[j] = ct.index(lv, j);
    }
    __device__ __forceinline__ void finish(const float (&v)[4][F], int ox, unsigned pair_mask, float out_scale, float* dst) const {
        float acc[F];
#pragma unroll
        for (int f = 0; f < F; ++f) acc[f] = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float wk = pair_weight<3>(w, ox, j);
#pragma unroll
            for (int f = 0; f < F; ++f) acc[f] = __fmaf_rn(wk, v[j][f], acc[f]);
        }
#pragma unroll
        for (int f = 0; f < F; ++f) {
            acc[f] = __fadd_rn(acc[f], __shfl_xor_sync(pair_mask, acc[f], 1));
            if (out_scale != 1.0f) acc[f] = __fmul_rn(acc[f], out_scale);
        }
        if constexpr (F == 1) {
            if (ox == 0) dst[0] = acc[0];
        } else {
#pragma unroll
            for (int f = 0; f < F / 2; ++f) dst[ox * (F / 2) + f] = ox ? acc[F / 2 + f] : acc[f];
        }
    }
};

template <typename TT, int F, bool POW2>
__global__ void __launch_bounds__(kThreads, 3)
grid4d_field_forward_kernel(const __grid_constant__ EncodeParams P, const __grid_constant__ FieldParams Fp) {
    __shared__ LevelParams s_lv[kMaxGroup];
    __shared__ float4 s_xyzt[kFieldPoints];
    __shared__ double s_lat[kFieldTile][2], s_lon[kFieldTile][2];
    extern __shared__ float s_out[];                               // [256 points][G * F + F]: rows padded against bank conflicts
    const int G = P.group_size;
    const int group = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - group * P.n_tiles;
    const int first = group * G;
    const int n_lv = min(G, P.n_levels - first);
    if (threadIdx.x < n_lv) s_lv[threadIdx.x] = P.levels[first + threadIdx.x];

    const int tiles_lon = (Fp.n_lon + kFieldTile - 1) / kFieldTile;
    const int ti = tile / tiles_lon, tj = tile - ti * tiles_lon;
    // the trigonometry depends on the row (latitude) or the column (longitude) only: 16 + 16 evaluations per patch
    if (threadIdx.x < kFieldTile) {
        const int gi = min(ti * kFieldTile + static_cast<int>(threadIdx.x), Fp.n_lat - 1);
        const LatTerms t = geodetic_lat_terms(linspace_at(Fp.lat_start, Fp.lat_stop, Fp.lat_step, Fp.n_lat, gi), Fp.altitude);
        s_lat[threadIdx.x][0] = t.a1;
        s_lat[threadIdx.x][1] = t.z;
    } else if (threadIdx.x >= 32 && threadIdx.x < 32 + kFieldTile) {
        const int j = threadIdx.x - 32;
        const int gj = min(tj * kFieldTile + j, Fp.n_lon - 1);
        const LonTerms t = geodetic_lon_terms(linspace_at(Fp.lon_start, Fp.lon_stop, Fp.lon_step, Fp.n_lon, gj));
        s_lon[j][0] = t.co;
        s_lon[j][1] = t.so;
    }
    __syncthreads();
    {
        const int li = threadIdx.x >> 4, lj = threadIdx.x & 15;
        const int gi = ti * kFieldTile + li, gj = tj * kFieldTile + lj;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gi < Fp.n_lat && gj < Fp.n_lon) {
            const double a1 = s_lat[li][0];
            c.x = norm_coord(a1 * s_lon[lj][0], Fp.bbox_min[0], Fp.bbox_span[0]);      // ecef_from_geodetic, factored
            c.y = norm_coord(a1 * s_lon[lj][1], Fp.bbox_min[1], Fp.bbox_span[1]);
            c.z = norm_coord(s_lat[li][1], Fp.bbox_min[2], Fp.bbox_span[2]);
            c.w = Fp.has_time ? norm_coord(Fp.time, Fp.t_min, Fp.t_span) : 0.0f;
            if (Fp.xyzt_out && group == 0) reinterpret_cast<float4*>(Fp.xyzt_out)[static_cast<long long>(gi) * Fp.n_lon + gj] = c;
        }
        s_xyzt[threadIdx.x] = c;
    }
    __syncthreads();

    const int ox = threadIdx.x & 1;
    const unsigned pair_mask = 3u << (threadIdx.x & 30);
    const L2Policies pol = make_l2_policies(P.l2_hints != 0);
    const int row_floats = G * F + F;
    // A warp owns LEVELS (warp, warp + 8, ...) and walks the patch's 256 points for each: the level's parameters stay in
    // registers (the first version re-read them from shared memory for every point: 153 instructions per lane and level,
    // ncu r02e), its 32 lanes are 16 neighbouring points x the x-pair, two such point rows per trip (8 gathers in flight).
    const int warp = threadIdx.x >> 5, lane_pt = (threadIdx.x & 31) >> 1;
#pragma unroll 1
    for (int l = warp; l < n_lv; l += kThreads / 32) {
        const LevelParams lv = s_lv[l];
        const int d0 = lv.dim[0], d1 = lv.dim[1], d2 = lv.dim[2];
#pragma unroll 1
        for (int pt0 = 0; pt0 < kFieldPoints; pt0 += 32) {
            const int pa = pt0 + lane_pt, pb = pt0 + 16 + lane_pt;
            FieldLevel<F, POW2> A, B;
            A.prepare(lv, s_xyzt[pa], d0, d1, d2, ox);
            B.prepare(lv, s_xyzt[pb], d0, d1, d2, ox);
            float va[4][F], vb[4][F];
#pragma unroll
            for (int j = 0; j < 4; ++j) load_entry_hint<TT, F>(lv.table, A.idx[j], va[j], pol.keep);
#pragma unroll
            for (int j = 0; j < 4; ++j) load_entry_hint<TT, F>(lv.table, B.idx[j], vb[j], pol.keep);
            A.finish(va, ox, pair_mask, P.out_scale, s_out + pa * row_floats + l * F);
            B.finish(vb, ox, pair_mask, P.out_scale, s_out + pb * row_floats + l * F);
        }
    }
    __syncthreads();
    // rows leave the CTA one point per warp trip: consecutive lanes write consecutive floats of the point's slice
    const int n_out = n_lv * F;
    const int lane = threadIdx.x & 31;
    for (int c0 = 0; c0 < n_out; c0 += 32) {
        const int c = c0 + lane;
        const int l = c / F;
        const unsigned long long col = c < n_out ? s_lv[l].out_col + static_cast<unsigned>(c - l * F) : 0ull;
        for (int pt = warp; pt < kFieldPoints; pt += kThreads / 32) {
            const int gi = ti * kFieldTile + (pt >> 4), gj = tj * kFieldTile + (pt & 15);
            if (c < n_out && gi < Fp.n_lat && gj < Fp.n_lon)
                __stcs(P.out + (static_cast<long long>(gi) * Fp.n_lon + gj) * P.out_stride + col, s_out[pt * row_floats + c]);
        }
    }
}

template <typename TT, bool POW2>
static cudaError_t field_f(const EncodeParams& P, const FieldParams& Fp, int F, unsigned blocks, size_t smem, cudaStream_t s) {
    switch (F) {
        case 1: grid4d_field_forward_kernel<TT, 1, POW2><<<blocks, kThreads, smem, s>>>(P, Fp); break;
        case 2: grid4d_field_forward_kernel<TT, 2, POW2><<<blocks, kThreads, smem, s>>>(P, Fp); break;
        case 4: grid4d_field_forward_kernel<TT, 4, POW2><<<blocks, kThreads, smem, s>>>(P, Fp); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

int field_max_group(int F) { return 32 / F; }           // 256 points x (G x F + F) floats <= 36 KB of shared memory

cudaError_t launch_field_forward(EncodeParams P, const FieldParams& Fp, bool half_tables, int F, bool pow2, cudaStream_t s) {
    const long long tiles = static_cast<long long>((Fp.n_lat + kFieldTile - 1) / kFieldTile) * ((Fp.n_lon + kFieldTile - 1) / kFieldTile);
    if (tiles * P.n_groups >= (1ll << 31)) return cudaErrorInvalidValue;
    P.n_tiles = static_cast<int>(tiles);
    const unsigned blocks = static_cast<unsigned>(tiles * P.n_groups);
    const size_t smem = static_cast<size_t>(kFieldPoints) * (P.group_size * F + F) * sizeof(float);
    if (half_tables) return pow2 ? field_f<__half, true>(P, Fp, F, blocks, smem, s) : field_f<__half, false>(P, Fp, F, blocks, smem, s);
    return pow2 ? field_f<float, true>(P, Fp, F, blocks, smem, s) : field_f<float, false>(P, Fp, F, blocks, smem, s);
}

cudaError_t launch_field_xyzt(const FieldParams& Fp, float* xyzt, cudaStream_t s) {
    const long long n = static_cast<long long>(Fp.n_lat) * Fp.n_lon;
    if (n == 0) return cudaSuccess;
    field_xyzt_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(Fp, xyzt);
    return cudaGetLastError();
}

}  // namespace de
