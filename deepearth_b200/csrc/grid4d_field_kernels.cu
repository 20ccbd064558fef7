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
//     instruction land in the same 128 B lines and the L1 pipeline is charged one wavefront instead of sixteen;
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

template <typename TT, int F, bool POW2>
__global__ void __launch_bounds__(kThreads, 4)
grid4d_field_forward_kernel(const __grid_constant__ EncodeParams P, const __grid_constant__ FieldParams Fp) {
    constexpr int D = 3, J = PairGeom<D>::kCombos;
    __shared__ LevelParams s_lv[kMaxGroup];
    __shared__ float4 s_xyzt[kFieldPoints];
    extern __shared__ float s_out[];                               // [256 points][G * F]
    const int G = P.group_size;
    const int group = blockIdx.x / P.n_tiles;
    const int tile = blockIdx.x - group * P.n_tiles;
    const int first = group * G;
    const int n_lv = min(G, P.n_levels - first);
    if (threadIdx.x < n_lv) s_lv[threadIdx.x] = P.levels[first + threadIdx.x];

    const int tiles_lon = (Fp.n_lon + kFieldTile - 1) / kFieldTile;
    const int ti = tile / tiles_lon, tj = tile - ti * tiles_lon;
    {
        const int gi = ti * kFieldTile + (threadIdx.x >> 4), gj = tj * kFieldTile + (threadIdx.x & 15);
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gi < Fp.n_lat && gj < Fp.n_lon) {
            c = field_point(Fp, gi, gj);
            if (Fp.xyzt_out && group == 0) reinterpret_cast<float4*>(Fp.xyzt_out)[static_cast<long long>(gi) * Fp.n_lon + gj] = c;
        }
        s_xyzt[threadIdx.x] = c;
    }
    __syncthreads();

    const int ox = threadIdx.x & 1;
    const unsigned pair_mask = 3u << (threadIdx.x & 30);
    const L2Policies pol = make_l2_policies(P.l2_hints != 0);
    const int row_floats = G * F;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const int pt = (threadIdx.x >> 1) + 128 * pass;           // a warp: 16 consecutive points of one patch row
        const float4 c = s_xyzt[pt];
#pragma unroll 2
        for (int l = 0; l < n_lv; ++l) {
            const LevelParams& lv = s_lv[l];
            float x[D];
#pragma unroll
            for (int d = 0; d < D; ++d) x[d] = pick4(c, lv.dim[d]);
            long long g[D];
            float w[D];
            cell_of<D>(lv, x, g, w);
            CellTerms<D, POW2> ct;
            ct.set(lv, g, ox);
            uint32_t idx[J];
#pragma unroll
            for (int j = 0; j < J; ++j) idx[j] = ct.index(lv, j);
            float v[J][F];
#pragma unroll
            for (int j = 0; j < J; ++j) load_entry_hint<TT, F>(lv.table, idx[j], v[j], pol.keep);
            float acc[F];
#pragma unroll
            for (int f = 0; f < F; ++f) acc[f] = 0.0f;
#pragma unroll
            for (int j = 0; j < J; ++j) {
                const float wk = pair_weight<D>(w, ox, j);
#pragma unroll
                for (int f = 0; f < F; ++f) acc[f] = __fmaf_rn(wk, v[j][f], acc[f]);
            }
#pragma unroll
            for (int f = 0; f < F; ++f) {
                acc[f] = __fadd_rn(acc[f], __shfl_xor_sync(pair_mask, acc[f], 1));
                if (P.out_scale != 1.0f) acc[f] = __fmul_rn(acc[f], P.out_scale);
            }
            float* dst = s_out + pt * row_floats + l * F;
            if constexpr (F == 1) {
                if (ox == 0) dst[0] = acc[0];
            } else {
#pragma unroll
                for (int f = 0; f < F / 2; ++f) dst[ox * (F / 2) + f] = ox ? acc[F / 2 + f] : acc[f];
            }
        }
    }
    __syncthreads();
    // rows leave the CTA piecewise: consecutive threads write consecutive floats of a point's slice
    const int n_out = n_lv * F;
    for (int e = threadIdx.x; e < kFieldPoints * n_out; e += kThreads) {
        const int pt = e / n_out, c = e - pt * n_out;
        const int gi = ti * kFieldTile + (pt >> 4), gj = tj * kFieldTile + (pt & 15);
        if (gi >= Fp.n_lat || gj >= Fp.n_lon) continue;
        const int l = c / F;
        const long long p = static_cast<long long>(gi) * Fp.n_lon + gj;
        __stcs(P.out + p * P.out_stride + s_lv[l].out_col + (c - l * F), s_out[pt * row_floats + c]);
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

int field_max_group(int F) { return 32 / F; }           // 256 points x G x F floats = 32 KB of shared memory

cudaError_t launch_field_forward(EncodeParams P, const FieldParams& Fp, bool half_tables, int F, bool pow2, cudaStream_t s) {
    const long long tiles = static_cast<long long>((Fp.n_lat + kFieldTile - 1) / kFieldTile) * ((Fp.n_lon + kFieldTile - 1) / kFieldTile);
    if (tiles * P.n_groups >= (1ll << 31)) return cudaErrorInvalidValue;
    P.n_tiles = static_cast<int>(tiles);
    const unsigned blocks = static_cast<unsigned>(tiles * P.n_groups);
    const size_t smem = static_cast<size_t>(kFieldPoints) * P.group_size * F * sizeof(float);
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
