// geo2xyz.cu -- fused geodetic prologue (sm_100a): WGS-84 geodetic -> ECEF (fp64) -> bounding-box
// normalisation -> packed fp32 (x,y,z,t) rows, one pass, one thread per point.
// Replaces ~20 ATen launches of GeospatialConverter.geodetic_to_xyz + xyz_to_norm
// (/root/reference/geospatial/geo2xyz.py:146-160, :327-332; utils.py:54-77).
// Compiled with -fmad=false so every fp64 operation rounds once, in the reference's order.
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "geo_math.cuh"
#include "launch.h"

namespace de {

struct GeoParams {
    const void* geo;
    const void* t;
    float* xyzt;
    double* ecef;
    double* minmax;
    long long n;
    double bbox_min[3];
    double bbox_span[3];
    double t_min, t_span;
    int geo_f64, t_f64;
};

__device__ __forceinline__ void ecef_of(const GeoParams& P, long long p, double (&xyz)[3]) {
    double lat, lon, alt;
    if (P.geo_f64) {
        const double* g = static_cast<const double*>(P.geo) + 3 * p;
        lat = __ldg(g); lon = __ldg(g + 1); alt = __ldg(g + 2);
    } else {
        const float* g = static_cast<const float*>(P.geo) + 3 * p;
        lat = __ldg(g); lon = __ldg(g + 1); alt = __ldg(g + 2);   // _as_fp64 (:147)
    }
    ecef_from_geodetic(lat, lon, alt, xyz);
}

__global__ void __launch_bounds__(256) geo2xyzt_kernel(const __grid_constant__ GeoParams P) {
    const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= P.n) return;
    double xyz[3];
    ecef_of(P, p, xyz);
    float4 o;
    float* of = reinterpret_cast<float*>(&o);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        if (P.ecef) P.ecef[3 * p + d] = xyz[d];
        of[d] = norm_coord(xyz[d], P.bbox_min[d], P.bbox_span[d]);
    }
    double tv = 0.0;
    if (P.t) {
        const double tt = P.t_f64 ? __ldg(static_cast<const double*>(P.t) + p)
                                  : static_cast<double>(__ldg(static_cast<const float*>(P.t) + p));
        o.w = norm_coord(tt, P.t_min, P.t_span);                         // SPECIFICATIONS.MD:765
    } else {
        o.w = static_cast<float>(tv);
    }
    reinterpret_cast<float4*>(P.xyzt)[p] = o;
}

__device__ __forceinline__ void atomic_min_f64(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (v < __longlong_as_double(old)) {
        const unsigned long long seen = atomicCAS(a, old, __double_as_longlong(v));
        if (seen == old) break;
        old = seen;
    }
}
__device__ __forceinline__ void atomic_max_f64(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (v > __longlong_as_double(old)) {
        const unsigned long long seen = atomicCAS(a, old, __double_as_longlong(v));
        if (seen == old) break;
        old = seen;
    }
}

// BoundingBox.from_points (data_structures.py:95-100) folded into a running box (geo2xyz.py:305-320).
__global__ void __launch_bounds__(256) geo_bbox_kernel(const __grid_constant__ GeoParams P) {
    double lo[3] = {CUDART_INF, CUDART_INF, CUDART_INF};
    double hi[3] = {-CUDART_INF, -CUDART_INF, -CUDART_INF};
    for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < P.n;
         p += static_cast<long long>(gridDim.x) * blockDim.x) {
        double xyz[3];
        ecef_of(P, p, xyz);
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            lo[d] = fmin(lo[d], xyz[d]);
            hi[d] = fmax(hi[d], xyz[d]);
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            lo[d] = fmin(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], off));
            hi[d] = fmax(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], off));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            atomic_min_f64(P.minmax + d, lo[d]);
            atomic_max_f64(P.minmax + 3 + d, hi[d]);
        }
    }
}

// ---- timestamps from date columns (SPECIFICATIONS.MD:740-768, `_process_temporal`) ----------------------------------
// datetime(year, month, day, hour, minute, int(second)).timestamp() with the microseconds int((second % 1) * 1e6) added,
// evaluated for UTC: whole seconds since 1970-01-01 by the proleptic-Gregorian day count (exact integers), then
// `seconds + microseconds / 1e6` in fp64 as CPython's naive-datetime path does; t_norm = (ts - t_min) / (t_max - t_min).
struct TimeParams {
    const void* comp;         // (n, n_cols) columns year[, month, day, hour, minute, second]
    double* ts_out;           // (n,) fp64 epoch seconds, or NULL
    float* tn_out;            // (n,) fp32 normalised time, or NULL
    long long n;
    long long stride;
    int n_cols, is_f64;
    int def_month, def_day, def_hour;
    double t_min, t_span;
};

__global__ void __launch_bounds__(256) time_components_kernel(const TimeParams P) {
    const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= P.n) return;
    double c[6] = {1970.0, static_cast<double>(P.def_month), static_cast<double>(P.def_day), static_cast<double>(P.def_hour), 0.0, 0.0};
    for (int k = 0; k < P.n_cols && k < 6; ++k)
        c[k] = P.is_f64 ? __ldg(static_cast<const double*>(P.comp) + p * P.stride + k)
                        : static_cast<double>(__ldg(static_cast<const int*>(P.comp) + p * P.stride + k));
    long long y = static_cast<long long>(c[0]);                   // int(row['year']) ...
    const long long m = static_cast<long long>(c[1]), d = static_cast<long long>(c[2]);
    const long long hh = static_cast<long long>(c[3]), mi = static_cast<long long>(c[4]), ss = static_cast<long long>(c[5]);
    y -= (m <= 2);
    const long long era = (y >= 0 ? y : y - 399) / 400;
    const long long yoe = y - era * 400;
    const long long doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const long long doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    const long long days = era * 146097 + doe - 719468;
    const long long secs = days * 86400 + hh * 3600 + mi * 60 + ss;
    long long us = 0;
    if (P.is_f64 && P.n_cols >= 6) us = static_cast<long long>(fmod(c[5], 1.0) * 1e6);    // int((second % 1) * 1e6)
    const double ts = static_cast<double>(secs) + static_cast<double>(us) / 1e6;
    if (P.ts_out) P.ts_out[p] = ts;
    if (P.tn_out) P.tn_out[p] = static_cast<float>((ts - P.t_min) / P.t_span);             // SPECIFICATIONS.MD:765
}

cudaError_t launch_time_components(const void* comp, bool is_f64, long long n, long long stride, int n_cols, int def_month,
                                   int def_day, int def_hour, double t_min, double t_span, double* ts_out, float* tn_out,
                                   cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    TimeParams P{comp, ts_out, tn_out, n, stride, n_cols, is_f64 ? 1 : 0, def_month, def_day, def_hour, t_min, t_span};
    time_components_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_geo2xyzt(const void* geo, bool geo_f64, const void* t, bool t_f64, const double* bbox_min,
                            const double* bbox_span, const double* t_range, float* xyzt, double* ecef,
                            long long n, cudaStream_t s) {
    GeoParams P{};
    P.geo = geo; P.t = t; P.xyzt = xyzt; P.ecef = ecef; P.minmax = nullptr; P.n = n;
    for (int d = 0; d < 3; ++d) { P.bbox_min[d] = bbox_min[d]; P.bbox_span[d] = bbox_span[d]; }
    P.t_min = t_range ? t_range[0] : 0.0;
    P.t_span = t_range ? t_range[1] - t_range[0] : 1.0;
    P.geo_f64 = geo_f64; P.t_f64 = t_f64;
    if (n == 0) return cudaSuccess;
    const unsigned blocks = static_cast<unsigned>((n + 255) / 256);
    geo2xyzt_kernel<<<blocks, 256, 0, s>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_geo_bbox(const void* geo, bool geo_f64, double* minmax, long long n, cudaStream_t s) {
    GeoParams P{};
    P.geo = geo; P.minmax = minmax; P.n = n; P.geo_f64 = geo_f64;
    if (n == 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    geo_bbox_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(P);
    return cudaGetLastError();
}

}  // namespace de
