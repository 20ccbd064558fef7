// geo_math.cuh -- the fp64 arithmetic of the geodetic prologue, shared by geo2xyz.cu and grid4d_field_kernels.cu.
// Both translation units are compiled with -fmad=false: every operation rounds once, in the reference's order
// (/root/reference/geospatial/geo2xyz.py:146-160, :327-332; utils.py:54-77), so the two kernels agree bit for bit.
#pragma once
#include <cuda_runtime.h>

namespace de {

__device__ __forceinline__ void ecef_from_geodetic(double lat, double lon, double alt, double (&xyz)[3]) {
    constexpr double A = 6378137.0;                       // geo2xyz.py:98
    constexpr double FL = 1.0 / 298.257223563;            // :99
    constexpr double E2 = 2.0 * FL - FL * FL;             // :100
    constexpr double D2R = 0.017453292519943295769236907684886;  // torch.deg2rad
    double sl, cl, so, co;
    sincos(lat * D2R, &sl, &cl);
    sincos(lon * D2R, &so, &co);
    const double N = A / sqrt(1.0 - E2 * (sl * sl));      // :153
    xyz[0] = ((N + alt) * cl) * co;                       // :157
    xyz[1] = ((N + alt) * cl) * so;                       // :158
    xyz[2] = (N * (1.0 - E2) + alt) * sl;                 // :159
}

// (v - lo) / span in fp64, rounded once to fp32; _safe_div: |span| < 1e-9 -> 0.5 (utils.py:54-77)
__device__ __forceinline__ float norm_coord(double v, double lo, double span) {
    return static_cast<float>(fabs(span) < 1e-9 ? 0.5 : (v - lo) / span);
}

}  // namespace de
