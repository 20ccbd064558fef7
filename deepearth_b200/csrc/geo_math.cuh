// geo_math.cuh -- the fp64 arithmetic of the geodetic prologue, shared by geo2xyz.cu and grid4d_field_kernels.cu.
// Both translation units are compiled with -fmad=false: every operation rounds once, in the reference's order
// (/root/reference/geospatial/geo2xyz.py:146-160, :327-332; utils.py:54-77), so the two kernels agree bit for bit.
#pragma once
#include <cuda_runtime.h>

namespace de {

// The latitude- and longitude-only factors of the ECEF position.  ecef_from_geodetic() is their product; the dense-field
// kernel evaluates them once per grid row / column (16 + 16 per 256-point patch) -- same operations, same order, same bits.
struct LatTerms { double a1, z; };      // a1 = (N + alt) cos(lat),  z = (N (1 - e^2) + alt) sin(lat)
struct LonTerms { double co, so; };
__device__ __forceinline__ LatTerms geodetic_lat_terms(double lat, double alt) {
    constexpr double A = 6378137.0;                       // geo2xyz.py:98
    constexpr double FL = 1.0 / 298.257223563;            // :99
    constexpr double E2 = 2.0 * FL - FL * FL;             // :100
    constexpr double D2R = 0.017453292519943295769236907684886;  // torch.deg2rad
    double sl, cl;
    sincos(lat * D2R, &sl, &cl);
    const double N = A / sqrt(1.0 - E2 * (sl * sl));      // :153
    LatTerms t;
    t.a1 = (N + alt) * cl;                                // :157-158, first factor
    t.z = (N * (1.0 - E2) + alt) * sl;                    // :159
    return t;
}
__device__ __forceinline__ LonTerms geodetic_lon_terms(double lon) {
    constexpr double D2R = 0.017453292519943295769236907684886;
    LonTerms t;
    sincos(lon * D2R, &t.so, &t.co);
    return t;
}
__device__ __forceinline__ void ecef_from_geodetic(double lat, double lon, double alt, double (&xyz)[3]) {
    const LatTerms la = geodetic_lat_terms(lat, alt);
    const LonTerms lo = geodetic_lon_terms(lon);
    xyz[0] = la.a1 * lo.co;                               // :157
    xyz[1] = la.a1 * lo.so;                               // :158
    xyz[2] = la.z;                                        // :159
}

// (v - lo) / span in fp64, rounded once to fp32; _safe_div: |span| < 1e-9 -> 0.5 (utils.py:54-77)
__device__ __forceinline__ float norm_coord(double v, double lo, double span) {
    return static_cast<float>(fabs(span) < 1e-9 ? 0.5 : (v - lo) / span);
}

}  // namespace de
