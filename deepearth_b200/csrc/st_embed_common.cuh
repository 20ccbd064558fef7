// st_embed_common.cuh -- nn.GELU() (erf form) and its derivative in fp32, shared by st_embed.cu and st_embed_tc.cu.
#pragma once

namespace de {

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}

}  // namespace de
