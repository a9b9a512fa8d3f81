// rbf_math.cuh -- the Gaussian RBF arithmetic shared by the stand-alone K3 kernel (dgnn_rbf.cu) and the fused epilogue
// of the fill passes (row_sort.cuh).  Reference: src/graph/edge_features.cpp:7-24
//   g[k] = 1/(sigma*sqrt(2*pi)) * exp(-0.5 * (k*dr - d)^2 / sigma^2),  sigma = r_c / 3
// evaluated in fp32 as 2^(coef * x^2 + log2(norm)), x = k*dr - d: one FADD, one FMUL, one FFMA and one MUFU.EX2 per
// feature.  The exponent lies in (-6.5 + log2 norm, log2 norm]: never denormal, so the bare ex2.approx.ftz is safe and
// the range fix-up exp2f() carries (FSETP + 2 predicated FMUL) is dropped.  Measured error vs the fp64 oracle < 1e-6
// relative (tolerance 1e-5).
#pragma once

#include <cuda_runtime.h>

#include "dgnn_internal.h"

__device__ __forceinline__ float dg_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// four consecutive bins k .. k+3 of the edge at distance d
__device__ __forceinline__ float4 dg_rbf4(float d, int k, const DgRbfEpi& P) {
    const float c0 = (float)k * P.dr;
    const float x0 = c0 - d;
    const float x1 = (c0 + P.dr) - d;
    const float x2 = (c0 + 2.0f * P.dr) - d;
    const float x3 = (c0 + 3.0f * P.dr) - d;
    float4 v;
    v.x = dg_ex2(fmaf(P.coef * x0, x0, P.lognorm));
    v.y = dg_ex2(fmaf(P.coef * x1, x1, P.lognorm));
    v.z = dg_ex2(fmaf(P.coef * x2, x2, P.lognorm));
    v.w = dg_ex2(fmaf(P.coef * x3, x3, P.lognorm));
    return v;
}

__device__ __forceinline__ float dg_rbf1(float d, int k, const DgRbfEpi& P) {
    const float x = (float)k * P.dr - d;
    return dg_ex2(fmaf(P.coef * x, x, P.lognorm));
}
