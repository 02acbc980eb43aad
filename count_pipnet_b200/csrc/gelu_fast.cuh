// bf16-mode GELU for the fc1 epilogue: exact-erf GELU (torchvision/models/convnext.py:56, nn.GELU()) approximated by
// a degree-8 polynomial in u = 2 x^2 / C^2 - 1 evaluated with packed fp32 FMAs, two elements per instruction.
// Maximum absolute error 4.2e-5 (tools/fit_gelu.py), two orders of magnitude below the bf16 rounding of the result
// that follows. fp32 mode keeps erff (gemm_tcgen05.cuh).
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

// tools/fit_gelu.py: C = 4.5, degree 8, max |gelu_fast - gelu| = 4.134e-05 at x = 0.591
constexpr float GELU_C = 4.5f;
#define CPN_GELU_P0 0.313804163f
#define CPN_GELU_P1 -0.154357051f
#define CPN_GELU_P2 0.109652691f
#define CPN_GELU_P3 -0.0809524889f
#define CPN_GELU_P4 0.0550852001f
#define CPN_GELU_P5 -0.0339379044f
#define CPN_GELU_P6 0.0248944433f
#define CPN_GELU_P7 -0.0183107892f
#define CPN_GELU_P8 0.00634496979f

__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rd;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2*>(&rd);
}

// gelu of both halves of x.
__device__ __forceinline__ float2 gelu_fast2(float2 x) {
    float2 xc;
    xc.x = fminf(fmaxf(x.x, -GELU_C), GELU_C);
    xc.y = fminf(fmaxf(x.y, -GELU_C), GELU_C);
    const float k = 2.0f / (GELU_C * GELU_C);
    const float2 u = ffma2(fmul2(xc, make_float2(k, k)), xc, make_float2(-1.f, -1.f));
    float2 r = make_float2(CPN_GELU_P8, CPN_GELU_P8);
    r = ffma2(r, u, make_float2(CPN_GELU_P7, CPN_GELU_P7));
    r = ffma2(r, u, make_float2(CPN_GELU_P6, CPN_GELU_P6));
    r = ffma2(r, u, make_float2(CPN_GELU_P5, CPN_GELU_P5));
    r = ffma2(r, u, make_float2(CPN_GELU_P4, CPN_GELU_P4));
    r = ffma2(r, u, make_float2(CPN_GELU_P3, CPN_GELU_P3));
    r = ffma2(r, u, make_float2(CPN_GELU_P2, CPN_GELU_P2));
    r = ffma2(r, u, make_float2(CPN_GELU_P1, CPN_GELU_P1));
    r = ffma2(r, u, make_float2(CPN_GELU_P0, CPN_GELU_P0));
    const float2 phi = ffma2(fmul2(xc, make_float2(0.5f, 0.5f)), r, make_float2(0.5f, 0.5f));
    return fmul2(x, phi);
}

}  // namespace cpn
