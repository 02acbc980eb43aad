// bf16-mode GELU for the fc1 epilogue: exact-erf GELU (torchvision/models/convnext.py:56, nn.GELU()) approximated by
// a degree-8 polynomial in u = 2 x^2 / C^2 - 1 evaluated with packed fp32 FMAs, two elements per instruction.
// Maximum absolute error 4.2e-5 (tools/fit_gelu.py), two orders of magnitude below the bf16 rounding of the result
// that follows. fp32 mode keeps erff (gemm_tcgen05.cuh).
#pragma once
#include "device_utils.cuh"

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

__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rd;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2*>(&rd);
}
__device__ __forceinline__ float fma_sat(float a, float b, float c) {
    float d;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// Clamp-free variant used by the TMA epilogue: Phi(x) = sat(0.5 + x * R(x^2)), R of degree 6 in x^2 (Horner, packed),
// the [0,1] clamp being the .sat modifier of the last FMA; beyond the fitted range |x| <= 3.9375 the polynomial keeps
// |x * R| >= 0.5 with the sign of x up to overflow (checked by tools/fit_gelu2.py), so no input clamp is needed.
// 10 instructions per element pair instead of 17. Maximum absolute error 1.94e-4 (at x = 0.65, i.e. 4e-4 relative),
// below the half-ulp of the bf16 rounding that follows (2e-3 relative).
// tools/fit_gelu2.py: fit on |x| <= 3.9375, degree 6 in x^2
#define CPN_GELU_R0 0.397860432f
#define CPN_GELU_R1 -0.0645769476f
#define CPN_GELU_R2 0.0087859431f
#define CPN_GELU_R3 -0.0008187387f
#define CPN_GELU_R4 4.86081859e-05f
#define CPN_GELU_R5 -1.63859936e-06f
#define CPN_GELU_R6 2.36860857e-08f

__device__ __forceinline__ float2 gelu_sat2(float2 x) {
    const float2 v = fmul2(x, x);
    float2 r = ffma2(make_float2(CPN_GELU_R6, CPN_GELU_R6), v, make_float2(CPN_GELU_R5, CPN_GELU_R5));
    r = ffma2(r, v, make_float2(CPN_GELU_R4, CPN_GELU_R4));
    r = ffma2(r, v, make_float2(CPN_GELU_R3, CPN_GELU_R3));
    r = ffma2(r, v, make_float2(CPN_GELU_R2, CPN_GELU_R2));
    r = ffma2(r, v, make_float2(CPN_GELU_R1, CPN_GELU_R1));
    r = ffma2(r, v, make_float2(CPN_GELU_R0, CPN_GELU_R0));
    const float2 phi = make_float2(fma_sat(x.x, r.x, 0.5f), fma_sat(x.y, r.y, 0.5f));
    return fmul2(x, phi);
}


// fp32-mode GELU: the exact erf form 0.5 x (1 + erf(x / sqrt 2)) = 0.5 x + |x| (0.5 - 0.5 erfc(|x| / sqrt 2)) with
// erfc(z) = (a1 t + .. + a5 t^5) exp(-z^2), t = 1 / (1 + p z) (Abramowitz & Stegun 7.1.26, |error| <= 1.5e-7). Evaluated on
// both halves with packed FMAs, two MUFU.RCP and two MUFU.EX2; z is pre-scaled by sqrt(log2 e) so that exp(-z^2) is one
// ex2. Measured in fp32 emulation over [-12, 12]: max |error| 4.7e-7 (1.9e-7 relative to max(|x|, 1)) -- below the
// 1.2e-6 of torch's own fp32 gelu against fp64, and about half the instructions of erff (26 per element).
__device__ __forceinline__ float2 gelu_erfc2(float2 x) {
    constexpr float C1 = 0.84932180028801904272f;   // sqrt(log2 e) / sqrt 2
    constexpr float PP = 0.27273953202083960690f;   // 0.3275911 / sqrt(log2 e)
    const float2 ax = make_float2(fabsf(x.x), fabsf(x.y));
    const float2 z = fmul2(ax, make_float2(C1, C1));
    const float2 den = ffma2(z, make_float2(PP, PP), make_float2(1.f, 1.f));
    float2 t, ex;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(den.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(den.y));
    const float2 zz = fmul2(z, z);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex.x) : "f"(-zz.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex.y) : "f"(-zz.y));
    float2 q = ffma2(t, make_float2(1.061405429f, 1.061405429f), make_float2(-1.453152027f, -1.453152027f));
    q = ffma2(q, t, make_float2(1.421413741f, 1.421413741f));
    q = ffma2(q, t, make_float2(-0.284496736f, -0.284496736f));
    q = ffma2(q, t, make_float2(0.254829592f, 0.254829592f));
    q = fmul2(q, t);
    const float2 h = fmul2(q, ex);                                                     // erfc(|x| / sqrt 2)
    const float2 d = ffma2(h, make_float2(-0.5f, -0.5f), make_float2(0.5f, 0.5f));  // 0.5 - 0.5 erfc
    return ffma2(ax, d, fmul2(x, make_float2(0.5f, 0.5f)));
}

}  // namespace cpn
