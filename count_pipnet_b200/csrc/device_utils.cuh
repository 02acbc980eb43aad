// Small device helpers shared by the kernels: warp reductions, packed fp32x2 arithmetic, bf16 packing.
#pragma once
#include <type_traits>

#include "ptx.cuh"

namespace cpn {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Packed fp32x2 FMA (FFMA2 on sm_100a): d = a * b + c on both halves.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a);
    rb = *reinterpret_cast<unsigned long long*>(&b);
    rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}

// bf16(a) in the low half, bf16(b) in the high half (round to nearest even): one F2FP instead of two F2F + shift/or.
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    uint32_t d;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
    return d;
}
__device__ __forceinline__ float bf16_round(float a) { return __bfloat162float(__float2bfloat16_rn(a)); }
__device__ __forceinline__ uint32_t pack2(float a, float b) { return pack_bf16x2(a, b); }

// Compile-time loop: f(std::integral_constant<int, I>) for I in [0, N), so that I can be an asm immediate.
template <int I, int N>
struct StaticFor {
    template <class F>
    __device__ __forceinline__ static void run(F&& f) {
        f(std::integral_constant<int, I>{});
        StaticFor<I + 1, N>::run(f);
    }
};
template <int N>
struct StaticFor<N, N> {
    template <class F>
    __device__ __forceinline__ static void run(F&&) {}
};

}  // namespace cpn
