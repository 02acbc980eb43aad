// Microbenchmark: throughput of fma.rn.f32x2 (FFMA2) against scalar FFMA on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b),
                       rc = *reinterpret_cast<unsigned long long*>(&c), rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
template <int CH, bool PACKED>
__global__ void k(float* out, int iters, float s) {
    float2 acc[CH];
    float2 w = make_float2(s, s * 0.5f), x = make_float2(threadIdx.x * 1e-3f, 1.0f);
    for (int i = 0; i < CH; ++i) acc[i] = make_float2(i, -i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (PACKED) acc[i] = ffma2(x, w, acc[i]);
            else { acc[i].x = fmaf(x.x, w.x, acc[i].x); acc[i].y = fmaf(x.y, w.y, acc[i].y); }
        }
    }
    float r = 0;
    for (int i = 0; i < CH; ++i) r += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int CH, bool PACKED>
void run(const char* name, int threads) {
    float* out; cudaMalloc(&out, 148 * 8 * 1024 * 4);
    int iters = 20000;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<CH, PACKED><<<148, threads>>>(out, 100, 1.0001f);
    cudaEventRecord(a);
    k<CH, PACKED><<<148, threads>>>(out, iters, 1.0001f);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double fma = 148.0 * threads * iters * CH * 2;
    printf("%-28s threads=%4d chains=%2d: %.3f ms  %.2f TFMA/s (%.1f FMA/clk/SM @1.9GHz)\n", name, threads, CH, ms,
           fma / ms / 1e9, fma / ms / 1e9 * 1e12 / 148 / 1.9e9 / 1e0 / 1e0 * 1e-0 / 1.0);
    cudaFree(out);
}
int main() {
    run<16, true>("FFMA2", 128); run<16, false>("FFMA", 128);
    run<16, true>("FFMA2", 256); run<16, false>("FFMA", 256);
    run<16, true>("FFMA2", 512); run<16, false>("FFMA", 512);
    run<4, true>("FFMA2", 256); run<4, false>("FFMA", 256);
    run<28, true>("FFMA2", 192); run<28, false>("FFMA", 192);
    return 0;
}
