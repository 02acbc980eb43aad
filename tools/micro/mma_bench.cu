// Microbenchmark: throughput of the legacy warp-level tensor-core path (mma.sync) on sm_100a, to decide whether the
// depthwise 7x7 conv can run as banded-Toeplitz MMAs. Prints MMAs per SM per cycle-equivalent and TFLOP/s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/mma_bench tools/micro/mma_bench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

template <int NACC, bool TF32>
__global__ void mma_loop(float* out, int iters) {
    float acc[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
    uint32_t a0 = threadIdx.x * 0x3c003c00u, a1 = 0x3c003c00u, a2 = 0x3c003c00u, a3 = 0x3c003c00u;
    uint32_t b0 = 0x3c003c00u + blockIdx.x, b1 = 0x3c003c00u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) {
            if (TF32)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    if (s == 12345.678f) out[0] = s;
}

template <int NACC, bool TF32>
void run(int warps_per_sm, int sms, float* d) {
    const int iters = 4096;
    const int threads = 128;
    const int ctas = sms * warps_per_sm / 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    mma_loop<NACC, TF32><<<ctas, threads>>>(d, 16);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    mma_loop<NACC, TF32><<<ctas, threads>>>(d, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double mmas = static_cast<double>(ctas) * 4 * iters * NACC;
    const double flop = mmas * 2.0 * 16 * 8 * (TF32 ? 8 : 16);
    printf("%s NACC=%d warps/SM=%2d: %.3f ms  %.1f TFLOP/s  %.2f MMA/SM/ns\n", TF32 ? "tf32 m16n8k8 " : "bf16 m16n8k16", NACC,
           warps_per_sm, ms, flop / ms / 1e9, mmas / sms / (ms * 1e6));
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* d;
    cudaMalloc(&d, 4);
    for (int w : {4, 8, 16, 32}) run<8, false>(w, sms, d);
    for (int w : {4, 8, 16}) run<14, false>(w, sms, d);
    for (int w : {4, 8, 16, 32}) run<8, true>(w, sms, d);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
