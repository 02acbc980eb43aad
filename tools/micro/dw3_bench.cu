// Microbenchmark for a "lanes over space" depthwise 7x7: a lane owns one channel and a block of RB output rows x 4 output
// columns, its 49 taps in registers (scalar FFMA), inputs read from a channel-major shared-memory tile with LDS.128.
// Question: what fraction of the FP32 FMA peak does the compute phase reach at 12-16 warps per SM (<= 128 registers)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/dw3_bench tools/micro/dw3_bench.cu
#include <cuda_runtime.h>
#include <cstdio>

constexpr int TW = 4;

template <int RB, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
dw3_compute(const float* __restrict__ w49, float* __restrict__ sink, int iters, int NS, int NB, int wrow, int plane_in,
            int plane_out) {
    extern __shared__ __align__(16) float smem[];
    const int U = NS * NB;
    const int CH = THREADS / U;
    float* tin = smem;
    float* tout = smem + CH * plane_in;
    for (int i = threadIdx.x; i < CH * plane_in; i += THREADS) tin[i] = 1e-3f * static_cast<float>(i % 97);
    __syncthreads();
    const int c = threadIdx.x / U, u = threadIdx.x % U, k = u / NS, s = u % NS;
    if (c >= CH) return;
    float wr[49];
#pragma unroll
    for (int i = 0; i < 49; ++i) wr[i] = __ldg(w49 + (blockIdx.x % 7) * 49 + i);
    const float* src0 = tin + c * plane_in + k * RB * wrow + 4 * s;
    float* dst0 = tout + c * plane_out + k * RB * (NS * TW) + 4 * s;
    float total = 0.f;
    for (int it = 0; it < iters; ++it) {
        float acc[RB][TW];
#pragma unroll
        for (int o = 0; o < RB; ++o)
#pragma unroll
            for (int t = 0; t < TW; ++t) acc[o][t] = 0.5f;
#pragma unroll
        for (int i = 0; i < RB + 6; ++i) {
            float x[12];
            const float4* p = reinterpret_cast<const float4*>(src0 + i * wrow);
            const float4 a = p[0], b = p[1], d = p[2];
            x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
            x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
            x[8] = d.x; x[9] = d.y; x[10] = d.z; x[11] = d.w;
#pragma unroll
            for (int kh = 0; kh < 7; ++kh) {
                const int o = i - kh;
                if (o < 0 || o >= RB) continue;
#pragma unroll
                for (int kw = 0; kw < 7; ++kw)
#pragma unroll
                    for (int t = 0; t < TW; ++t) acc[o][t] = fmaf(x[1 + t + kw], wr[kh * 7 + kw], acc[o][t]);
            }
        }
#pragma unroll
        for (int o = 0; o < RB; ++o)
            *reinterpret_cast<float4*>(dst0 + o * (NS * TW)) = make_float4(acc[o][0], acc[o][1], acc[o][2], acc[o][3]);
        total += acc[0][0];
        src0 += (it & 1) ? -4 : 4;  // keep the loads from being hoisted
    }
    if (total == 1234.5f) sink[threadIdx.x] = total;
}

template <int RB, int THREADS, int MINB>
void run(const float* w, float* sink, int NS, int NB) {
    const int U = NS * NB, CH = THREADS / U;
    const int wrow = NS * TW + 8 + 4;
    const int plane_in = (NB * RB + 6) * wrow + 8, plane_out = NB * RB * NS * TW;
    const size_t smem = static_cast<size_t>(CH) * (plane_in + plane_out) * 4;
    auto kern = dw3_compute<RB, THREADS, MINB>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, THREADS, smem);
    const int iters = 2000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = 148 * (nb > 0 ? nb : 1);
    kern<<<grid, THREADS, smem>>>(w, sink, 10, NS, NB, wrow, plane_in, plane_out);
    cudaError_t e = cudaDeviceSynchronize();
    cudaEventRecord(e0);
    kern<<<grid, THREADS, smem>>>(w, sink, iters, NS, NB, wrow, plane_in, plane_out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fma = static_cast<double>(grid) * CH * U * iters * RB * TW * 49;
    printf("RB=%d threads=%d (%d used) ctas/SM=%d regs=%d smem=%zu KB NS=%d NB=%d: %.3f ms  %.2f TFMA/s = %.1f FMA/clk/SM at 1.965 GHz  (%s)\n",
           RB, THREADS, CH * U, nb, fa.numRegs, smem >> 10, NS, NB, ms, fma / ms / 1e9, fma / ms / 1e9 * 1e12 / 148 / 1.965e9 * 1e-3 * 1e3 / 1e3 * 1e0,
           cudaGetErrorString(e));
}

int main() {
    float *w, *sink;
    cudaMalloc(&w, 7 * 49 * 4);
    cudaMalloc(&sink, 4096);
    cudaMemset(w, 0, 7 * 49 * 4);
    run<7, 448, 1>(w, sink, 14, 2);   // 56-wide maps: 14 strips x 2 row blocks x 16 channels
    run<7, 448, 1>(w, sink, 7, 4);    // 28-wide maps
    run<7, 224, 2>(w, sink, 7, 4);    // 8 channels per CTA, two CTAs per SM
    run<7, 224, 2>(w, sink, 14, 2);
    run<7, 512, 1>(w, sink, 8, 4);    // 32 units x 16 channels
    run<7, 256, 2>(w, sink, 8, 4);
    run<6, 448, 1>(w, sink, 14, 2);
    run<4, 512, 1>(w, sink, 8, 4);
    return 0;
}
