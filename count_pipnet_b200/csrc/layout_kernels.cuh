// Layout / gather kernels: NCHW<->NHWC transposes at the module boundary, and the im2col / pooling / subsample
// gathers that turn the ResNet convolutions of features/resnet_features.py into rows for the tcgen05 GEMM.
#pragma once
#include "device_utils.cuh"

namespace cpn {

// [B, HW, C] -> [B, C, HW] (and back) through a 32x33 shared-memory tile; grid = (ceil(HW/32), ceil(C/32), B).
__global__ void __launch_bounds__(256)
transpose_last2_kernel(const float* __restrict__ in, float* __restrict__ out, int R, int Ccols) {
    // in: [B, R, Ccols], out: [B, Ccols, R]
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const float* ib = in + static_cast<size_t>(b) * R * Ccols;
    float* ob = out + static_cast<size_t>(b) * R * Ccols;
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        tile[i][tx] = (r < R && c < Ccols) ? ib[static_cast<size_t>(r) * Ccols + c] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;
        if (r < R && c < Ccols) ob[static_cast<size_t>(c) * R + r] = tile[tx][i];
    }
}

// im2col over a channels-last bf16 map, 8 channels (16 B) per thread.
__global__ void __launch_bounds__(256)
im2col_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int B, int H, int W, int C, int k, int stride,
                   int pad, int Ho, int Wo, int Kpad) {
    const int c8 = C >> 3, k8 = Kpad >> 3;
    const long long total = static_cast<long long>(B) * Ho * Wo * k8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int col = static_cast<int>(i % k8);
        const long long row = i / k8;
        const int wo = static_cast<int>(row % Wo);
        const int ho = static_cast<int>((row / Wo) % Ho);
        const int b = static_cast<int>(row / (static_cast<long long>(Wo) * Ho));
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        const int tap = col / c8, cc = col % c8;
        if (tap < k * k) {
            const int kh = tap / k, kw = tap % k;
            const int hi = ho * stride - pad + kh, wi = wo * stride - pad + kw;
            if (hi >= 0 && hi < H && wi >= 0 && wi < W)
                v = __ldg(x + ((static_cast<size_t>(b) * H + hi) * W + wi) * c8 + cc);
        }
        out[i] = v;
    }
}

// im2col straight from the fp32 NCHW network input; column = c*k*k + kh*k + kw (the torch weight flattening).
// A thread produces 8 consecutive columns of one row (16-byte stores of the hi and lo planes); the per-column tap offsets
// come from a small shared-memory table instead of divisions per element.
__global__ void __launch_bounds__(256)
im2col_nchw_f32_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo,
                       int B, int C, int H, int W, int k, int stride, int pad, int Ho, int Wo, int Kpad) {
    extern __shared__ int s_tab[];  // [Kpad] offset c*H*W + kh*W + kw, then [Kpad] (kh << 16 | kw) or -1 for padding columns
    int* s_off = s_tab;
    int* s_khw = s_tab + Kpad;
    const int kk = k * k;
    for (int col = threadIdx.x; col < Kpad; col += blockDim.x) {
        if (col < C * kk) {
            const int c = col / kk, r = col - c * kk;
            const int kh = r / k, kw = r - kh * k;
            s_off[col] = (c * H + kh) * W + kw;
            s_khw[col] = (kh << 16) | kw;
        } else {
            s_off[col] = 0;
            s_khw[col] = -1;
        }
    }
    __syncthreads();
    const int groups = Kpad >> 3;
    const long long total = static_cast<long long>(B) * Ho * Wo * groups;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int grp = static_cast<int>(i % groups);
        const long long row = i / groups;
        const int wo = static_cast<int>(row % Wo);
        const int rest = static_cast<int>(row / Wo);
        const int ho = rest % Ho, b = rest / Ho;
        const int h0 = ho * stride - pad, w0 = wo * stride - pad;
        const float* xb = x + static_cast<size_t>(b) * C * H * W + static_cast<long long>(h0) * W + w0;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int col = grp * 8 + e;
            const int khw = s_khw[col];
            const int hi = h0 + (khw >> 16), wi = w0 + (khw & 0xffff);
            v[e] = (khw >= 0 && hi >= 0 && hi < H && wi >= 0 && wi < W) ? __ldg(xb + s_off[col]) : 0.f;
        }
        uint4 hi4, lo4;
        hi4.x = pack_bf16x2(v[0], v[1]); hi4.y = pack_bf16x2(v[2], v[3]);
        hi4.z = pack_bf16x2(v[4], v[5]); hi4.w = pack_bf16x2(v[6], v[7]);
        const size_t o = static_cast<size_t>(row) * Kpad + grp * 8;
        *reinterpret_cast<uint4*>(out + o) = hi4;
        if (out_lo != nullptr) {
            lo4.x = pack_bf16x2(v[0] - __uint_as_float(hi4.x << 16), v[1] - __uint_as_float(hi4.x & 0xffff0000u));
            lo4.y = pack_bf16x2(v[2] - __uint_as_float(hi4.y << 16), v[3] - __uint_as_float(hi4.y & 0xffff0000u));
            lo4.z = pack_bf16x2(v[4] - __uint_as_float(hi4.z << 16), v[5] - __uint_as_float(hi4.z & 0xffff0000u));
            lo4.w = pack_bf16x2(v[6] - __uint_as_float(hi4.w << 16), v[7] - __uint_as_float(hi4.w & 0xffff0000u));
            *reinterpret_cast<uint4*>(out_lo + o) = lo4;
        }
    }
}

// MaxPool2d(kernel 3, stride 2, pad 1) on a channels-last fp32 map; writes fp32 and/or the bf16 hi (+lo) planes.
__global__ void __launch_bounds__(256)
maxpool3s2_kernel(const float* __restrict__ x, float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_hi,
                  __nv_bfloat16* __restrict__ out_lo, int B, int H, int W, int C, int Ho, int Wo) {
    const long long total = static_cast<long long>(B) * Ho * Wo * C;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int c = static_cast<int>(i % C);
        const long long row = i / C;
        const int wo = static_cast<int>(row % Wo);
        const int ho = static_cast<int>((row / Wo) % Ho);
        const int b = static_cast<int>(row / (static_cast<long long>(Wo) * Ho));
        float m = -INFINITY;
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * 2 - 1 + kh;
            if (hi < 0 || hi >= H) continue;
            for (int kw = 0; kw < 3; ++kw) {
                const int wi = wo * 2 - 1 + kw;
                if (wi < 0 || wi >= W) continue;
                m = fmaxf(m, __ldg(x + ((static_cast<size_t>(b) * H + hi) * W + wi) * C + c));
            }
        }
        if (out_f32 != nullptr) out_f32[i] = m;
        if (out_hi != nullptr) {
            const __nv_bfloat16 h = __float2bfloat16_rn(m);
            out_hi[i] = h;
            if (out_lo != nullptr) out_lo[i] = __float2bfloat16_rn(m - __bfloat162float(h));
        }
    }
}

// The same pooling on a channels-last bf16 map, 8 channels (16 bytes) per thread. Rounding to bf16 is monotonic, so
// max(bf16(a), bf16(b)) == bf16(max(a, b)): pooling the bf16 stem output gives the bits the fp32 pooling + rounding gives.
__global__ void __launch_bounds__(256)
maxpool3s2_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int B, int H, int W, int C, int Ho, int Wo) {
    const int c8 = C >> 3;
    const long long total = static_cast<long long>(B) * Ho * Wo * c8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cc = static_cast<int>(i % c8);
        const long long row = i / c8;
        const int wo = static_cast<int>(row % Wo);
        const int ho = static_cast<int>((row / Wo) % Ho);
        const int b = static_cast<int>(row / (static_cast<long long>(Wo) * Ho));
        const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
        __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * 2 - 1 + kh;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int wi = wo * 2 - 1 + kw;
                if (wi < 0 || wi >= W) continue;
                const uint4 v = __ldg(x + ((static_cast<size_t>(b) * H + hi) * W + wi) * c8 + cc);
                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) m[e] = __hmax2(m[e], *reinterpret_cast<const __nv_bfloat162*>(&w4[e]));
            }
        }
        uint4 o;
        o.x = *reinterpret_cast<uint32_t*>(&m[0]); o.y = *reinterpret_cast<uint32_t*>(&m[1]);
        o.z = *reinterpret_cast<uint32_t*>(&m[2]); o.w = *reinterpret_cast<uint32_t*>(&m[3]);
        out[i] = o;
    }
}

// out[b, ho, wo, :] = x[b, 2*ho, 2*wo, :]
__global__ void __launch_bounds__(256)
subsample2_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int B, int H, int W, int C, int Ho, int Wo) {
    const int c8 = C >> 3;
    const long long total = static_cast<long long>(B) * Ho * Wo * c8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cc = static_cast<int>(i % c8);
        const long long row = i / c8;
        const int wo = static_cast<int>(row % Wo);
        const int ho = static_cast<int>((row / Wo) % Ho);
        const int b = static_cast<int>(row / (static_cast<long long>(Wo) * Ho));
        out[i] = __ldg(x + ((static_cast<size_t>(b) * H + 2 * ho) * W + 2 * wo) * c8 + cc);
    }
}

}  // namespace cpn
