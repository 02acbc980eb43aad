// Depthwise 7x7 conv (pad 3) + bias of a CNBlock (torchvision/models/convnext.py:52), bf16 mode.
//
//   x [B,H,W,C] fp32 (residual stream)  ->  a [B*H*W, C] bf16 = round(conv + bias)            (GEMM A operand)
//                                           stats [B*H*W, C/32, 2] fp32 = (sum, sum of squares) of the ROUNDED
//                                           values over each 32-channel group
//
// The LayerNorm that follows the conv in the reference (convnext.py:53-54) is not applied here: it is folded into
// the fc1 GEMM (gemm_tma_epi.cuh, EPI_LN_GELU_BF16), whose epilogue needs only the per-token mean and rstd that it
// rebuilds from these partial sums. That removes every cross-warp dependency from this kernel:
//
//   warp   = (32 consecutive channels) x (strip of TW columns) x (row segment) of one image; lane = channel, so every
//            load is one full 128-byte line and the 49 weights of the lane's channel stay in registers;
//   march  : the warp walks down the input rows keeping the 7 in-flight output rows in registers (7*TW accumulators),
//            so an input row is loaded once (TW+6 values, prefetched one row ahead) and feeds 49*TW FMAs;
//            filter rows whose output row lies outside the segment are skipped (warp-uniform branches);
//   stats  : 2*TW values per finished row are reduced over the 32 lanes with a transposing butterfly
//            (4+2+1+1+1 shuffles instead of 8*5) - a fixed summation order, so results are run-to-run identical.
//
// ~110 registers per thread -> 16 warps per SM; warps never synchronise with each other.
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

constexpr int DWS_WARPS = 4;  // warps per CTA (independent units)

template <int C, int TW>
__global__ void __launch_bounds__(DWS_WARPS * 32, 4)
dwconv7_stats_kernel(const float* __restrict__ x, const float* __restrict__ w49, const float* __restrict__ bias,
                     __nv_bfloat16* __restrict__ out, float* __restrict__ stats, int H, int W, int nstrips,
                     int rows_per_seg) {
    static_assert(C % 32 == 0, "one lane per channel, whole warps");
    static_assert(TW == 4, "the statistics butterfly is written for 2*TW = 8 values");
    constexpr int NPART = C / 32;
    constexpr int NIN = TW + 6;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int unit = blockIdx.x * DWS_WARPS + warp;
    if (unit >= nstrips * NPART) return;
    const int cg = unit / nstrips;
    const int strip = unit - cg * nstrips;
    const int c = cg * 32 + lane;
    const int w0 = strip * TW;
    const int h0 = blockIdx.y * rows_per_seg;
    const int nout = min(rows_per_seg, H - h0);
    const int b = blockIdx.z;

    float wr[49];
#pragma unroll
    for (int k = 0; k < 49; ++k) wr[k] = __ldg(w49 + k * C + c);
    const float bs = __ldg(bias + c);

    // column validity of the TW+6 inputs of a row (bit i: column w0-3+i inside the image)
    unsigned colmask = 0;
#pragma unroll
    for (int i = 0; i < NIN; ++i)
        if (w0 - 3 + i >= 0 && w0 - 3 + i < W) colmask |= 1u << i;
    const bool edge = colmask != ((1u << NIN) - 1u);

    const float* xb = x + static_cast<size_t>(b) * H * W * C + c;
    const int nrows_in = nout + 6;

    float cur[NIN], nxt[NIN];
    auto load_row = [&](float (&dst)[NIN], int r) {
        const int hi = h0 - 3 + r;
        if (hi < 0 || hi >= H) return;  // never read: the FMAs of an outside row are skipped
        const float* p = xb + (static_cast<long long>(hi) * W + (w0 - 3)) * C;
        if (edge) {
#pragma unroll
            for (int i = 0; i < NIN; ++i) dst[i] = ((colmask >> i) & 1u) ? __ldg(p + i * C) : 0.f;
        } else {
#pragma unroll
            for (int i = 0; i < NIN; ++i) dst[i] = __ldg(p + i * C);
        }
    };
#pragma unroll
    for (int i = 0; i < NIN; ++i) cur[i] = nxt[i] = 0.f;
    load_row(cur, 0);

    float acc[7][TW];
#pragma unroll
    for (int s = 0; s < 7; ++s)
#pragma unroll
        for (int t = 0; t < TW; ++t) acc[s][t] = 0.f;

    const bool hi16 = (lane & 16) != 0, hi8 = (lane & 8) != 0, hi4 = (lane & 4) != 0;

    for (int rb = 0; rb < nrows_in; rb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int r = rb + j;
            if (r >= nrows_in) break;
            if (r + 1 < nrows_in) load_row(nxt, r + 1);
            const int hi = h0 - 3 + r;
            const bool row_ok = hi >= 0 && hi < H;
            // input row r feeds output row o = r - kh (segment-local) through filter row kh; kh == 0 opens the row
#pragma unroll
            for (int kh = 0; kh < 7; ++kh) {
                const int slot = (j - kh + 7) % 7;
                const int o = r - kh;
                if (kh == 0) {
#pragma unroll
                    for (int t = 0; t < TW; ++t) acc[slot][t] = bs;
                }
                if (row_ok && o >= 0 && o < nout) {
#pragma unroll
                    for (int kw = 0; kw < 7; ++kw)
#pragma unroll
                        for (int t = 0; t < TW; ++t) acc[slot][t] = fmaf(cur[t + kw], wr[kh * 7 + kw], acc[slot][t]);
                }
            }
            if (r >= 6) {
                // output row r - 6 is complete (slot (j + 1) % 7)
                const int slot = (j + 1) % 7;
                const int ho = h0 + r - 6;
                const size_t pix0 = (static_cast<size_t>(b) * H + ho) * W + w0;
                float a[2 * TW];
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    const __nv_bfloat16 hb = __float2bfloat16_rn(acc[slot][t]);
                    const float vr = __bfloat162float(hb);
                    if (w0 + t < W) out[(pix0 + t) * C + c] = hb;
                    a[2 * t] = vr;
                    a[2 * t + 1] = vr * vr;
                }
                // transposing butterfly: afterwards every lane holds the 32-lane total of value index lane >> 2
                float b4[4], c2[2];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float send = hi16 ? a[i] : a[i + 4];
                    const float keep = hi16 ? a[i + 4] : a[i];
                    b4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float send = hi8 ? b4[i] : b4[i + 2];
                    const float keep = hi8 ? b4[i + 2] : b4[i];
                    c2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
                float d = (hi4 ? c2[1] : c2[0]) + __shfl_xor_sync(0xffffffffu, hi4 ? c2[0] : c2[1], 4);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                const int idx = lane >> 2;  // = 2 * t + which
                if ((lane & 3) == 0 && w0 + (idx >> 1) < W)
                    stats[((pix0 + (idx >> 1)) * NPART + cg) * 2 + (idx & 1)] = d;
            }
#pragma unroll
            for (int i = 0; i < NIN; ++i) cur[i] = nxt[i];
        }
    }
}

}  // namespace cpn
