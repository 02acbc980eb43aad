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
//   lane   = one channel PAIR, so the 49 FMAs of a tap are packed fma.rn.f32x2 (FFMA2: one issue slot per two FMAs,
//            which leaves issue slots for the loads, the rounding and the statistics) and every access is 8 bytes;
//   half-warp (16 lanes) = 32 consecutive channels x a strip of TW columns x a row segment of one image; the two
//            halves of a warp are consecutive (strip, channel-group) units, so any C that is a multiple of 32 fills
//            whole warps (C = 96 -> 3 half-warps per strip);
//   march  : a lane walks down the input rows keeping the 7 in-flight output rows in registers (7*TW packed
//            accumulators) and its 49 packed weights in registers; an input row is loaded once (TW+6 values,
//            prefetched one row ahead into the other of two buffers) and feeds 49*TW packed FMAs; filter rows whose
//            output row lies outside the segment are skipped (warp-uniform branches);
//   stats  : 2*TW values per finished row are reduced over the 16 lanes of the half-warp with a transposing
//            butterfly (4+2+1+1 shuffles instead of 8*4) - a fixed summation order, so results are run-to-run
//            identical.
//
// ~230 registers per thread -> 8 warps per SM, each with 28 independent accumulator chains.
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

constexpr int DWS_WARPS = 2;  // warps per CTA (independent of each other)
constexpr int DWS_TW = 4;     // output columns per lane
constexpr int DWS_PF = 4;     // L2 prefetch distance in rows

__device__ __forceinline__ float2 ldg_f2(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }

template <int C>
__global__ void __launch_bounds__(DWS_WARPS * 32, 4)
dwconv7_stats_kernel(const float* __restrict__ x, const float* __restrict__ w49, const float* __restrict__ bias,
                     __nv_bfloat16* __restrict__ out, float* __restrict__ stats, int H, int W, int nstrips,
                     int rows_per_seg, int pf) {
    static_assert(C % 32 == 0, "32-channel groups");
    constexpr int TW = DWS_TW;
    constexpr int NPART = C / 32;
    constexpr int NIN = TW + 6;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int hl = lane & 15;
    const int nunits = nstrips * NPART;
    int unit = (blockIdx.x * DWS_WARPS + warp) * 2 + (lane >> 4);
    if (unit - (lane >> 4) >= nunits) return;  // whole warp beyond the end
    const bool live = unit < nunits;          // the upper half of the last warp may have no unit
    if (!live) unit = nunits - 1;             // compute a duplicate, store nothing
    const int strip = unit / NPART;
    const int cg = unit - strip * NPART;
    const int c0 = cg * 32 + 2 * hl;
    const int w0 = strip * TW;
    const int h0 = blockIdx.y * rows_per_seg;
    const int nout = min(rows_per_seg, H - h0);
    const int b = blockIdx.z;

    float2 wr[49];
#pragma unroll
    for (int k = 0; k < 49; ++k) wr[k] = ldg_f2(w49 + k * C + c0);
    const float2 bs = ldg_f2(bias + c0);

    // column validity of the TW+6 inputs of a row (bit i: column w0-3+i inside the image)
    unsigned colmask = 0;
#pragma unroll
    for (int i = 0; i < NIN; ++i)
        if (w0 - 3 + i >= 0 && w0 - 3 + i < W) colmask |= 1u << i;
    const bool edge = __any_sync(0xffffffffu, colmask != ((1u << NIN) - 1u));

    const float* xb = x + static_cast<size_t>(b) * H * W * C + c0 + static_cast<long long>(w0 - 3) * C;
    const int nrows_in = nout + 6;

    float2 bufa[NIN], bufb[NIN];
    auto load_row = [&](float2 (&dst)[NIN], int r) {
        const int hi = h0 - 3 + r;
        if (hi < 0 || hi >= H) return;  // never read: the FMAs of an outside row are skipped
        const float* p = xb + static_cast<long long>(hi) * W * C;
        if (edge) {
#pragma unroll
            for (int i = 0; i < NIN; ++i)
                if ((colmask >> i) & 1u) dst[i] = ldg_f2(p + i * C);  // outside columns keep their initial zero
        } else {
#pragma unroll
            for (int i = 0; i < NIN; ++i) dst[i] = ldg_f2(p + i * C);
        }
    };
    // L2 prefetch DWS_PF rows ahead (one 128-byte line per half-warp and column): the register prefetch above covers an
    // L2 hit (~250 cycles), not a DRAM miss
    auto prefetch_row = [&](int r) {
        const int hi = h0 - 3 + r;
        if (hi < 0 || hi >= H || r >= nrows_in || pf <= 0) return;
        const float* p = xb + static_cast<long long>(hi) * W * C;
        if (hl == 0) {
#pragma unroll
            for (int i = 0; i < NIN; ++i)
                if ((colmask >> i) & 1u) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + i * C));
        }
    };
#pragma unroll
    for (int i = 0; i < NIN; ++i) bufa[i] = bufb[i] = make_float2(0.f, 0.f);
    load_row(bufa, 0);
    for (int r = 1; r < pf; ++r) prefetch_row(r);

    float2 acc[7][TW];
#pragma unroll
    for (int s = 0; s < 7; ++s)
#pragma unroll
        for (int t = 0; t < TW; ++t) acc[s][t] = make_float2(0.f, 0.f);

    const bool hi8 = (lane & 8) != 0, hi4 = (lane & 4) != 0, hi2 = (lane & 2) != 0;

    for (int rb = 0; rb < nrows_in; rb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int r = rb + j;
            if (r >= nrows_in) break;
            // rows alternate between the two input buffers; 7 is odd, so the roles are restored by a copy at the
            // end of each group of 7 (below) instead of after every row
            float2(&cur)[NIN] = (j & 1) ? bufb : bufa;
            float2(&nxt)[NIN] = (j & 1) ? bufa : bufb;
            if (r + 1 < nrows_in) load_row(nxt, r + 1);
            prefetch_row(r + pf);
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
                        for (int t = 0; t < TW; ++t) acc[slot][t] = ffma2(cur[t + kw], wr[kh * 7 + kw], acc[slot][t]);
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
                    const uint32_t pk = pack_bf16x2(acc[slot][t].x, acc[slot][t].y);
                    if (live && w0 + t < W) *reinterpret_cast<uint32_t*>(out + (pix0 + t) * C + c0) = pk;
                    const float v0 = __uint_as_float(pk << 16), v1 = __uint_as_float(pk & 0xffff0000u);
                    a[2 * t] = v0 + v1;
                    a[2 * t + 1] = fmaf(v1, v1, v0 * v0);
                }
                // transposing butterfly over the half-warp: afterwards lane hl holds the total of value index hl >> 1
                float b4[4], c2[2];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float send = hi8 ? a[i] : a[i + 4];
                    const float keep = hi8 ? a[i + 4] : a[i];
                    b4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float send = hi4 ? b4[i] : b4[i + 2];
                    const float keep = hi4 ? b4[i + 2] : b4[i];
                    c2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                }
                float d = (hi2 ? c2[1] : c2[0]) + __shfl_xor_sync(0xffffffffu, hi2 ? c2[0] : c2[1], 2);
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                const int idx = hl >> 1;  // = 2 * t + which
                if (live && (lane & 1) == 0 && w0 + (idx >> 1) < W)
                    stats[((pix0 + (idx >> 1)) * NPART + cg) * 2 + (idx & 1)] = d;
            }
        }
        // after 7 rows the next row (already prefetched) sits in bufb; the next group starts on bufa
#pragma unroll
        for (int i = 0; i < NIN; ++i) bufa[i] = bufb[i];
    }
}

}  // namespace cpn
