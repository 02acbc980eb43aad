// Depthwise 7x7 (pad 3) + bias + per-token LayerNorm of a ConvNeXt block.
//
//   * thread = (column strip g, channel pair): its 49 packed weights live in registers and it marches down the rows
//     keeping the 7 in-flight output rows of its TW tokens in registers, so every value read from shared memory feeds
//     7*TW packed FMAs (fma.rn.f32x2);
//   * input rows (G*TW+6 tokens x C channels, fp32) arrive in a shared-memory ring through TMA 4-D box loads issued by
//     thread 0; the zero padding of the convolution is the TMA out-of-bounds fill, and rows outside the image are
//     neither loaded nor computed;
//   * LayerNorm statistics: each thread adds its two channels, a packed butterfly reduces the 2*TW sums of a warp in
//     9 shuffles, one shared-memory atomic per warp and value finishes the row, and an mbarrier arrival publishes it.
//     The row is normalised two rows later (split-phase barrier: nobody ever waits for the slowest warp) from the
//     values the thread kept in registers, rounded to bf16 and stored 128 B per warp and token.
//
// Reference arithmetic: CNBlock.block[0] (Conv2d 7x7 groups=C) -> Permute -> LayerNorm(C, eps 1e-6),
// torchvision/models/convnext.py:52-54.  Layout: x [B,H,W,C] fp32 channels-last, out [B*H*W, C] bf16.
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

constexpr int DW2_MAX_STAGES = 8;
constexpr int DW2_NSTAT = 16;  // ring of per-row statistics slots (see the zeroing protocol in the kernel)
constexpr int DW2_MAXG = 4;

struct Dw2Params {
    const float* w49;   // [49, C]
    const float* bias;  // [C]
    const float* ln_w;  // [C]
    const float* ln_b;  // [C]
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;  // or null
    int H, W, C;
    int G;             // column strips per CTA
    int gthreads;      // threads per strip: C/2 rounded up to a whole number of warps
    int CB;            // channels per TMA box
    int nbox;          // C / CB
    int box_bytes;     // (G*TW+6) * CB * 4, rounded up to 128
    int stage_bytes;   // nbox * box_bytes
    int nst;           // ring depth
    int rows_per_cta;  // output rows per CTA (H when the grid has one row chunk)
    float eps;
};

// One packed cvt (F2FP): low half = a, high half = b.
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// Butterfly that reduces 8 per-lane values over the warp in 9 shuffles. On return `d` holds, in every lane, the
// warp-wide sum of value number ((lane >> 2) & 7) with its bits reversed: index = 4*bit4 + 2*bit3 + bit2.
__device__ __forceinline__ float warp_reduce8(const float (&a)[8], int lane) {
    const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
    float k4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = b4 ? a[i] : a[i + 4];
        const float keep = b4 ? a[i + 4] : a[i];
        k4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float k2[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float send = b3 ? k4[i] : k4[i + 2];
        const float keep = b3 ? k4[i + 2] : k4[i];
        k2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    const float send = b2 ? k2[0] : k2[1];
    float d = (b2 ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, send, 4);
    d += __shfl_xor_sync(0xffffffffu, d, 2);
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    return d;
}

template <int TW, int NT>
__global__ void __launch_bounds__(NT, 1)
dwconv7_ln_tma_kernel(const __grid_constant__ CUtensorMap tmap_x, const Dw2Params p) {
    extern __shared__ uint8_t dw2_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dw2_smem_raw) + 127) & ~uintptr_t(127));
    const int C = p.C;
    const int cols = p.G * TW;  // output tokens per row handled by this CTA
    uint8_t* in_ring = smem;
    float* stats = reinterpret_cast<float*>(smem + static_cast<size_t>(p.nst) * p.stage_bytes);  // [NSTAT][MAXG][8]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(stats + DW2_NSTAT * DW2_MAXG * 8);
    uint64_t* empty_bar = full_bar + DW2_MAX_STAGES;
    uint64_t* stat_bar = empty_bar + DW2_MAX_STAGES;

    const int tid = threadIdx.x, lane = tid & 31;
    const int w0 = blockIdx.x * cols;
    const int h0 = blockIdx.y * p.rows_per_cta;
    const int nout = min(p.rows_per_cta, p.H - h0);
    const int b = blockIdx.z;
    const int nrows_in = nout + 6;

    for (int i = tid; i < DW2_NSTAT * DW2_MAXG * 8; i += NT) stats[i] = 0.f;
    if (tid == 0) {
        ptx::prefetch_tmap(&tmap_x);
        for (int s = 0; s < p.nst; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], NT / 32);
        }
        for (int s = 0; s < DW2_NSTAT; ++s) ptx::mbar_init(&stat_bar[s], NT / 32);
        ptx::fence_mbar_init();
    }
    __syncthreads();

    const int half = C >> 1;
    const int g = tid / p.gthreads;
    const int cp = tid - g * p.gthreads;
    const bool active = g < p.G && cp < half;
    const int c0 = active ? 2 * cp : 0;
    const int gs = active ? g : 0;
    const int bx = c0 / p.CB;
    const int coff = c0 - bx * p.CB;
    float2 wr[49];
#pragma unroll
    for (int k = 0; k < 49; ++k)
        wr[k] = active ? __ldg(reinterpret_cast<const float2*>(p.w49 + static_cast<size_t>(k) * C + c0))
                       : make_float2(0.f, 0.f);
    const float2 bs = active ? __ldg(reinterpret_cast<const float2*>(p.bias + c0)) : make_float2(0.f, 0.f);
    const float2 gam = active ? __ldg(reinterpret_cast<const float2*>(p.ln_w + c0)) : make_float2(0.f, 0.f);
    const float2 bet = active ? __ldg(reinterpret_cast<const float2*>(p.ln_b + c0)) : make_float2(0.f, 0.f);
    float2 acc[6][TW];
#pragma unroll
    for (int s = 0; s < 6; ++s)
#pragma unroll
        for (int t = 0; t < TW; ++t) acc[s][t] = bs;
    float2 held[2][TW];  // finished rows waiting for their statistics: [0] = previous row, [1] = the one before
#pragma unroll
    for (int t = 0; t < TW; ++t) held[0][t] = held[1][t] = make_float2(0.f, 0.f);

    const uint32_t in_off = static_cast<uint32_t>(bx * p.box_bytes + (gs * TW * p.CB + coff) * 4);
    const uint32_t tok_stride = static_cast<uint32_t>(p.CB * 4);
    const uint32_t in_ring_u32 = ptx::smem_u32(in_ring);
    const uint32_t stats_u32 = ptx::smem_u32(stats);
    const int h_end = h0 + nout;
    const float inv_c = 1.0f / static_cast<float>(C);
    const int ridx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);  // value this lane ends up with
    const int wtok0 = w0 + gs * TW;
    int stage = 0;
    uint32_t phase = 0;

    // Thread 0 is also the TMA producer: it refills every ring stage that all threads have released (non-blocking
    // test before each row; a blocking wait only if the row it is about to consume was never issued).
    const uint32_t tx = static_cast<uint32_t>(p.nbox) * (cols + 6) * p.CB * 4;
    int prod_r = 0, pstage = 0, issued = 0, consumed = 0;
    uint32_t pphase = 0;
    auto produce = [&]() {
        while (prod_r < nrows_in) {
            const int hi = h0 - 3 + prod_r;
            if (hi < 0 || hi >= p.H) { ++prod_r; continue; }
            if (issued <= consumed) ptx::mbar_wait(&empty_bar[pstage], pphase ^ 1);
            else if (!ptx::mbar_test(&empty_bar[pstage], pphase ^ 1)) break;
            ptx::mbar_arrive_expect_tx(&full_bar[pstage], tx);
            uint8_t* dst = in_ring + static_cast<size_t>(pstage) * p.stage_bytes;
            for (int bxi = 0; bxi < p.nbox; ++bxi)
                ptx::tma_load_4d(dst + bxi * p.box_bytes, &tmap_x, bxi * p.CB, w0 - 3, hi, b, &full_bar[pstage]);
            if (++pstage == p.nst) { pstage = 0; pphase ^= 1; }
            ++prod_r;
            ++issued;
        }
    };

    // Normalise and store output row `o` (relative to h0) from `v`, once its statistics are complete.
    auto finish_row = [&](int o, const float2 (&v)[TW]) {
        const int ss = o % DW2_NSTAT;
        ptx::mbar_wait(&stat_bar[ss], static_cast<uint32_t>((o / DW2_NSTAT) & 1));
        const uint32_t sp = stats_u32 + static_cast<uint32_t>((ss * DW2_MAXG + gs) * 8) * 4;
        float st[8];
        const float4 s0 = ptx::lds_f4(sp);
        st[0] = s0.x; st[1] = s0.y; st[2] = s0.z; st[3] = s0.w;
        if (TW > 2) {
            const float4 s1 = ptx::lds_f4(sp + 16);
            st[4] = s1.x; st[5] = s1.y; st[6] = s1.z; st[7] = s1.w;
        }
        const size_t orow = (static_cast<size_t>(b) * p.H + (h0 + o)) * p.W;
#pragma unroll
        for (int t = 0; t < TW; ++t) {
            const float mean = st[2 * t] * inv_c;
            const float var = fmaxf(st[2 * t + 1] * inv_c - mean * mean, 0.f);
            const float rstd = rsqrtf(var + p.eps);
            const int wo = wtok0 + t;
            if (active && wo < p.W) {
                const float y0 = (v[t].x - mean) * rstd * gam.x + bet.x;
                const float y1 = (v[t].y - mean) * rstd * gam.y + bet.y;
                const size_t off = (orow + wo) * C + c0;
                *reinterpret_cast<uint32_t*>(p.out_hi + off) = pack2(y0, y1);
                if (p.out_lo != nullptr)
                    *reinterpret_cast<uint32_t*>(p.out_lo + off) = pack2(y0 - bf16_round(y0), y1 - bf16_round(y1));
            }
        }
    };

    // acc[s] holds the partial sums of output row hi - 3 + s (s = 0..5) when input row hi is about to be processed.
    // Row hi adds filter row kh = 6 - s to it and the result moves to position s - 1 (the FMA chain reads acc[s] and
    // writes acc[s - 1], so the shift costs nothing); position 0 completes and leaves as `done`, a fresh row starts
    // from the bias at position 5. One copy of the 49*TW FMAs serves every row: the loop is not unrolled.
#pragma unroll 1
    for (int r = 0; r < nrows_in; ++r) {
        const int hi = h0 - 3 + r;
        float2 done[TW];
        if (hi >= 0 && hi < p.H) {
            if (tid == 0) produce();
            ptx::mbar_wait(&full_bar[stage], phase);
            const uint32_t src = in_ring_u32 + static_cast<uint32_t>(stage) * p.stage_bytes + in_off;
            float2 in[TW + 6];
#pragma unroll
            for (int i = 0; i < TW + 6; ++i) in[i] = ptx::lds_f2(src + i * tok_stride);
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&empty_bar[stage]);  // one arrival per warp
            ++consumed;
            if (++stage == p.nst) { stage = 0; phase ^= 1; }
#pragma unroll
            for (int s = 0; s < 7; ++s) {
                const int kh = 6 - s;
                const int ho = hi - 3 + s;
                if (ho >= h0 && ho < h_end) {
#pragma unroll
                    for (int t = 0; t < TW; ++t) {
                        float2 a = (s == 6) ? bs : acc[s < 6 ? s : 0][t];
#pragma unroll
                        for (int kw = 0; kw < 7; ++kw) a = ffma2(in[t + kw], wr[kh * 7 + kw], a);
                        if (s == 0) done[t] = a;
                        else acc[s - 1][t] = a;
                    }
                } else if (s == 0) {
#pragma unroll
                    for (int t = 0; t < TW; ++t) done[t] = acc[0][t];
                }
            }
        } else {
#pragma unroll
            for (int t = 0; t < TW; ++t) {
                done[t] = acc[0][t];
#pragma unroll
                for (int s = 1; s < 6; ++s) acc[s - 1][t] = acc[s][t];
                acc[5][t] = bs;
            }
        }
        if (r >= 6) {
            const int o = r - 6;  // output row (relative to h0) that just completed
            float part[8];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                if (t < TW) {
                    part[2 * t] = done[t].x + done[t].y;
                    part[2 * t + 1] = fmaf(done[t].x, done[t].x, done[t].y * done[t].y);
                } else {
                    part[2 * t] = 0.f;
                    part[2 * t + 1] = 0.f;
                }
            }
            const float d = warp_reduce8(part, lane);
            const int ss = o % DW2_NSTAT;
            if ((lane & 3) == 0 && ridx < 2 * TW) atomicAdd(&stats[(ss * DW2_MAXG + gs) * 8 + ridx], d);
            // Recycle the slot of row o + 3 (last used by row o - 13). Safe: this thread has passed the barrier of
            // row o - 3 (finish_row, previous iteration), so every thread completed iteration o - 4 and with it all
            // reads of rows <= o - 6. Ordered: a thread adds to row o + 3 only after it waited for the barrier of
            // row o (finish_row in its iteration o + 2), which includes the arrival just below.
            if (tid < DW2_MAXG * 8) stats[((o + 3) % DW2_NSTAT) * DW2_MAXG * 8 + tid] = 0.f;
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&stat_bar[ss]);
            if (o >= 2) finish_row(o - 2, held[1]);
#pragma unroll
            for (int t = 0; t < TW; ++t) {
                held[1][t] = held[0][t];
                held[0][t] = done[t];
            }
        }
    }
    if (nout >= 2) finish_row(nout - 2, held[1]);
    finish_row(nout - 1, held[0]);
}

}  // namespace cpn
