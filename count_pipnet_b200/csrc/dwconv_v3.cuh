// Depthwise 7x7 conv (pad 3) + bias of a CNBlock (torchvision/models/convnext.py:52), third generation: lanes over
// CHANNELS, warps over SPACE, every FMA a useful one.
//
//   x [B,H,W,C] fp32 (residual stream)  ->  hi [B*H*W, C] bf16 = round(conv + bias)             (GEMM A operand)
//                                           lo [B*H*W, C] bf16 = round(v - hi)                   (fp32 mode only)
//                                           stats [B*H*W, C/32, 2] fp32 = (sum, sum of squares) over each 32-channel
//                                           group of the values the GEMM will see (hi, or hi + lo)
//
// Why: dwconv7_stats_kernel (lane = channel pair, 98 tap registers, 192 registers per thread, 2 warps per scheduler) keeps
// the FMA pipe 48 % busy and spends a fifth of its FMAs on warm-up rows; tools/micro/dw3_bench.cu measured 78-90 % of the
// scalar-FFMA peak (26-30 of 33.4 TFMA/s) for the loop below at 14-16 warps per SM and <= 128 registers.
//   * an item = (32-channel group, image, tile of NB = 2 row blocks x NS <= 7 column strips); a WARP owns one unit = a
//     block of 7 output rows x 4 output columns, a LANE one of the 32 channels: 49 taps + 28 accumulators + 10 inputs in
//     registers, scalar FFMA. The 13 input rows of a block are fully unrolled with only the valid (input row, filter row)
//     pairs: no warm-up rows;
//   * the input of an item arrives with ONE TMA box [rows][columns][32 channels] (zero-filled outside the image) into a
//     two-deep ring; pixel stride 128 B, so the 32 lanes of an LDS.32 hit the 32 banks;
//   * 16 warps (4 per scheduler, the whole register file at 128 registers) for the <= 14 units of an item: the unit ->
//     warp assignment rotates by the number of spare warps per item, so every scheduler carries the same load over two
//     items (14 fixed warps + a producer warp left two schedulers 25 % idle: profiles/r02_dwconv_v3.md). The first spare
//     warp of an item is that item's producer: it waits until the slot of the item after next is released and issues
//     its box;
//   * a warp releases its ring slot after its last input read and then writes its 28 x 32 results through a PRIVATE
//     shared-memory scratch in two rounds of 14 pixels (row stride 36 floats: conflict-free both ways): lane L < 28
//     reads back 16 channels of pixel L / 2, rounds them to bf16, stores 32 contiguous bytes and sums the rounded values
//     and their squares; the two lanes of a pixel add their sums (the LayerNorm statistics, same format as the other
//     depthwise kernels). Only __syncwarp is involved: the warps of a CTA drift apart, so one warp's output phase
//     overlaps the others' FMAs -- the first version of this kernel (out tile shared by the CTA, two __syncthreads per
//     item) lost 40 % of its time there and was slower than the row-marching kernel;
//   * CTAs are persistent over contiguous item ranges, items ordered (channel group, image, row tile, column tile), so
//     the taps are loaded once or twice per CTA.
#pragma once
#include "device_utils.cuh"

namespace cpn {

constexpr int DW3_RB = 7;         // output rows per unit
constexpr int DW3_TW = 4;         // output columns per unit
constexpr int DW3_NB = 2;         // row blocks per tile
constexpr int DW3_MAX_NS = 7;     // column strips per tile
constexpr int DW3_SROW = 36;      // floats per pixel row of a warp's scratch (32 channels + 4: bank rotation)
constexpr int DW3_WARPS = 16;      // 4 per scheduler
constexpr int DW3_HALF = DW3_RB * DW3_TW / 2;            // pixels per output round
constexpr int DW3_BROW = 40;       // bf16 elements per pixel row of the bf16-mode scratch (32 channels + 8: 80-byte stride)
constexpr int DW3_SCRATCH = DW3_RB * DW3_TW * DW3_BROW / 2;  // floats per warp: 28 x 80 B (>= the 14 x 144 B of fp32 mode)
static_assert(DW3_SCRATCH >= DW3_HALF * DW3_SROW, "the scratch serves both modes");

struct Dw3Params {
    const float* w49;   // [49, C]
    const float* bias;  // [C]
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;  // nullptr in bf16 mode
    float* stats;           // [B*H*W, C/32, 2]
    int B, H, W, C;
    int NS;           // column strips per tile; units per item = NS * DW3_NB <= 14
    int wbox;         // staged columns per row = 4 NS + 6
    int rows_in;      // DW3_NB * 7 + 6
    int nct, nrt;     // column / row tiles per image
    int nitems;       // (C / 32) * B * nrt * nct
    int stage_bytes;  // bytes of one ring slot = bytes of one TMA box (multiple of 128)
};

struct Dw3Item {
    int ct, rt, b, grp;
};
__device__ __forceinline__ Dw3Item dw3_decode(const Dw3Params& p, int item) {
    Dw3Item t;
    t.ct = item % p.nct;
    int rem = item / p.nct;
    t.rt = rem % p.nrt;
    rem /= p.nrt;
    t.b = rem % p.B;
    t.grp = rem / p.B;
    return t;
}
__device__ __forceinline__ void dw3_next(const Dw3Params& p, Dw3Item& t) {
    if (++t.ct == p.nct) {
        t.ct = 0;
        if (++t.rt == p.nrt) {
            t.rt = 0;
            if (++t.b == p.B) { t.b = 0; ++t.grp; }
        }
    }
}

__device__ __forceinline__ void dw3_st256(void* dst, const uint32_t (&w)[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}

template <bool SPLIT>
__global__ void __launch_bounds__(32 * DW3_WARPS, 1)
dwconv7_v3_kernel(const __grid_constant__ CUtensorMap tmap_x, const Dw3Params p) {
    constexpr int RB = DW3_RB, TW = DW3_TW, HP = DW3_HALF;
    // no integer round trip on the base pointer: the compiler must keep seeing shared memory (LDS / STS, not generic LD / ST)
    extern __shared__ __align__(128) float dw3_smem[];
    const int stage_floats = p.stage_bytes >> 2;
    const int nunits = p.NS * DW3_NB;
    const int spare = DW3_WARPS - nunits;  // >= 2
    float* ring = dw3_smem;
    float* scratch = dw3_smem + 2 * stage_floats;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(scratch + DW3_WARPS * DW3_SCRATCH);
    uint64_t* empty_bar = full_bar + 2;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j_lo = static_cast<int>(static_cast<long long>(blockIdx.x) * p.nitems / gridDim.x);
    const int j_hi = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * p.nitems / gridDim.x);
    const int n = j_hi - j_lo;

    if (threadIdx.x == 0) {
        ptx::prefetch_tmap(&tmap_x);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], nunits);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();

    auto issue = [&](int j) {  // one thread: the box of item j into slot j & 1
        const Dw3Item it = dw3_decode(p, j_lo + j);
        const int slot = j & 1;
        if (j >= 2) ptx::mbar_wait(&empty_bar[slot], ((j >> 1) - 1) & 1);
        ptx::mbar_arrive_expect_tx(&full_bar[slot], static_cast<uint32_t>(p.stage_bytes));
        ptx::tma_load_4d(ring + slot * stage_floats, &tmap_x, it.grp * 32, it.ct * p.NS * TW - 3, it.rt * DW3_NB * RB - 3,
                         it.b, &full_bar[slot]);
    };
    if (threadIdx.x == 0 && n > 0) issue(0);

    const int in_row = p.wbox * 32;
    float* my = scratch + warp * DW3_SCRATCH;
    const int npart = p.C >> 5;
    // output role of lane L < 28: channels 16 (L & 1) .. + 15 of pixel L / 2 of the round
    const int pl = lane >> 1, hf = lane & 1;
    const int plc = pl < DW3_HALF ? pl : 0;  // lanes 28-31 read a valid row and store nothing

    float wr[49];
    float bs = 0.f;
    int cur_grp = -1;
    Dw3Item t = dw3_decode(p, j_lo);
    int rot = warp;  // unit of this warp for item j = (warp + spare * j) mod 16

    for (int j = 0; j < n; ++j, rot = (rot + spare) & (DW3_WARPS - 1), dw3_next(p, t)) {
        const int u = rot;
        if (u >= nunits) {
            // spare warp of this item; the first one loads the next item's box
            if (u == nunits && lane == 0 && j + 1 < n) issue(j + 1);
            continue;
        }
        if (t.grp != cur_grp) {
            cur_grp = t.grp;
#pragma unroll
            for (int i = 0; i < 49; ++i) wr[i] = __ldg(p.w49 + i * p.C + t.grp * 32 + lane);
            bs = __ldg(p.bias + t.grp * 32 + lane);
        }
        const int s = u % p.NS, k = u / p.NS;
        const int slot = j & 1;
        ptx::mbar_wait(&full_bar[slot], (j >> 1) & 1);

        // ------------------------------------------------------------ 7 x 4 outputs of one channel
        float acc[RB][TW];
#pragma unroll
        for (int o = 0; o < RB; ++o)
#pragma unroll
            for (int tt = 0; tt < TW; ++tt) acc[o][tt] = bs;
        {
            const float* src = ring + slot * stage_floats + (k * RB * p.wbox + TW * s) * 32 + lane;
#pragma unroll
            for (int i = 0; i < RB + 6; ++i) {
                float xv[TW + 6];
#pragma unroll
                for (int q = 0; q < TW + 6; ++q) xv[q] = src[q * 32];
                src += in_row;
#pragma unroll
                for (int kh = 0; kh < 7; ++kh) {
                    const int o = i - kh;
                    if (o < 0 || o >= RB) continue;
#pragma unroll
                    for (int kw = 0; kw < 7; ++kw)
#pragma unroll
                        for (int tt = 0; tt < TW; ++tt) acc[o][tt] = fmaf(xv[tt + kw], wr[kh * 7 + kw], acc[o][tt]);
                }
            }
        }
        __syncwarp();  // every lane has its inputs in registers (and has finished reading the scratch of its last item)
        if (lane == 0) ptx::mbar_arrive(&empty_bar[slot]);

        const int h_base = (t.rt * DW3_NB + k) * RB, w_base = (t.ct * p.NS + s) * TW;
        if (!SPLIT) {
            // -------------------------------------------------------- bf16 mode: ONE round through a bf16 scratch
            // lane = channel rounds its 28 values and parks them as [pixel][32 x bf16] (pixel stride 80 B: the 16-byte
            // reads of consecutive pixels rotate through the banks); lane L < 28 then owns pixel L: 64 contiguous bytes,
            // already packed, and the sums over its 32 channels. Bit-identical to the two-round fp32 path below.
            __nv_bfloat16* myb = reinterpret_cast<__nv_bfloat16*>(my);
#pragma unroll
            for (int q = 0; q < RB * TW; ++q) myb[q * DW3_BROW + lane] = __float2bfloat16_rn(acc[q / TW][q % TW]);
            __syncwarp();
            const int h = h_base + (lane >> 2), w = w_base + (lane & 3);
            if (lane < RB * TW && h < p.H && w < p.W) {
                uint32_t hw[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint4 f = *reinterpret_cast<const uint4*>(myb + lane * DW3_BROW + 8 * q);
                    hw[4 * q] = f.x; hw[4 * q + 1] = f.y; hw[4 * q + 2] = f.z; hw[4 * q + 3] = f.w;
                }
                // the summation order of the two-round path: two chains per 16-channel half, halves added last
                float sum[4] = {0.f, 0.f, 0.f, 0.f}, sq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const float r0 = __uint_as_float(hw[q] << 16), r1 = __uint_as_float(hw[q] & 0xffff0000u);
                    const int c4 = 2 * (q >> 3) + (q & 1);
                    sum[c4] += r0 + r1;
                    sq[c4] = fmaf(r0, r0, sq[c4]);
                    sq[c4] = fmaf(r1, r1, sq[c4]);
                }
                const size_t pix = (static_cast<size_t>(t.b) * p.H + h) * p.W + w;
                __nv_bfloat16* dst = p.out_hi + pix * p.C + t.grp * 32;
                const uint32_t(&h0)[8] = *reinterpret_cast<const uint32_t(*)[8]>(&hw[0]);
                const uint32_t(&h1)[8] = *reinterpret_cast<const uint32_t(*)[8]>(&hw[8]);
                dw3_st256(dst, h0);
                dw3_st256(dst + 16, h1);
                reinterpret_cast<float2*>(p.stats)[pix * npart + t.grp] =
                    make_float2((sum[0] + sum[1]) + (sum[2] + sum[3]), (sq[0] + sq[1]) + (sq[2] + sq[3]));
            }
            continue;
        }
        // ------------------------------------------------------------ fp32 mode: transpose through the scratch, 2 x 14 pixels
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            if (r == 1) __syncwarp();  // round 0 has been read
#pragma unroll
            for (int q = 0; q < HP; ++q) my[q * DW3_SROW + lane] = acc[(r * HP + q) / TW][(r * HP + q) % TW];
            __syncwarp();
            const int pi = r * HP + pl;
            const int h = h_base + (pi >> 2), w = w_base + (pi & 3);
            const bool live = lane < 2 * HP && h < p.H && w < p.W;
            float v[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 f = *reinterpret_cast<const float4*>(my + plc * DW3_SROW + hf * 16 + 4 * q);
                v[4 * q] = f.x; v[4 * q + 1] = f.y; v[4 * q + 2] = f.z; v[4 * q + 3] = f.w;
            }
            uint32_t hw[8], lw[8];
            float sum[2] = {0.f, 0.f}, sq[2] = {0.f, 0.f};  // two chains, fixed order
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                hw[q] = pack_bf16x2(v[2 * q], v[2 * q + 1]);
                float r0 = __uint_as_float(hw[q] << 16), r1 = __uint_as_float(hw[q] & 0xffff0000u);
                if (SPLIT) {
                    lw[q] = pack_bf16x2(v[2 * q] - r0, v[2 * q + 1] - r1);
                    r0 += __uint_as_float(lw[q] << 16);  // what the three-pass GEMM sees: hi + lo
                    r1 += __uint_as_float(lw[q] & 0xffff0000u);
                }
                sum[q & 1] += r0 + r1;
                sq[q & 1] = fmaf(r0, r0, sq[q & 1]);
                sq[q & 1] = fmaf(r1, r1, sq[q & 1]);
            }
            float ssum = sum[0] + sum[1], ssq = sq[0] + sq[1];
            ssum += __shfl_xor_sync(0xffffffffu, ssum, 1);  // lower 16 channels + upper 16 channels (commutative: same bits)
            ssq += __shfl_xor_sync(0xffffffffu, ssq, 1);
            if (live) {
                const size_t pix = (static_cast<size_t>(t.b) * p.H + h) * p.W + w;
                // one 256-bit store per lane: a whole 32-byte sector per request (two 128-bit stores would each write half
                // of two sectors)
                dw3_st256(p.out_hi + pix * p.C + t.grp * 32 + hf * 16, hw);
                if (SPLIT) dw3_st256(p.out_lo + pix * p.C + t.grp * 32 + hf * 16, lw);
                if (hf == 0) reinterpret_cast<float2*>(p.stats)[pix * npart + t.grp] = make_float2(ssum, ssq);
            }
        }
    }
}

}  // namespace cpn
