// Depthwise 7x7 conv (pad 3) + bias of a CNBlock (torchvision/models/convnext.py:52) on the TENSOR cores.
//
//   x [B,H,W,C] fp32 (residual stream)  ->  hi [B*H*W, C] bf16 = round(conv + bias)             (GEMM A operand)
//                                           lo [B*H*W, C] bf16 = round(v - hi)                   (fp32 mode only)
//                                           stats [B*H*W, C/32, 2] fp32 = (sum, sum of squares) over each 32-channel
//                                           group of the values the GEMM will see (hi, or hi + lo)
//
// Why: on the FMA pipe the conv costs 49 FMAs per output and both CUDA-core kernels (dwconv_stats.cuh: taps in registers,
// 8 warps per SM; dwconv_v2.cuh: taps in shared memory) end at 25-36 % of that pipe (profiles/r02_dwconv.md); it was
// 31 % of the benchmark step. Per channel the conv is a banded-Toeplitz product, and the warp-level tensor-core path
// (mma.sync m16n8k16, measured 550 TFLOP/s on B200: tools/micro/mma_bench.cu) does it with 10x fewer instructions:
//
//   D[m][n] += A[m][k] * B_kh[k][n]      m = strip (image, 8 output columns), n = column inside the strip,
//                                         k = input column inside the strip's 16-column window (8 s - 3 + k),
//                                         B_kh[k][n] = w[kh][k - n] for 0 <= k - n <= 6, else 0 (same for every strip)
//
// one MMA per (channel, filter row kh, 16 strips): 128 outputs x 7 taps. A warp owns two channels of the CTA's 32-channel
// group: their B fragments (7 filter rows x 2 registers) stay in registers, and it marches down the input rows with the 7
// in-flight output rows of both channels in 56 accumulator registers, exactly like the CUDA-core kernels.
// Precision: the fp32 input is split x = hi + lo (two bf16 planes, 16 mantissa bits) and both planes are multiplied
// (x is the residual stream; rounding it to bf16 alone costs 8e-3 on the pooled scores, measured with the oracle); the
// taps are bf16 in bf16 mode (2 MMAs per filter row) and hi + lo in fp32 mode (3 MMAs: hi*hi + lo*hi + hi*lo), fp32
// accumulation throughout.
//
// Data path per CTA (16 warps, one CTA per SM, persistent over jobs = (32-channel group, image group, column block, row
// segment)):
//   TMA ring     : input rows [columns x 32 channels] fp32, 128-byte swizzle, zero-filled outside the map (thread 0 issues)
//   convert      : all warps, lane = channel: fp32 row -> per-channel bf16 planes (hi, lo) [channel][image][column pair]
//                  (the layout the A fragments read with one 32-bit load), double-buffered
//   MMA          : per warp, its two channels, filter rows guarded so that row segments cost no extra MMAs
//   epilogue     : a completed output row goes to a staging tile [channel][strip][column pair] (bf16 pairs, + lo plane)
//   store        : warp w takes strip w: lane = channel, so the 32-channel LayerNorm partial sums are a fixed-order warp
//                  butterfly and the rows go out as 64-byte channel-contiguous segments
//   ONE block barrier per input row separates the phases (planes and staging are double-buffered).
#pragma once
#include <type_traits>

#include "device_utils.cuh"

namespace cpn {

constexpr int DWM_WARPS = 16;
constexpr int DWM_THREADS = 32 * DWM_WARPS;
constexpr int DWM_NSTAGE = 3;  // raw input rows in flight
constexpr int DWM_SCS = 65;    // staging words per channel (64 + 1: conflict-free across channels)

struct DwmParams {
    const float* w49;   // [49, C]
    const float* bias;  // [C]
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;  // nullptr in bf16 mode
    float* stats;           // [B*H*W, C/32, 2]
    int B, H, W, C;
    int sb;            // strips (8 output columns) per image and column block, <= 16
    int ncb;           // column blocks per row
    int nimg;          // images per job = 16 / sb
    int nip;           // image groups = ceil(B / nimg)
    int rows_per_seg;  // output rows per row segment
    int nseg;
    int njobs;         // (C / 32) * nip * ncb * nseg
    int rw;            // words (column pairs) per image row in a plane = 4 sb + 4
    int cs;            // plane words per channel = nimg * rw rounded up to odd
    int box_tx;        // bytes one TMA box delivers = (8 sb + 8) * 128
    int box_stride;    // bytes between the boxes of one stage (multiple of 1024)
    int stage_bytes;   // nimg * box_stride
};

struct DwmJob {
    int grp, ip, cb, h0, nout;
};
__device__ __forceinline__ DwmJob dwm_decode(const DwmParams& p, int job) {
    DwmJob j;
    const int seg = job % p.nseg;
    int rem = job / p.nseg;
    j.cb = rem % p.ncb;
    rem /= p.ncb;
    j.ip = rem % p.nip;
    j.grp = rem / p.nip;
    j.h0 = seg * p.rows_per_seg;
    j.nout = min(p.rows_per_seg, p.H - j.h0);
    return j;
}

// D (fp32, 16x8) += A (bf16, 16x16, row) * B (bf16, 16x8, col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

// Shared memory carve-up (bytes from the 1024-aligned base)
struct DwmSmem {
    int raw, planes, staging, blo, bars, total;
};
__host__ __device__ inline DwmSmem dwm_smem(const DwmParams& p, bool split) {
    DwmSmem s;
    s.raw = 0;
    s.planes = DWM_NSTAGE * p.stage_bytes;                 // [2 buffers][hi, lo][32 channels][cs] words
    s.staging = s.planes + 2 * 2 * 32 * p.cs * 4;          // [2 buffers][hi, lo][32 channels][DWM_SCS] words
    s.blo = s.staging + 2 * 2 * 32 * DWM_SCS * 4;          // fp32 mode: lo planes of the B fragments
    s.bars = s.blo + (split ? 32 * 7 * 2 * 32 * 4 : 0);    // [channel][kh][reg][lane] words
    s.total = s.bars + DWM_NSTAGE * 8 + 64;
    return s;
}

template <bool SPLIT>
__global__ void __launch_bounds__(DWM_THREADS, 1)
dwconv7_mma_kernel(const __grid_constant__ CUtensorMap tmap_x, const DwmParams p) {
    constexpr int NS = DWM_NSTAGE;
    extern __shared__ uint8_t dwm_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dwm_smem_raw) + 1023) & ~uintptr_t(1023));
    const DwmSmem L = dwm_smem(p, SPLIT);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bars);
    const uint32_t raw0 = ptx::smem_u32(smem + L.raw);
    const uint32_t planes0 = ptx::smem_u32(smem + L.planes);
    const uint32_t staging0 = ptx::smem_u32(smem + L.staging);
    const uint32_t blo0 = ptx::smem_u32(smem + L.blo);
    const uint32_t plane_bytes = static_cast<uint32_t>(32 * p.cs * 4);   // one plane (hi or lo) of one buffer
    constexpr uint32_t STG_BYTES = 32 * DWM_SCS * 4;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int job_lo = static_cast<int>(static_cast<long long>(blockIdx.x) * p.njobs / gridDim.x);
    const int job_hi = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * p.njobs / gridDim.x);
    const int nstrips = p.nimg * p.sb;  // valid strips (<= 16)
    const int NPART = p.C >> 5;

    if (threadIdx.x == 0) {
        ptx::prefetch_tmap(&tmap_x);
        for (int s = 0; s < NS; ++s) ptx::mbar_init(&full_bar[s], 1);
        ptx::fence_mbar_init();
    }
    __syncthreads();
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();

    // ---- producer state (thread 0 only): the stream of in-image input rows over this CTA's jobs
    int pj = job_lo, phr = 0, phr_end = 0, pstage = 0;
    bool pinit = false;
    DwmJob pjb;
    auto produce_next = [&]() {
        // issues the TMA loads of the next input row of the stream into ring slot pstage (known to be free)
        for (;;) {
            if (!pinit) {
                if (pj >= job_hi) return;
                pjb = dwm_decode(p, pj);
                phr = max(pjb.h0 - 3, 0);
                phr_end = min(pjb.h0 + pjb.nout + 3, p.H);
                pinit = true;
            }
            if (phr < phr_end) break;
            ++pj;
            pinit = false;
        }
        ptx::mbar_arrive_expect_tx(&full_bar[pstage], static_cast<uint32_t>(p.nimg * p.box_tx));
        uint8_t* dst = smem + L.raw + pstage * p.stage_bytes;
        const int col0 = pjb.cb * p.sb * 8 - 3;
        for (int i = 0; i < p.nimg; ++i)  // images beyond B and columns outside the map arrive as zeros
            ptx::tma_load_4d(dst + i * p.box_stride, &tmap_x, pjb.grp * 32, col0, phr, pjb.ip * p.nimg + i, &full_bar[pstage]);
        ++phr;
        if (++pstage == NS) pstage = 0;
    };
    if (threadIdx.x == 0)
        for (int s = 0; s < NS; ++s) produce_next();

    // ---- consumer state
    int cstage = 0;
    uint32_t cphase = 0;
    // A-fragment word offsets of this lane inside a channel's plane: strips m = g and g + 8 (idle strips are clamped to a
    // valid one: their results are never stored)
    const int m0 = min(g, nstrips - 1), m1 = min(g + 8, nstrips - 1);
    const uint32_t ao0 = static_cast<uint32_t>(((m0 / p.sb) * p.rw + 4 * (m0 % p.sb) + t) * 4);
    const uint32_t ao1 = static_cast<uint32_t>(((m1 / p.sb) * p.rw + 4 * (m1 % p.sb) + t) * 4);
    const uint32_t ch_off[2] = {static_cast<uint32_t>((2 * warp) * p.cs * 4), static_cast<uint32_t>((2 * warp + 1) * p.cs * 4)};

    uint32_t bhi[2][7][2];  // B fragments of this warp's two channels
    float bias2[2] = {0.f, 0.f};
    int cur_grp = -1;

    // fp32 row in ring slot `stage` -> bf16 planes (hi, lo) of buffer `buf`; lane = channel
    auto convert_row = [&](int buf) {
        ptx::mbar_wait(&full_bar[cstage], cphase);
        const uint32_t rbase = raw0 + static_cast<uint32_t>(cstage * p.stage_bytes);
        const uint32_t phi = planes0 + static_cast<uint32_t>(buf) * 2u * plane_bytes + static_cast<uint32_t>(lane * p.cs * 4);
        const uint32_t plo = phi + plane_bytes;
        const int items = p.nimg * p.rw;
        const uint32_t cchunk = static_cast<uint32_t>(lane >> 2), cin = static_cast<uint32_t>((lane & 3) * 4);
        for (int it = warp; it < items; it += DWM_WARPS) {
            const int i = it / p.rw, pp = it - i * p.rw;
            const uint32_t q = static_cast<uint32_t>(2 * pp);
            const uint32_t b0 = rbase + static_cast<uint32_t>(i * p.box_stride) + q * 128u;
            // TMA 128-byte swizzle: 16-byte chunk index XOR (row-of-128-bytes index & 7); rows q and q + 1
            const float x0 = lds_f32(b0 + (((cchunk ^ (q & 7u)) << 4) | cin));
            const float x1 = lds_f32(b0 + 128u + (((cchunk ^ ((q + 1u) & 7u)) << 4) | cin));
            const uint32_t h = pack_bf16x2(x0, x1);
            const uint32_t l = pack_bf16x2(x0 - __uint_as_float(h << 16), x1 - __uint_as_float(h & 0xffff0000u));
            const uint32_t o = static_cast<uint32_t>((i * p.rw + pp) * 4);
            sts_u32(phi + o, h);
            sts_u32(plo + o, l);
        }
        if (++cstage == NS) { cstage = 0; cphase ^= 1; }
    };

    for (int job = job_lo; job < job_hi; ++job) {
        const DwmJob jb = dwm_decode(p, job);
        if (jb.grp != cur_grp) {
            // B fragments of the new channel group: b0 = B[2t, 2t+1][n = g], b1 = B[2t+8, 2t+9][g], B[k][n] = w[kh][k - n]
            cur_grp = jb.grp;
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                const int c = jb.grp * 32 + 2 * warp + ch;
                bias2[ch] = __ldg(p.bias + c);
#pragma unroll
                for (int kh = 0; kh < 7; ++kh) {
                    float wv[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int k = 2 * t + (e & 1) + (e >> 1) * 8;
                        const int kw = k - g;
                        wv[e] = (kw >= 0 && kw < 7) ? __ldg(p.w49 + (kh * 7 + kw) * p.C + c) : 0.f;
                    }
                    const uint32_t h0 = pack_bf16x2(wv[0], wv[1]), h1 = pack_bf16x2(wv[2], wv[3]);
                    bhi[ch][kh][0] = h0;
                    bhi[ch][kh][1] = h1;
                    if (SPLIT) {
                        const uint32_t l0 = pack_bf16x2(wv[0] - __uint_as_float(h0 << 16), wv[1] - __uint_as_float(h0 & 0xffff0000u));
                        const uint32_t l1 = pack_bf16x2(wv[2] - __uint_as_float(h1 << 16), wv[3] - __uint_as_float(h1 & 0xffff0000u));
                        // only this warp reads these words back, and always after the block barrier of the first row
                        const uint32_t a = blo0 + static_cast<uint32_t>((((2 * warp + ch) * 7 + kh) * 2) * 32 + lane) * 4u;
                        sts_u32(a, l0);
                        sts_u32(a + 128u, l1);
                    }
                }
            }
        }
        // output geometry of the strip this warp stores (store phase): strip index = warp
        const int simg = warp / p.sb, sstr = warp - simg * p.sb;
        const int sb_img = jb.ip * p.nimg + simg;
        const bool strip_ok = warp < nstrips && sb_img < p.B;
        const int scol = (jb.cb * p.sb + sstr) * 8;
        const int total_r = jb.nout + 6;  // input rows h0 - 3 .. h0 + nout + 2, inside or outside the image

        float acc[7][2][4];

        // ---- one finished output row (staging buffer sbuf, segment-local row o): stats + coalesced stores of strip `warp`
        auto store_row = [&](int sbuf, int o) {
            if (!strip_ok) return;  // warp-uniform
            const uint32_t shi = staging0 + static_cast<uint32_t>(sbuf) * 2u * STG_BYTES + static_cast<uint32_t>(lane * DWM_SCS * 4) +
                                 static_cast<uint32_t>(warp * 16);
            uint32_t wh[4], wl[4];
            float a[16];
#pragma unroll
            for (int tt = 0; tt < 4; ++tt) {
                wh[tt] = lds_u32(shi + tt * 4);
                float v0 = __uint_as_float(wh[tt] << 16), v1 = __uint_as_float(wh[tt] & 0xffff0000u);
                if (SPLIT) {
                    wl[tt] = lds_u32(shi + STG_BYTES + tt * 4);
                    v0 += __uint_as_float(wl[tt] << 16);
                    v1 += __uint_as_float(wl[tt] & 0xffff0000u);
                }
                a[4 * tt] = v0;
                a[4 * tt + 1] = v0 * v0;
                a[4 * tt + 2] = v1;
                a[4 * tt + 3] = v1 * v1;
            }
            // transposing butterfly over the 32 channels: afterwards lane l holds the total of value l >> 1
            // (value 2 * pixel + which, pixel = column inside the strip); fixed order -> run-to-run identical
            const bool b16 = (lane & 16) != 0, b8 = (lane & 8) != 0, b4 = (lane & 4) != 0, b2 = (lane & 2) != 0;
            float r8[8], r4[4], r2[2];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float send = b16 ? a[i] : a[i + 8], keep = b16 ? a[i + 8] : a[i];
                r8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float send = b8 ? r8[i] : r8[i + 4], keep = b8 ? r8[i + 4] : r8[i];
                r4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const float send = b4 ? r4[i] : r4[i + 2], keep = b4 ? r4[i + 2] : r4[i];
                r2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            float d = (b2 ? r2[1] : r2[0]) + __shfl_xor_sync(0xffffffffu, b2 ? r2[0] : r2[1], 2);
            d += __shfl_xor_sync(0xffffffffu, d, 1);
            const size_t pix0 = (static_cast<size_t>(sb_img) * p.H + (jb.h0 + o)) * p.W + scol;
            {
                const int v = lane >> 1, px = v >> 1;
                if ((lane & 1) == 0 && scol + px < p.W) p.stats[((pix0 + px) * NPART + jb.grp) * 2 + (v & 1)] = d;
            }
            // pair exchange: even lanes end up with (channel, channel + 1) of the even column, odd lanes with
            // (channel - 1, channel) of the odd column -> 32-bit stores, 64 contiguous bytes per column
            const bool odd = (lane & 1) != 0;
#pragma unroll
            for (int tt = 0; tt < 4; ++tt) {
                const int col = scol + 2 * tt + (odd ? 1 : 0);
                const size_t off = (pix0 + 2 * tt + (odd ? 1 : 0)) * p.C + jb.grp * 32 + (lane & ~1);
                const uint32_t oh = __shfl_xor_sync(0xffffffffu, wh[tt], 1);
                const uint32_t outh = odd ? __byte_perm(oh, wh[tt], 0x7632) : __byte_perm(wh[tt], oh, 0x5410);
                if (col < p.W) *reinterpret_cast<uint32_t*>(p.out_hi + off) = outh;
                if (SPLIT) {
                    const uint32_t ol = __shfl_xor_sync(0xffffffffu, wl[tt], 1);
                    const uint32_t outl = odd ? __byte_perm(ol, wl[tt], 0x7632) : __byte_perm(wl[tt], ol, 0x5410);
                    if (col < p.W) *reinterpret_cast<uint32_t*>(p.out_lo + off) = outl;
                }
            }
        };

        // ---- one input row r (J = accumulator slot the row opens; filter row kh accumulates into slot (J - kh) mod 7,
        // slot (J + 1) mod 7 holds the output row r - 6 that this row completes)
        auto do_row = [&](auto jc, const int r) {
            constexpr int j = decltype(jc)::value;
            const int hr = jb.h0 - 3 + r;
            const int buf = r & 1;
#pragma unroll
            for (int ch = 0; ch < 2; ++ch)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[j][ch][e] = bias2[ch];  // bias is the first addend of output row r
            if (hr >= 0 && hr < p.H) {
                const uint32_t pb = planes0 + static_cast<uint32_t>(buf) * 2u * plane_bytes;
                uint32_t ahi[2][4], alo[2][4];
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    const uint32_t b = pb + ch_off[ch];
                    ahi[ch][0] = lds_u32(b + ao0);
                    ahi[ch][1] = lds_u32(b + ao1);
                    ahi[ch][2] = lds_u32(b + ao0 + 16);
                    ahi[ch][3] = lds_u32(b + ao1 + 16);
                    alo[ch][0] = lds_u32(b + plane_bytes + ao0);
                    alo[ch][1] = lds_u32(b + plane_bytes + ao1);
                    alo[ch][2] = lds_u32(b + plane_bytes + ao0 + 16);
                    alo[ch][3] = lds_u32(b + plane_bytes + ao1 + 16);
                }
                // filter row kh feeds output row r - kh, which must lie in [0, nout)
                const int klo = max(0, r - jb.nout + 1), khi = min(6, r);
#pragma unroll
                for (int kh = 0; kh < 7; ++kh) {
                    if (kh >= klo && kh <= khi) {
                        const int s = (j - kh + 7) % 7;
#pragma unroll
                        for (int ch = 0; ch < 2; ++ch) {
                            mma_bf16_16816(acc[s][ch], ahi[ch], bhi[ch][kh][0], bhi[ch][kh][1]);
                            mma_bf16_16816(acc[s][ch], alo[ch], bhi[ch][kh][0], bhi[ch][kh][1]);
                            if (SPLIT) {
                                const uint32_t a = blo0 + static_cast<uint32_t>((((2 * warp + ch) * 7 + kh) * 2) * 32 + lane) * 4u;
                                mma_bf16_16816(acc[s][ch], ahi[ch], lds_u32(a), lds_u32(a + 128u));
                            }
                        }
                    }
                }
            }
            if (r >= 6) {
                // output row r - 6 is complete: this warp's two channels go to the staging tile as column pairs
                const int s = (j + 1) % 7;
                const uint32_t sh = staging0 + static_cast<uint32_t>(buf) * 2u * STG_BYTES + static_cast<uint32_t>(lane * 4);
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    const uint32_t sc = sh + static_cast<uint32_t>((2 * warp + ch) * DWM_SCS * 4);
                    const uint32_t h0 = pack_bf16x2(acc[s][ch][0], acc[s][ch][1]);  // strip g, columns 2t, 2t + 1
                    const uint32_t h1 = pack_bf16x2(acc[s][ch][2], acc[s][ch][3]);  // strip g + 8
                    sts_u32(sc, h0);
                    sts_u32(sc + 128u, h1);
                    if (SPLIT) {
                        sts_u32(sc + STG_BYTES, pack_bf16x2(acc[s][ch][0] - __uint_as_float(h0 << 16),
                                                            acc[s][ch][1] - __uint_as_float(h0 & 0xffff0000u)));
                        sts_u32(sc + STG_BYTES + 128u, pack_bf16x2(acc[s][ch][2] - __uint_as_float(h1 << 16),
                                                                   acc[s][ch][3] - __uint_as_float(h1 & 0xffff0000u)));
                    }
                }
            }
            // next input row -> the other plane buffer (its last readers finished before the previous barrier)
            const int hn = hr + 1;
            const bool conv = r + 1 < total_r && hn >= 0 && hn < p.H;
            if (conv) convert_row(buf ^ 1);
            // the output row completed by the PREVIOUS input row is stored now (its staging buffer is complete)
            if (r >= 7) store_row(buf ^ 1, r - 7);
            asm volatile("bar.sync 1, %0;" ::"n"(DWM_THREADS) : "memory");
            if (conv && threadIdx.x == 0) produce_next();  // the ring slot just converted is free again
        };

        // prologue: the first input row (if it lies inside the image) goes to plane buffer 0
        {
            const int h_first = jb.h0 - 3;
            const bool conv = h_first >= 0 && h_first < p.H;  // row r = 0
            if (conv) convert_row(0);
            asm volatile("bar.sync 1, %0;" ::"n"(DWM_THREADS) : "memory");
            if (conv && threadIdx.x == 0) produce_next();
        }
        for (int rb = 0; rb < total_r; rb += 7) {
            bool more = true;
            StaticFor<0, 7>::run([&](auto jc) {
                constexpr int j = decltype(jc)::value;
                if (more && rb + j < total_r) do_row(jc, rb + j);
                else more = false;
            });
        }
        // the last output row of the job (completed by input row total_r - 1, staged in buffer (total_r - 1) & 1)
        store_row((total_r - 1) & 1, total_r - 7);
    }
}

}  // namespace cpn
