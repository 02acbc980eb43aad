// Depthwise 7x7 conv (pad 3) + bias of a CNBlock (torchvision/models/convnext.py:52), second generation.
//
//   x [B,H,W,C] fp32 (residual stream)  ->  hi [B*H*W, C] bf16 = round(conv + bias)             (GEMM A operand)
//                                           lo [B*H*W, C] bf16 = round(v - hi)                   (fp32 mode only)
//                                           stats [B*H*W, C/32, 2] fp32 = (sum, sum of squares) over each 32-channel
//                                           group of the values the GEMM will see (hi, or hi + lo)
// The LayerNorm that follows the conv (convnext.py:53-54) is folded into the fc1 GEMM, which rebuilds mean / rstd per
// token from these partial sums (gemm_tma_epi.cuh, mlp_fused.cuh).
//
// Why a second kernel: dwconv7_stats_kernel keeps the 49 filter taps of a channel pair in 98 registers per lane, which
// caps residency at 8 warps per SM (192 registers); ncu showed the FMA pipe half idle inside a wave and the last of
// 2.3-4.5 waves a third full (profiles/r01_ncu_full_v10.md). Here
//   * the taps live in SHARED memory, laid out so that a warp's 16-byte loads of one tap pair are two conflict-free
//     128-byte wavefronts; a lane keeps only accumulators (7 in-flight output rows x 4 columns x channel pair) and one
//     input row: <= 128 registers, 14-16 resident warps per SM;
//   * lane = (channel pair cp of 8, slot q of 4); slots 2u / 2u+1 are the two 16-channel halves of unit u, a unit being
//     (image, group of 4 output columns) of the CTA's 32-channel group. The 16 lanes of a unit hold all 32 channels of
//     their tokens, so the LayerNorm partial sums are reduced inside the warp with a fixed-order butterfly;
//   * input rows arrive by TMA (4-D tensor map over [B,H,W,C], box = 32 channels x all columns of the job + 6 halo
//     columns, zero-filled outside the image, 128-byte swizzle so that the lanes' 8-byte reads spread over all banks)
//     into a 6-deep mbarrier ring filled by a producer warp that runs ahead ACROSS job boundaries;
//   * the CTAs are persistent over jobs = (32-channel group, image group, column block, row segment), each CTA taking a
//     contiguous range, so the tail of the grid is one short job instead of a partial wave. Row segments cost only
//     their halo rows' loads: filter rows whose output row lies outside the segment are skipped (uniform branches).
#pragma once
#include <type_traits>

#include "device_utils.cuh"

namespace cpn {

constexpr int DW2_TW = 4;        // output columns per lane
constexpr int DW2_NSTAGE = 6;    // input rows in flight
constexpr int DW2_MAX_WARPS = 7; // consumer warps (two units each); + 1 producer warp = 256 threads
constexpr int DW2_WBYTES = 7 * 4 * 2 * 8 * 16;  // taps of one 32-channel group: [kh][tap pair j][half][cp][16 bytes]

struct Dw2Params {
    const float* w49;   // [49, C]
    const float* bias;  // [C]
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;  // nullptr in bf16 mode
    float* stats;           // [B*H*W, C/32, 2]
    int B, H, W, C;
    int ngb;           // column groups (of DW2_TW columns) per column block
    int ncb;           // column blocks per row
    int nimg;          // images per job
    int nip;           // image groups = ceil(B / nimg)
    int units;         // nimg * ngb <= 2 * DW2_MAX_WARPS
    int nwc;           // consumer warps = ceil(units / 2)
    int rows_per_seg;  // output rows per row segment
    int nseg;
    int njobs;         // (C / 32) * nip * ncb * nseg
    int box_stride;    // bytes between the boxes of one stage (multiple of 1024)
    int box_tx;        // bytes one TMA box delivers: (ngb * DW2_TW + 6) * 128
    int stage_bytes;   // nimg * box_stride
};

struct Dw2Job {
    int grp, ip, cb, h0, nout;
};
__device__ __forceinline__ Dw2Job dw2_decode(const Dw2Params& p, int job) {
    Dw2Job j;
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

template <bool SPLIT, bool ROLL>
__global__ void __launch_bounds__(32 * (DW2_MAX_WARPS + 1), 2)
dwconv7_v2_kernel(const __grid_constant__ CUtensorMap tmap_x, const Dw2Params p) {
    constexpr int TW = DW2_TW, NS = DW2_NSTAGE, NIN = TW + 6;
    static_assert(TW == 4, "the swizzle arithmetic below uses (4 g + i) & 7 == ((g & 1) << 2) ^ (i & 7)");
    extern __shared__ uint8_t dw2_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dw2_smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* ring = smem;
    uint8_t* wsm = ring + NS * p.stage_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(wsm + DW2_WBYTES);
    uint64_t* empty_bar = full_bar + NS;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // contiguous job range of this CTA (consecutive jobs share the channel group, i.e. the taps in shared memory)
    const int job_lo = static_cast<int>(static_cast<long long>(blockIdx.x) * p.njobs / gridDim.x);
    const int job_hi = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * p.njobs / gridDim.x);

    if (threadIdx.x == 0) {
        ptx::prefetch_tmap(&tmap_x);
        for (int s = 0; s < NS; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], p.nwc);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();

    if (warp == p.nwc) {
        // ------------------------------------------------------------ producer: one thread streams the input rows
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int job = job_lo; job < job_hi; ++job) {
                const Dw2Job jb = dw2_decode(p, job);
                const int r_lo = max(jb.h0 - 3, 0), r_hi = min(jb.h0 + jb.nout + 3, p.H);
                const int col0 = jb.cb * p.ngb * TW - 3;
                for (int hr = r_lo; hr < r_hi; ++hr) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], static_cast<uint32_t>(p.nimg * p.box_tx));
                    uint8_t* dst = ring + stage * p.stage_bytes;
                    for (int i = 0; i < p.nimg; ++i)  // images beyond B and columns / rows outside the map arrive as zeros
                        ptx::tma_load_4d(dst + i * p.box_stride, &tmap_x, jb.grp * 32, col0, hr, jb.ip * p.nimg + i,
                                         &full_bar[stage]);
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
            }
        }
        return;
    }

    // ---------------------------------------------------------------- consumers
    const int cp = lane & 7, q = lane >> 3, slot = q >> 1, half = q & 1;
    const int NPART = p.C >> 5;
    const bool hi8 = (lane & 8) != 0, hi4 = (lane & 4) != 0, hi2 = (lane & 2) != 0;
    const int sidx = (lane & 15) >> 1;  // statistics value this lane ends up holding: 2 * t + which
    const uint32_t wlane = ptx::smem_u32(wsm) + static_cast<uint32_t>(half * 128 + cp * 16);
    const uint32_t ring0 = ptx::smem_u32(ring);
    int stage = 0;
    uint32_t phase = 0;
    int cur_grp = -1;
    float2 bs = make_float2(0.f, 0.f);

    for (int job = job_lo; job < job_hi; ++job) {
        const Dw2Job jb = dw2_decode(p, job);
        if (jb.grp != cur_grp) {
            // new 32-channel group: everyone has finished with the old taps, then refill (rare: jobs are contiguous)
            asm volatile("bar.sync 1, %0;" ::"r"(p.nwc * 32) : "memory");
            float* wf = reinterpret_cast<float*>(wsm);
            for (int idx = threadIdx.x; idx < DW2_WBYTES / 4; idx += p.nwc * 32) {
                const int e = idx & 3, c8 = (idx >> 2) & 7, hf = (idx >> 5) & 1, j = (idx >> 6) & 3, kh = idx >> 8;
                const int kw = 2 * j + (e >> 1);
                wf[idx] = kw < 7 ? __ldg(p.w49 + (kh * 7 + kw) * p.C + jb.grp * 32 + hf * 16 + 2 * c8 + (e & 1)) : 0.f;
            }
            asm volatile("bar.sync 1, %0;" ::"r"(p.nwc * 32) : "memory");
            cur_grp = jb.grp;
            bs = __ldg(reinterpret_cast<const float2*>(p.bias + jb.grp * 32 + half * 16 + 2 * cp));
        }
        int unit = warp * 2 + slot;
        const bool live = unit < p.units;
        if (!live) unit = p.units - 1;  // compute a duplicate, store nothing
        const int img = unit / p.ngb, g = unit - img * p.ngb;
        const int b = jb.ip * p.nimg + img;
        const bool valid = live && b < p.B;
        const int w0 = (jb.cb * p.ngb + g) * TW;  // first output column of this lane
        const int c0 = jb.grp * 32 + half * 16 + 2 * cp;
        // byte offset of this lane's (column 0, channel pair) inside a stage: the 16-byte chunk index is XORed with the
        // row-of-128-bytes index (= box column) & 7 by the TMA swizzle
        const uint32_t chunkx = static_cast<uint32_t>((half * 4 + (cp >> 1)) ^ ((g & 1) << 2));
        const uint32_t laneoff = static_cast<uint32_t>(img * p.box_stride + g * (TW * 128)) + (chunkx << 4) +
                                 static_cast<uint32_t>((cp & 1) * 8);
        unsigned store_mask = 0;
#pragma unroll
        for (int t = 0; t < TW; ++t)
            if (valid && w0 + t < p.W) store_mask |= 1u << t;
        const size_t pix00 = (static_cast<size_t>(min(b, p.B - 1)) * p.H + jb.h0) * p.W + w0;
        __nv_bfloat16* out_p = p.out_hi + pix00 * p.C + c0;
        __nv_bfloat16* lo_p = SPLIT ? p.out_lo + pix00 * p.C + c0 : nullptr;
        float* st_p = p.stats + ((pix00 + (sidx >> 1)) * NPART + jb.grp) * 2 + (sidx & 1);
        const bool st_ok = (lane & 1) == 0 && ((store_mask >> (sidx >> 1)) & 1u) != 0;
        const long long row_elems = static_cast<long long>(p.W) * p.C;
        const long long st_row = static_cast<long long>(p.W) * NPART * 2;

        float2 acc[7][TW];
        const int total_r = jb.nout + 6;  // input rows h0 - 3 .. h0 + nout + 2, in or outside the image
        // One input row. J is the accumulator slot that the row opens (its output row r through filter row 0); filter
        // row kh accumulates into slot (J - kh) mod 7 and slot (J + 1) mod 7 holds the output row r - 6 this row completes.
        auto do_row = [&](auto jc, const int r) {
            constexpr int j = decltype(jc)::value;
            const int hr = jb.h0 - 3 + r;
#pragma unroll
            for (int t = 0; t < TW; ++t) acc[j][t] = bs;  // bias is the first addend of the output row
            if (hr >= 0 && hr < p.H) {
                ptx::mbar_wait(&full_bar[stage], phase);
                const uint32_t base = ring0 + static_cast<uint32_t>(stage * p.stage_bytes) + laneoff;
                float2 cur[NIN];
#pragma unroll
                for (int i = 0; i < NIN; ++i) cur[i] = ptx::lds_f2((base ^ static_cast<uint32_t>((i & 7) << 4)) + i * 128);
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&empty_bar[stage]);  // the row is in registers: free the slot
                if (++stage == NS) { stage = 0; phase ^= 1; }
                // filter row kh feeds output row r - kh, which must lie in [0, nout)
                const int klo = max(0, r - jb.nout + 1), khi = min(6, r);
#pragma unroll
                for (int kh = 0; kh < 7; ++kh) {
                    if (kh >= klo && kh <= khi) {
                        const int s = (j - kh + 7) % 7;
                        const uint32_t wa = wlane + kh * (4 * 256);
                        const float4 w01 = ptx::lds_f4(wa), w23 = ptx::lds_f4(wa + 256), w45 = ptx::lds_f4(wa + 512);
                        const float2 w6 = ptx::lds_f2(wa + 768);
                        const float2 wk[7] = {make_float2(w01.x, w01.y), make_float2(w01.z, w01.w),
                                              make_float2(w23.x, w23.y), make_float2(w23.z, w23.w),
                                              make_float2(w45.x, w45.y), make_float2(w45.z, w45.w), w6};
#pragma unroll
                        for (int kw = 0; kw < 7; ++kw)
#pragma unroll
                            for (int t = 0; t < TW; ++t) acc[s][t] = ffma2(cur[t + kw], wk[kw], acc[s][t]);
                    }
                }
            }
            if (r >= 6) {
                // output row r - 6 is complete
                const int s = (j + 1) % 7;
                float a[8];
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    const float v0 = acc[s][t].x, v1 = acc[s][t].y;
                    const uint32_t pk = pack_bf16x2(v0, v1);
                    if ((store_mask >> t) & 1u) *reinterpret_cast<uint32_t*>(out_p + t * p.C) = pk;
                    float s0 = __uint_as_float(pk << 16), s1 = __uint_as_float(pk & 0xffff0000u);
                    if (SPLIT) {
                        const uint32_t pl = pack_bf16x2(v0 - s0, v1 - s1);
                        if ((store_mask >> t) & 1u) *reinterpret_cast<uint32_t*>(lo_p + t * p.C) = pl;
                        s0 += __uint_as_float(pl << 16);  // what the three-pass GEMM sees: hi + lo
                        s1 += __uint_as_float(pl & 0xffff0000u);
                    }
                    a[2 * t] = s0 + s1;
                    a[2 * t + 1] = fmaf(s1, s1, s0 * s0);
                }
                // transposing butterfly over the unit's 16 lanes (32 channels): afterwards lane hl holds the total of
                // value index hl >> 1; fixed order, so results are run-to-run identical
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
                if (st_ok) *st_p = d;
                out_p += row_elems;
                if (SPLIT) lo_p += row_elems;
                st_p += st_row;
            }
        };
        if (ROLL) {
            // compact loop (fits the instruction cache): every row uses slot 6 for the row it opens and slot 0 for the one
            // it completes, and the slots move down by one afterwards (48 register moves per row)
#pragma unroll 1
            for (int r = 0; r < total_r; ++r) {
                do_row(std::integral_constant<int, 6>{}, r);
#pragma unroll
                for (int sl = 0; sl < 6; ++sl)
#pragma unroll
                    for (int t = 0; t < TW; ++t) acc[sl][t] = acc[sl + 1][t];
            }
        } else {
            // slots renamed instead of moved: seven copies of the row body (76 KB of code)
            for (int rb = 0; rb < total_r; rb += 7) {
                bool more = true;
                StaticFor<0, 7>::run([&](auto jc) {
                    constexpr int j = decltype(jc)::value;
                    if (more && rb + j < total_r) do_row(jc, rb + j);
                    else more = false;
                });
            }
        }
    }
}

}  // namespace cpn
