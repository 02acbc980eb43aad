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
//            halves of a warp are neighbouring strips (so the W halo is shared in L1), and any C that is a multiple
//            of 32 fills whole warps;
//   march  : a lane walks down the input rows keeping the 7 in-flight output rows in registers (7*TW packed
//            accumulators) and its 49 packed weights in registers; an input row (TW+6 values per lane) feeds 49*TW
//            packed FMAs. The body is branch-free: rows outside the image arrive as zeros, and filter rows whose
//            output row lies outside the segment accumulate into slots that are never stored;
//   loads  : cp.async (8 bytes per lane and column, zero-filled outside the image) into a private shared-memory ring
//            DWS_DEPTH rows ahead - a DRAM-latency-deep prefetch that costs no registers; the lane reads its own
//            data back with LDS.64, so no barrier is involved;
//   stats  : 2*TW values per finished row are reduced over the 16 lanes of the half-warp with a transposing
//            butterfly (4+2+1+1 shuffles instead of 8*4) - a fixed summation order, so results are run-to-run
//            identical.
#pragma once
#include <type_traits>

#include "convnext_kernels.cuh"

namespace cpn {

constexpr int DWS_WARPS = 2;   // warps per CTA (independent of each other)
// Residency (ncu, profiles/r01_ncu_full_v10.md): a warp's registers sit in ONE of the SM's four 16K-register partitions, so at
// 192 registers per thread 2 warps fit per scheduler = 8 warps = 4 CTAs per SM (not the 5 that 65536 / (64 * 192) suggests);
// B=64 then runs 2.4 / 2.4 / 2.3 / 4.5 waves per stage. 3 warps per scheduler need <= 168 registers (TW = 3 variant below).
constexpr int DWS_CTAS_PER_SM = 4;
constexpr int DWS_TW = 4;      // output columns per lane
constexpr int DWS_DEPTH = 3;   // rows in flight ahead of the one being consumed
constexpr int DWS_NSLOT = DWS_DEPTH + 1;

__device__ __forceinline__ float2 ldg_f2(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }

// SPLIT (fp32 mode): the value is written as hi + lo bf16 planes and the statistics are those of hi + lo.
template <int C, int MAXREG, int TW = DWS_TW, bool SPLIT = false>
__global__ void __launch_bounds__(DWS_WARPS * 32) __maxnreg__(MAXREG)
dwconv7_stats_kernel(const float* __restrict__ x, const float* __restrict__ w49, const float* __restrict__ bias,
                     __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo, float* __restrict__ stats, int H,
                     int W, int nstrips, int rows_per_seg) {
    static_assert(C % 32 == 0, "32-channel groups");
    static_assert(TW >= 1 && TW <= 4, "the statistics butterfly carries 2 * TW <= 8 values");
    constexpr int NPART = C / 32;
    constexpr int NIN = TW + 6;
    constexpr int NT = DWS_WARPS * 32;
    __shared__ __align__(16) float2 ring[DWS_NSLOT][NIN][NT];  // [slot][column][thread]: private per thread

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int hl = lane & 15;
    const int nunits = nstrips * NPART;
    int unit = (blockIdx.x * DWS_WARPS + warp) * 2 + (lane >> 4);
    if (unit - (lane >> 4) >= nunits) return;  // whole warp beyond the end
    const bool live = unit < nunits;          // the upper half of the last warp may have no unit
    if (!live) unit = nunits - 1;             // compute a duplicate, store nothing
    const int cg = unit / nstrips;
    const int strip = unit - cg * nstrips;
    const int c0 = cg * 32 + 2 * hl;
    const int w0 = strip * TW;
    const int h0 = blockIdx.y * rows_per_seg;
    const int nout = min(rows_per_seg, H - h0);
    const int b = blockIdx.z;

    float2 wr[49];
#pragma unroll
    for (int k = 0; k < 49; ++k) wr[k] = ldg_f2(w49 + k * C + c0);
    const float2 bs = ldg_f2(bias + c0);
    ptx::pdl_wait();  // weights above are constants of the plan; the activations come from the previous kernel
    ptx::pdl_launch_dependents();

    // column validity of the TW+6 inputs of a row (bit i: column w0-3+i inside the image)
    unsigned colmask = 0;
#pragma unroll
    for (int i = 0; i < NIN; ++i)
        if (w0 - 3 + i >= 0 && w0 - 3 + i < W) colmask |= 1u << i;

    const float* xb = x + static_cast<size_t>(b) * H * W * C + c0;
    const int nrows_in = nout + 6;
    const uint32_t ring0 = ptx::smem_u32(&ring[0][0][threadIdx.x]);
    constexpr uint32_t COL_STRIDE = NT * 8, SLOT_STRIDE = NIN * NT * 8;

    // cp.async of the next input row into ring slot `slot`; rows outside the image are not fetched (their FMAs are
    // skipped), outside columns are zero-filled: src-size 0 reads nothing, so their out-of-image addresses are never
    // dereferenced. Always commits a group so that group counting stays uniform.
    const bool edge = __any_sync(0xffffffffu, colmask != ((1u << NIN) - 1u));
    constexpr long long ROW_ELEMS_PER_W = C;  // elements per pixel
    const long long row_stride = static_cast<long long>(W) * ROW_ELEMS_PER_W;
    const float* p_issue = xb + static_cast<long long>(w0 - 3) * C + static_cast<long long>(h0 - 3) * row_stride;
    int hi_issue = h0 - 3;                  // image row of the next row to fetch
    const int hi_end = h0 - 3 + nrows_in;   // one past the last input row of this segment
    auto issue_next = [&](int slot) {
        if (hi_issue >= 0 && hi_issue < min(H, hi_end)) {
            const uint32_t dst = ring0 + slot * SLOT_STRIDE;
            const float* p = p_issue;
            if (edge) {
                StaticFor<0, NIN>::run([&](auto ic) {
                    constexpr int i = decltype(ic)::value;
                    asm volatile("cp.async.ca.shared.global [%0 + %3], [%1 + %4], 8, %2;" ::"r"(dst), "l"(p),
                                 "r"(((colmask >> i) & 1u) * 8u), "n"(i * COL_STRIDE), "n"(i * C * 4)
                                 : "memory");
                });
            } else {
                StaticFor<0, NIN>::run([&](auto ic) {
                    constexpr int i = decltype(ic)::value;
                    asm volatile("cp.async.ca.shared.global [%0 + %2], [%1 + %3], 8;" ::"r"(dst), "l"(p),
                                 "n"(i * COL_STRIDE), "n"(i * C * 4)
                                 : "memory");
                });
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        ++hi_issue;
        p_issue += row_stride;
    };
#pragma unroll
    for (int r = 0; r < DWS_DEPTH; ++r) issue_next(r);

    float2 acc[7][TW];
#pragma unroll
    for (int s = 0; s < 7; ++s)
#pragma unroll
        for (int t = 0; t < TW; ++t) acc[s][t] = make_float2(0.f, 0.f);

    const bool hi8 = (lane & 8) != 0, hi4 = (lane & 4) != 0, hi2 = (lane & 2) != 0;
    int slot = 0;          // ring slot of row r
    int hi_cur = h0 - 3;   // image row of input row r
    // running output pointers (advance by one image row per finished output row)
    const size_t pix00 = (static_cast<size_t>(b) * H + h0) * W + w0;
    __nv_bfloat16* out_p = out + pix00 * C + c0;
    __nv_bfloat16* lo_p = SPLIT ? out_lo + pix00 * C + c0 : nullptr;
    const int sidx = hl >> 1;  // statistics value this lane ends up holding: 2 * t + which
    float* st_p = stats + ((pix00 + (sidx >> 1)) * NPART + cg) * 2 + (sidx & 1);
    const bool st_ok = live && (lane & 1) == 0 && (sidx >> 1) < TW && w0 + (sidx >> 1) < W;
    unsigned store_mask = 0;
#pragma unroll
    for (int t = 0; t < TW; ++t)
        if (live && w0 + t < W) store_mask |= 1u << t;

    for (int rb = 0; rb < nrows_in; rb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int r = rb + j;
            if (r >= nrows_in) break;
            {
                int ps = slot + DWS_DEPTH;
                if (ps >= DWS_NSLOT) ps -= DWS_NSLOT;
                issue_next(ps);  // overwrites the slot of row r - 1, whose values were consumed above
            }
            asm volatile("cp.async.wait_group %0;" ::"n"(DWS_DEPTH) : "memory");  // row r has landed
            // input row r feeds output row o = r - kh (segment-local) through filter row kh; kh == 0 opens the row
            // (bias as the first addend). Output rows with o < 0 or o >= nout collect values that are never stored: the
            // body stays branch-free with 28 independent accumulators between two uses of the same one. (Round 2 measured
            // the alternative -- filter rows outside the segment skipped by warp-uniform branches, finer row segments for
            // a fuller last wave --: 0.069-0.077 ms against 0.063 ms at 64 x 27 x 27 x 384, profiles/r02_dwconv.md.)
            if (hi_cur >= 0 && hi_cur < H) {
                float2 cur[NIN];
                const uint32_t src = ring0 + slot * SLOT_STRIDE;
                StaticFor<0, NIN>::run([&](auto ic) {
                    constexpr int i = decltype(ic)::value;
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2 + %3];"
                                 : "=f"(cur[i].x), "=f"(cur[i].y)
                                 : "r"(src), "n"(i * COL_STRIDE)
                                 : "memory");
                });
#pragma unroll
                for (int kw = 0; kw < 7; ++kw)
#pragma unroll
                    for (int kh = 0; kh < 7; ++kh) {
                        const int s = (j - kh + 7) % 7;
#pragma unroll
                        for (int t = 0; t < TW; ++t)
                            acc[s][t] = ffma2(cur[t + kw], wr[kh * 7 + kw], (kh == 0 && kw == 0) ? bs : acc[s][t]);
                    }
            } else {
                // a row of zero padding above or below the image: it only opens its output row
#pragma unroll
                for (int t = 0; t < TW; ++t) acc[j][t] = bs;
            }
            if (++slot == DWS_NSLOT) slot = 0;
            ++hi_cur;
            if (r >= 6) {
                // output row r - 6 is complete (slot (j + 1) % 7)
                const int s = (j + 1) % 7;
                float a[8];  // 2 * TW values, zero-padded to the butterfly's 8
#pragma unroll
                for (int i = 2 * TW; i < 8; ++i) a[i] = 0.f;
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    const uint32_t pk = pack_bf16x2(acc[s][t].x, acc[s][t].y);
                    if ((store_mask >> t) & 1u) *reinterpret_cast<uint32_t*>(out_p + t * C) = pk;
                    float v0 = __uint_as_float(pk << 16), v1 = __uint_as_float(pk & 0xffff0000u);
                    if (SPLIT) {
                        const uint32_t pl = pack_bf16x2(acc[s][t].x - v0, acc[s][t].y - v1);
                        if ((store_mask >> t) & 1u) *reinterpret_cast<uint32_t*>(lo_p + t * C) = pl;
                        v0 += __uint_as_float(pl << 16);  // what the three-pass GEMM sees: hi + lo
                        v1 += __uint_as_float(pl & 0xffff0000u);
                    }
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
                if (st_ok) *st_p = d;
                out_p += row_stride;
                if (SPLIT) lo_p += row_stride;
                st_p += static_cast<long long>(W) * NPART * 2;
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

}  // namespace cpn
