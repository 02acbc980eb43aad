// bf16-mode GEMMs of the ConvNeXt blocks: same TMA -> tcgen05 main loop as gemm_tcgen05.cuh, with an epilogue that
// moves whole 32-row x 128-byte blocks between registers, shared memory and HBM:
//
//   EPI_GELU_BF16      out_bf16 = gelu(acc + bias)                    CNBlock fc1 + GELU   (tv convnext.py:55-56)
//   EPI_SCALE_RES_F32  out_f32  = res_f32 + scale * (acc + bias)      CNBlock fc2, layer_scale, residual (:57,64-66)
//   EPI_BIAS_F32       out_f32  = acc + bias                          downsample conv 2x2 (:146-150)
//   EPI_LN_GELU_BF16   out_bf16 = gelu(rstd_m * (acc - mean_m * colsum) + bias)   LayerNorm folded into fc1 (:53-56):
//                      A holds the un-normalised depthwise-conv output, Bw = bf16(W1 * gamma), colsum[n] = sum_k Bw[n,k],
//                      bias = b1 + W1 * beta; mean_m / rstd_m come from the per-token partial sums written by
//                      dwconv7_stats_kernel (stats [M, nparts, 2]), summed here in a fixed order.
//
// Each of the 8 epilogue warps owns a 32-row quarter of the accumulator (its TMEM lanes) and every other column chunk.
// A chunk (64 bf16 or 32 fp32 columns = 128 B per row) is read with tcgen05.ld.x32, finished in registers, written
// to a 4 KB staging block in the 128-byte TMA swizzle (conflict-free 16-byte stores, lane = row) and sent to HBM with
// one TMA store; the fp32 residual comes in the same way (TMA load into the staging block, prefetched one chunk ahead
// and, for the first chunk of a tile, before the accumulator is even ready), so the epilogue issues no per-lane
// global loads or stores at all. out and res may be the same buffer (the residual stream is updated in place).
#pragma once
#include "gelu_fast.cuh"
#include "gemm_tcgen05.cuh"

namespace cpn {

enum GemmEpi : int { EPI_GELU_BF16 = 0, EPI_SCALE_RES_F32 = 1, EPI_BIAS_F32 = 2, EPI_LN_GELU_BF16 = 3 };
__host__ __device__ constexpr bool epi_out_bf16(int epi) { return epi == EPI_GELU_BF16 || epi == EPI_LN_GELU_BF16; }

constexpr int LN_MAX_PARTS = 24;  // K = C <= 768 channels in 32-channel groups

struct Gemm2Params {
    int M, N, K;
    int num_m_blocks, num_n_blocks, num_k_blocks;
    const float* bias;   // [N]
    const float* scale;  // [N] (EPI_SCALE_RES_F32)
    // EPI_LN_GELU_BF16
    const float* stats;   // [M, nparts, 2] partial (sum, sum of squares) over the K = C channels of each token
    const float* colsum;  // [N]
    int nparts;
    float inv_k;          // 1 / C
    float eps;
    // tile order: 0 = consecutive tiles walk the N blocks of one M block (co-running CTAs share the A tile),
    // 1 = consecutive tiles walk the M blocks of one N block (co-running CTAs share the B tile)
    int m_fastest;
    int dbg;  // -DCPN_DEV_KNOBS builds only (CPN_GEMM_DBG): bit 1 = MMA thread skips the MMAs, bit 2 = epilogue skips its TMA stores
};

// CG = 1: one CTA per 128 x BN tile. CG = 2: a CTA pair (cluster of two, tcgen05 cta_group::2) per 256 x BN tile;
// each CTA stages its own 128 rows of A and HALF of the B tile, so operand traffic from L2 per FLOP drops by a third.
template <int BN, int EPI, int CG = 1, int EW = GEMM_EPI_WARPS>
struct Gemm2Cfg {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = (BN / CG) * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int CHUNK_COLS = epi_out_bf16(EPI) ? 64 : 32;  // 128 bytes of output per row
    static constexpr int NCHUNK = BN / CHUNK_COLS;
    static constexpr int STG_BLOCK = 32 * 128;                        // one chunk: 32 rows x 128 B
    static constexpr int NGRP = EW / 4;                               // epilogue warps per 32-row TMEM lane quarter
    static constexpr int NBUF = EW == 8 ? 2 : 1;                      // staging blocks per epilogue warp (64 KB in total)
    static constexpr int THREADS = 128 + 32 * EW;
    static constexpr int STAGING_BYTES = EW * NBUF * STG_BLOCK;
    static constexpr int BAR_BYTES = (2 * 8 + 4 + 2 * EW) * 8 + 16;
    static_assert(EW == 8 || (EW == 16 && epi_out_bf16(EPI)), "16 epilogue warps: bf16-output epilogues only");
    static constexpr int BUDGET = 227 * 1024 - 1024 - STAGING_BYTES - BAR_BYTES;
    static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 8 ? 8 : (BUDGET / STAGE_BYTES);
    static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = TILE_BYTES + STAGING_BYTES + BAR_BYTES + 1024;
    static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                     : (2 * BN <= 256) ? 256 : 512;
    static_assert(BN % CHUNK_COLS == 0, "tile width must be a whole number of 128-byte output chunks");
    static_assert(STAGES >= 2, "not enough shared memory for a pipeline");
};

// EW = epilogue warps (8 or 16). 16 (bf16-output epilogues) puts four epilogue warps on every scheduler: the LN+GELU
// epilogue is issue-latency-bound with two (measured: fc1 of stage 3 takes 0.057 of its 0.071 ms with the MMAs and the
// stores switched off), and a 64-column chunk is processed as two 32-column halves to stay under 102 registers.
template <int BN, int EPI, int CG = 1, int EW = GEMM_EPI_WARPS>
__global__ void __launch_bounds__(128 + 32 * EW, 1)
gemm_tma_epi_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_out, const Gemm2Params p) {
    using S = Gemm2Cfg<BN, EPI, CG, EW>;
    constexpr int STAGES = S::STAGES;
    static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N constraint for M=128");
    static_assert(CG == 1 || (CG == 2 && BN % 32 == 0), "a CTA pair splits the B tile in two halves of BN/2 rows");
    const uint32_t rank = CG == 2 ? ptx::cluster_ctarank() : 0u;  // rank 0 of a pair issues the MMAs

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    auto stage_a = [&](int s) { return smem + s * S::STAGE_BYTES; };
    auto stage_b = [&](int s) { return smem + s * S::STAGE_BYTES + S::A_BYTES; };
    uint8_t* staging = smem + S::TILE_BYTES;  // 1024-aligned: STAGE_BYTES is a multiple of 1024
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + S::STAGING_BYTES);
    uint64_t* empty_bar = full_bar + 8;
    uint64_t* tmem_full_bar = empty_bar + 8;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint64_t* res_bar = tmem_empty_bar + 2;  // [epilogue warp][2]
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(res_bar + 2 * EW);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;  // m blocks are 128 * CG rows
    const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_b);
        ptx::prefetch_tmap(&tmap_out);
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tmem_full_bar[s], 1);
            ptx::mbar_init(&tmem_empty_bar[s], CG * EW);  // both CTAs' epilogues report to the leader
        }
        for (int s = 0; s < 2 * EW; ++s) ptx::mbar_init(&res_bar[s], 1);
        ptx::fence_mbar_init();
    } else if (warp == 2) {
        if (CG == 2) {
            ptx::tmem_alloc_2sm(tmem_holder, S::TMEM_COLS);
            ptx::tmem_relinquish_2sm();
        } else {
            ptx::tmem_alloc(tmem_holder, S::TMEM_COLS);
            ptx::tmem_relinquish();
        }
    }
    ptx::tc_fence_before();
    if (CG == 2) ptx::cluster_sync();  // the peer's barriers are initialised before anything arrives on them
    else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    // everything above is on-chip set-up; the operands come from the previous kernel of the stream
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                const int m_blk = p.m_fastest ? tile % p.num_m_blocks : tile / p.num_n_blocks;
                const int n_blk = p.m_fastest ? tile / p.num_m_blocks : tile % p.num_n_blocks;
                const int m0 = (m_blk * CG + static_cast<int>(rank)) * GEMM_BM;
                const int n0 = n_blk * BN + static_cast<int>(rank) * (BN / CG);
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (CG == 2) {
                        // both CTAs' bytes are counted on the leader's barrier, which the MMA thread waits on
                        if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * S::STAGE_BYTES);
                        ptx::tma_load_2d_2sm(stage_a(stage), &tmap_a, kb * GEMM_BK, m0, &full_bar[stage]);
                        ptx::tma_load_2d_2sm(stage_b(stage), &tmap_b, kb * GEMM_BK, n0, &full_bar[stage]);
                    } else {
                        ptx::mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                        ptx::tma_load_2d(stage_a(stage), &tmap_a, kb * GEMM_BK, m0, &full_bar[stage]);
                        ptx::tma_load_2d(stage_b(stage), &tmap_b, kb * GEMM_BK, n0, &full_bar[stage]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer: the whole warp walks the loop
        // (uniform control flow, so the descriptors live in uniform registers) and one elected lane issues
        if (rank == 0) {
            const uint32_t el = ptx::elect_one();
            constexpr uint32_t idesc = ptx::make_idesc_bf16(GEMM_BM * CG, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                ptx::mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + as * BN;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_a(stage)));
                    const uint64_t bdesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_b(stage)));
#pragma unroll
                    for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k) {
                        if (CPN_DBG(p.dbg, 2)) continue;
                        ptx::umma_bf16_if<CG>(el, tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    }
                    ptx::umma_commit_if<CG>(el, &empty_bar[stage]);  // frees the stage (in both CTAs of a pair)
                    if (kb == p.num_k_blocks - 1) ptx::umma_commit_if<CG>(el, &tmem_full_bar[as]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue (EW warps)
        constexpr int NGRP = S::NGRP, NBUF = S::NBUF;
        const int ew = warp - 4;
        const int q = warp & 3;    // TMEM lane quarter this warp may access = 32-row block of the tile
        const int half = ew >> 2;  // this warp takes chunks half, half + NGRP, ...
        const uint32_t stg = ptx::smem_u32(staging) + ew * NBUF * S::STG_BLOCK;
        uint8_t* stg_ptr = staging + ew * NBUF * S::STG_BLOCK;
        uint64_t* rbar = res_bar + 2 * ew;
        const uint32_t sw = static_cast<uint32_t>(lane & 7);          // swizzle phase of this lane's row
        const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;  // this lane's row inside a staging block
        uint32_t g = 0;  // chunks processed so far by this warp: buffer = g & 1, barrier parity = (g >> 1) & 1
        int it = 0;
        for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
            const int m_blk = p.m_fastest ? tile % p.num_m_blocks : tile / p.num_n_blocks;
            const int n_blk = p.m_fastest ? tile / p.num_m_blocks : tile % p.num_n_blocks;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            const int row0 = (m_blk * CG + static_cast<int>(rank)) * GEMM_BM + q * 32;
            const int col_tile = n_blk * BN;
            if (EPI == EPI_SCALE_RES_F32 && half < S::NCHUNK) {
                // residual of the first chunk: does not depend on the accumulator, so it travels while the MMAs run
                if (lane == 0) {
                    ptx::bulk_wait_read<0>();  // earlier stores have finished reading both staging blocks
                    const uint32_t b = g & 1;
                    ptx::mbar_arrive_expect_tx(&rbar[b], S::STG_BLOCK);
                    ptx::tma_load_2d(stg_ptr + b * S::STG_BLOCK, &tmap_out, col_tile + half * S::CHUNK_COLS, row0, &rbar[b]);
                }
                __syncwarp();
            }
            float ln_rs = 1.f, ln_nm = 0.f;  // rstd and -mean * rstd of this lane's row
            if (EPI == EPI_LN_GELU_BF16) {
                const int row = min(row0 + lane, p.M - 1);
                const float2* sp = reinterpret_cast<const float2*>(p.stats) + static_cast<size_t>(row) * p.nparts;
                // all partial sums in flight at once (the row is read once per tile), then a fixed-order sum
                float2 part[LN_MAX_PARTS];
#pragma unroll
                for (int i = 0; i < LN_MAX_PARTS; ++i)
                    part[i] = i < p.nparts ? __ldg(sp + i) : make_float2(0.f, 0.f);
                float s1 = 0.f, s2 = 0.f;
#pragma unroll
                for (int i = 0; i < LN_MAX_PARTS; ++i) {
                    s1 += part[i].x;
                    s2 += part[i].y;
                }
                const float mean = s1 * p.inv_k;
                const float var = fmaxf(s2 * p.inv_k - mean * mean, 0.f);
                ln_rs = rsqrtf(var + p.eps);
                ln_nm = -mean * ln_rs;
            }
            const float2 ln_rs2 = make_float2(ln_rs, ln_rs), ln_nm2 = make_float2(ln_nm, ln_nm);
            ptx::mbar_wait(&tmem_full_bar[as], aphase);
            ptx::tc_fence_after();
            const uint32_t taddr_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
            for (int c = half; c < S::NCHUNK; c += NGRP, ++g) {
                const uint32_t b = NBUF == 2 ? (g & 1) : 0;
                const uint32_t buf = stg + b * S::STG_BLOCK + row_off;
                const int col0 = col_tile + c * S::CHUNK_COLS;
                const bool last = c + NGRP >= S::NCHUNK;
                if (EPI == EPI_SCALE_RES_F32) {
                    if (!last) {
                        if (lane == 0) {
                            ptx::bulk_wait_read<0>();  // the store of the previous chunk has released the other block
                            ptx::mbar_arrive_expect_tx(&rbar[b ^ 1], S::STG_BLOCK);
                            ptx::tma_load_2d(stg_ptr + (b ^ 1) * S::STG_BLOCK, &tmap_out, col0 + 2 * S::CHUNK_COLS, row0,
                                             &rbar[b ^ 1]);
                        }
                        __syncwarp();
                    }
                } else {
                    if (lane == 0) ptx::bulk_wait_read<NBUF - 1>();  // the store that last used this block is done reading
                    __syncwarp();
                }
                uint32_t r0[32];
                ptx::tmem_ld32(taddr_row + c * S::CHUNK_COLS, r0);
                if (epi_out_bf16(EPI)) {
                    // 64 columns as two halves of 32: gelu(acc + bias) -> 2 x 4 x 16 bytes of bf16
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        if (hh == 1) ptx::tmem_ld32(taddr_row + c * S::CHUNK_COLS + 32, r0);
                        ptx::tmem_ld_wait();
                        if (last && hh == 1) {
                            ptx::tc_fence_before();
                            __syncwarp();
                            if (lane == 0) { if (CG == 2) ptx::mbar_arrive_leader(&tmem_empty_bar[as]); else ptx::mbar_arrive(&tmem_empty_bar[as]); }
                        }
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                        const uint32_t* src = r0 + 8 * j;
                        const int cj = col0 + 32 * hh + 8 * j;
                        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + cj));
                        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + cj + 4));
                        float2 v0 = make_float2(__uint_as_float(src[0]), __uint_as_float(src[1]));
                        float2 v1 = make_float2(__uint_as_float(src[2]), __uint_as_float(src[3]));
                        float2 v2 = make_float2(__uint_as_float(src[4]), __uint_as_float(src[5]));
                        float2 v3 = make_float2(__uint_as_float(src[6]), __uint_as_float(src[7]));
                        if (EPI == EPI_LN_GELU_BF16) {
                            // rstd * acc + (-mean * rstd * colsum + bias), two columns per instruction
                            const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.colsum + cj));
                            const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.colsum + cj + 4));
                            v0 = ffma2(v0, ln_rs2, ffma2(ln_nm2, make_float2(s0.x, s0.y), make_float2(b0.x, b0.y)));
                            v1 = ffma2(v1, ln_rs2, ffma2(ln_nm2, make_float2(s0.z, s0.w), make_float2(b0.z, b0.w)));
                            v2 = ffma2(v2, ln_rs2, ffma2(ln_nm2, make_float2(s1.x, s1.y), make_float2(b1.x, b1.y)));
                            v3 = ffma2(v3, ln_rs2, ffma2(ln_nm2, make_float2(s1.z, s1.w), make_float2(b1.z, b1.w)));
                        } else {
                            v0 = fadd2(v0, make_float2(b0.x, b0.y));
                            v1 = fadd2(v1, make_float2(b0.z, b0.w));
                            v2 = fadd2(v2, make_float2(b1.x, b1.y));
                            v3 = fadd2(v3, make_float2(b1.z, b1.w));
                        }
                        const float2 y0 = gelu_sat2(v0);
                        const float2 y1 = gelu_sat2(v1);
                        const float2 y2 = gelu_sat2(v2);
                        const float2 y3 = gelu_sat2(v3);
                        ptx::sts_u4(buf + ((static_cast<uint32_t>(4 * hh + j) ^ sw) << 4),
                                    make_uint4(pack2(y0.x, y0.y), pack2(y1.x, y1.y), pack2(y2.x, y2.y), pack2(y3.x, y3.y)));
                    }
                    }
                } else {
                    ptx::tmem_ld_wait();
                    if (last) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) { if (CG == 2) ptx::mbar_arrive_leader(&tmem_empty_bar[as]); else ptx::mbar_arrive(&tmem_empty_bar[as]); }
                    }
                    if (EPI == EPI_SCALE_RES_F32) ptx::mbar_wait(&rbar[b], (g >> 1) & 1);
                    // 32 columns -> res + scale * (acc + bias) (or acc + bias) -> 8 x 16 bytes of fp32
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint32_t a = buf + ((static_cast<uint32_t>(j) ^ sw) << 4);
                        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 4 * j));
                        float4 v = make_float4(__uint_as_float(r0[4 * j]) + b4.x, __uint_as_float(r0[4 * j + 1]) + b4.y,
                                               __uint_as_float(r0[4 * j + 2]) + b4.z, __uint_as_float(r0[4 * j + 3]) + b4.w);
                        if (EPI == EPI_SCALE_RES_F32) {
                            const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + col0 + 4 * j));
                            const float4 rs = ptx::lds_f4(a);
                            v.x = fmaf(v.x, s4.x, rs.x); v.y = fmaf(v.y, s4.y, rs.y);
                            v.z = fmaf(v.z, s4.z, rs.z); v.w = fmaf(v.w, s4.w, rs.w);
                        }
                        ptx::sts_u4(a, make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z),
                                                  __float_as_uint(v.w)));
                    }
                }
                ptx::fence_proxy_async();  // generic-proxy writes -> visible to the TMA store
                __syncwarp();
                if (lane == 0 && !CPN_DBG(p.dbg, 4)) {
                    ptx::tma_store_2d(&tmap_out, stg_ptr + b * S::STG_BLOCK, col0, row0);
                    ptx::bulk_commit();
                }
            }
            if (half >= S::NCHUNK) {  // a tile narrower than two chunks leaves this warp without work
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) { if (CG == 2) ptx::mbar_arrive_leader(&tmem_empty_bar[as]); else ptx::mbar_arrive(&tmem_empty_bar[as]); }
            }
        }
        if (lane == 0) ptx::bulk_wait<0>();  // all stores complete before the CTA (and its shared memory) goes away
        __syncwarp();
    }

    ptx::tc_fence_before();
    if (CG == 2) ptx::cluster_sync();  // neither CTA's shared memory / TMEM goes away while the pair still uses it
    else __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        if (CG == 2) ptx::tmem_dealloc_2sm(tmem_base, S::TMEM_COLS);
        else ptx::tmem_dealloc(tmem_base, S::TMEM_COLS);
    }
}

}  // namespace cpn
