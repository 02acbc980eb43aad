// fp32-mode ("bf16x3") GEMMs of the ConvNeXt blocks with the TMA epilogue of gemm_tma_epi.cuh.
//
// Every operand is a pair of bf16 planes (x ~= hi + lo, 16 mantissa bits) and a k-step issues three tcgen05 MMAs into the
// same fp32 accumulator: A_hi * B_hi + A_hi * B_lo + A_lo * B_hi (the dropped lo * lo term is 2^-18 relative). Measured
// against the fp32 reference: 2-6e-5 on pooled scores / logits (tests/test_parity_gpu.py), the 1e-4 bar of SURVEY 8(d).
//
//   EPI_LN_GELU_BF16   (hi, lo) = split(gelu_erf(rstd_m * (acc - mean_m * colsum) + bias))   LayerNorm folded into fc1
//                      (torchvision convnext.py:53-56): A = un-normalised depthwise-conv output planes, B = planes of
//                      W1 * gamma, colsum[n] = sum_k (B_hi + B_lo)[n,k], bias = b1 + W1 * beta, mean / rstd from the
//                      per-token partial sums of (hi + lo) written by dwconv7_v2_kernel<true>. GELU is the exact erf form.
//   EPI_SCALE_RES_F32  out_f32 = res_f32 + scale * (acc + bias)     fc2, layer_scale, residual (convnext.py:57, 64-66)
//   EPI_BIAS_F32       out_f32 = acc + bias                         downsample conv 2x2 (convnext.py:146-150)
//
// Pipeline: as gemm_tma_epi_kernel (TMA producer warp, one MMA-issuing thread, 8 epilogue warps, two TMEM accumulator
// stages, optional CTA pairs with cta_group::2). A stage holds [A_hi | A_lo | B_hi | B_lo] (64 KB for a 256-wide pair
// tile), so the epilogue staging is cut to ONE 4 KB block per warp to leave room for three stages: fp32 outputs move
// 32-column chunks (128 B per row, 128-byte swizzle), bf16 outputs 32-column half chunks as a hi block and a lo block of
// 32 rows x 64 B each (64-byte swizzle), both sent with TMA stores.
#pragma once
#include "gemm_tma_epi.cuh"

namespace cpn {

template <int BN, int EPI, int CG>
struct GemmSplitCfg {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = (BN / CG) * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = 2 * (A_BYTES + B_BYTES);  // [A_hi | A_lo | B_hi | B_lo]
    static constexpr int EW = GEMM_EPI_WARPS;
    static constexpr int NCHUNK = BN / 32;       // 32 accumulator columns per epilogue step
    static constexpr int STG_BLOCK = 32 * 128;   // per epilogue warp: 32 rows x 128 B, or hi + lo blocks of 32 x 64 B
    static constexpr int THREADS = 128 + 32 * EW;
    static constexpr int STAGING_BYTES = EW * STG_BLOCK;
    static constexpr int BAR_BYTES = (2 * 8 + 4 + EW) * 8 + 16;
    static constexpr int BUDGET = 227 * 1024 - 1024 - STAGING_BYTES - BAR_BYTES;
    static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 6 ? 6 : (BUDGET / STAGE_BYTES);
    static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = TILE_BYTES + STAGING_BYTES + BAR_BYTES + 1024;
    static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                     : (2 * BN <= 256) ? 256 : 512;
    static_assert(BN % 32 == 0, "tile width must be a whole number of 32-column chunks");
    static_assert(STAGES >= 2, "not enough shared memory for a pipeline");
    static_assert(B_BYTES % 1024 == 0, "operand blocks must keep the swizzle atom alignment");
};

template <int BN, int EPI, int CG>
__global__ void __launch_bounds__(128 + 32 * GEMM_EPI_WARPS, 1)
gemm_split_epi_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_a_lo,
                      const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_b_lo,
                      const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_out_lo,
                      const Gemm2Params p) {
    using S = GemmSplitCfg<BN, EPI, CG>;
    constexpr int STAGES = S::STAGES, EW = S::EW;
    static_assert(BN % 16 == 0 && BN >= 32 && BN <= 256, "UMMA N constraint for M=128");
    static_assert(CG == 1 || (CG == 2 && BN % 32 == 0), "a CTA pair splits the B tile in two halves of BN/2 rows");
    const uint32_t rank = CG == 2 ? ptx::cluster_ctarank() : 0u;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    auto stage_a = [&](int s) { return smem + s * S::STAGE_BYTES; };
    auto stage_a_lo = [&](int s) { return smem + s * S::STAGE_BYTES + S::A_BYTES; };
    auto stage_b = [&](int s) { return smem + s * S::STAGE_BYTES + 2 * S::A_BYTES; };
    auto stage_b_lo = [&](int s) { return smem + s * S::STAGE_BYTES + 2 * S::A_BYTES + S::B_BYTES; };
    uint8_t* staging = smem + S::TILE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + S::STAGING_BYTES);
    uint64_t* empty_bar = full_bar + 8;
    uint64_t* tmem_full_bar = empty_bar + 8;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint64_t* res_bar = tmem_empty_bar + 2;  // [epilogue warp]
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(res_bar + EW);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;  // m blocks are 128 * CG rows
    const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_a_lo);
        ptx::prefetch_tmap(&tmap_b);
        ptx::prefetch_tmap(&tmap_b_lo);
        ptx::prefetch_tmap(&tmap_out);
        if (epi_out_bf16(EPI)) ptx::prefetch_tmap(&tmap_out_lo);
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tmem_full_bar[s], 1);
            ptx::mbar_init(&tmem_empty_bar[s], CG * EW);
        }
        for (int s = 0; s < EW; ++s) ptx::mbar_init(&res_bar[s], 1);
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
    if (CG == 2) ptx::cluster_sync();
    else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
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
                    const int k0 = kb * GEMM_BK;
                    if (CG == 2) {
                        if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * S::STAGE_BYTES);
                        ptx::tma_load_2d_2sm(stage_a(stage), &tmap_a, k0, m0, &full_bar[stage]);
                        ptx::tma_load_2d_2sm(stage_b(stage), &tmap_b, k0, n0, &full_bar[stage]);
                        ptx::tma_load_2d_2sm(stage_a_lo(stage), &tmap_a_lo, k0, m0, &full_bar[stage]);
                        ptx::tma_load_2d_2sm(stage_b_lo(stage), &tmap_b_lo, k0, n0, &full_bar[stage]);
                    } else {
                        ptx::mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                        ptx::tma_load_2d(stage_a(stage), &tmap_a, k0, m0, &full_bar[stage]);
                        ptx::tma_load_2d(stage_b(stage), &tmap_b, k0, n0, &full_bar[stage]);
                        ptx::tma_load_2d(stage_a_lo(stage), &tmap_a_lo, k0, m0, &full_bar[stage]);
                        ptx::tma_load_2d(stage_b_lo(stage), &tmap_b_lo, k0, n0, &full_bar[stage]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (leader CTA; warp-uniform loop, one
        // elected lane issues): three MMAs per k-step into the same accumulator
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
                    const uint64_t ah = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_a(stage)));
                    const uint64_t al = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_a_lo(stage)));
                    const uint64_t bh = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_b(stage)));
                    const uint64_t bl = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_b_lo(stage)));
#pragma unroll
                    for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k) {
                        ptx::umma_bf16_if<CG>(el, tmem_d, ah + 2 * k, bh + 2 * k, idesc, (kb | k) != 0);
                        ptx::umma_bf16_if<CG>(el, tmem_d, ah + 2 * k, bl + 2 * k, idesc, 1u);
                        ptx::umma_bf16_if<CG>(el, tmem_d, al + 2 * k, bh + 2 * k, idesc, 1u);
                    }
                    ptx::umma_commit_if<CG>(el, &empty_bar[stage]);
                    if (kb == p.num_k_blocks - 1) ptx::umma_commit_if<CG>(el, &tmem_full_bar[as]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue (8 warps)
        const int ew = warp - 4;
        const int q = warp & 3;    // TMEM lane quarter = 32-row block of the tile
        const int grp = ew >> 2;   // this warp takes chunks grp, grp + 2, ...
        const uint32_t stg = ptx::smem_u32(staging) + ew * S::STG_BLOCK;
        uint8_t* stg_ptr = staging + ew * S::STG_BLOCK;
        uint64_t* rbar = res_bar + ew;
        auto release_tmem = [&](int as) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) ptx::mbar_arrive_leader(&tmem_empty_bar[as]); else ptx::mbar_arrive(&tmem_empty_bar[as]); }
        };
        uint32_t g = 0;  // residual loads waited for so far by this warp: barrier parity = g & 1
        int it = 0;
        for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
            const int m_blk = p.m_fastest ? tile % p.num_m_blocks : tile / p.num_n_blocks;
            const int n_blk = p.m_fastest ? tile / p.num_m_blocks : tile % p.num_n_blocks;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            const int row0 = (m_blk * CG + static_cast<int>(rank)) * GEMM_BM + q * 32;
            const int col_tile = n_blk * BN;
            if (EPI == EPI_SCALE_RES_F32 && grp < S::NCHUNK) {
                // residual of the first chunk: independent of the accumulator, so it travels while the MMAs run
                if (lane == 0) {
                    ptx::bulk_wait_read<0>();
                    ptx::mbar_arrive_expect_tx(rbar, S::STG_BLOCK);
                    ptx::tma_load_2d(stg_ptr, &tmap_out, col_tile + grp * 32, row0, rbar);
                }
                __syncwarp();
            }
            float ln_rs = 1.f, ln_nm = 0.f;  // rstd and -mean * rstd of this lane's row
            if (EPI == EPI_LN_GELU_BF16) {
                const int row = min(row0 + lane, p.M - 1);
                const float2* sp = reinterpret_cast<const float2*>(p.stats) + static_cast<size_t>(row) * p.nparts;
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
            ptx::mbar_wait(&tmem_full_bar[as], aphase);
            ptx::tc_fence_after();
            const uint32_t taddr_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
            for (int c = grp; c < S::NCHUNK; c += 2) {
                const int col0 = col_tile + c * 32;
                const bool last = c + 2 >= S::NCHUNK;
                uint32_t r0[32];
                ptx::tmem_ld32(taddr_row + c * 32, r0);
                if (EPI != EPI_SCALE_RES_F32) {
                    if (lane == 0) ptx::bulk_wait_read<0>();  // the stores that last used the staging block have read it
                    __syncwarp();
                }
                ptx::tmem_ld_wait();
                if (last) release_tmem(as);
                if (epi_out_bf16(EPI)) {
                    // 32 columns -> exact GELU -> hi block and lo block, 32 rows x 64 B each in the 64-byte TMA swizzle
                    const uint32_t sw = static_cast<uint32_t>((lane >> 1) & 3);
                    const uint32_t hi_row = stg + static_cast<uint32_t>(lane) * 64u;
                    const uint32_t lo_row = hi_row + 2048u;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int cj = col0 + 8 * j;
                        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + cj));
                        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + cj + 4));
                        const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.colsum + cj));
                        const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.colsum + cj + 4));
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        const float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
                        uint32_t ph[4], pl[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int i0 = 2 * e;
                            const float2 pre = make_float2(
                                fmaf(__uint_as_float(r0[8 * j + i0]), ln_rs, fmaf(ln_nm, ss[i0], bb[i0])),
                                fmaf(__uint_as_float(r0[8 * j + i0 + 1]), ln_rs, fmaf(ln_nm, ss[i0 + 1], bb[i0 + 1])));
                            const float2 ge = gelu_erfc2(pre);
                            const float v[2] = {ge.x, ge.y};
                            ph[e] = pack_bf16x2(v[0], v[1]);
                            pl[e] = pack_bf16x2(v[0] - __uint_as_float(ph[e] << 16),
                                                v[1] - __uint_as_float(ph[e] & 0xffff0000u));
                        }
                        const uint32_t off = (static_cast<uint32_t>(j) ^ sw) << 4;
                        ptx::sts_u4(hi_row + off, make_uint4(ph[0], ph[1], ph[2], ph[3]));
                        ptx::sts_u4(lo_row + off, make_uint4(pl[0], pl[1], pl[2], pl[3]));
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_2d(&tmap_out, stg_ptr, col0, row0);
                        ptx::tma_store_2d(&tmap_out_lo, stg_ptr + 2048, col0, row0);
                        ptx::bulk_commit();
                    }
                } else {
                    const uint32_t sw = static_cast<uint32_t>(lane & 7);
                    const uint32_t buf = stg + static_cast<uint32_t>(lane) * 128u;
                    if (EPI == EPI_SCALE_RES_F32) {
                        ptx::mbar_wait(rbar, g & 1);
                        ++g;
                    }
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
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_2d(&tmap_out, stg_ptr, col0, row0);
                        ptx::bulk_commit();
                        if (EPI == EPI_SCALE_RES_F32 && !last) {
                            // residual of this warp's next chunk: the block is free once the store has read it
                            ptx::bulk_wait_read<0>();
                            ptx::mbar_arrive_expect_tx(rbar, S::STG_BLOCK);
                            ptx::tma_load_2d(stg_ptr, &tmap_out, col0 + 64, row0, rbar);
                        }
                    }
                    __syncwarp();
                }
            }
            if (grp >= S::NCHUNK) release_tmem(as);  // a tile narrower than two chunks leaves this warp without work
        }
        if (lane == 0) ptx::bulk_wait<0>();
        __syncwarp();
    }

    ptx::tc_fence_before();
    if (CG == 2) ptx::cluster_sync();
    else __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        if (CG == 2) ptx::tmem_dealloc_2sm(tmem_base, S::TMEM_COLS);
        else ptx::tmem_dealloc(tmem_base, S::TMEM_COLS);
    }
}

}  // namespace cpn
