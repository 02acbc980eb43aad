// ResNet convolutions as tcgen05 GEMMs with folded BatchNorm (features/resnet_features.py:97-117 Bottleneck, :55-74
// BasicBlock, :174-210): bf16 operands, fp32 accumulation in TMEM, TMA epilogue.
//
//   A operand  plain   : [M, K] bf16 rows (1x1 convs: the activation itself), 2-D TMA box 64 x 128 rows;
//              implicit: the 3x3 (stride 1, pad 1) conv reads its taps straight from the channels-last activation
//                        [B, H, W, Cin] with 4-D TMA boxes {64 channels, W, Ht, 1}: an M tile is Ht full image rows
//                        (Ht * W <= 128 rows, a multiple of 16), k-block kb = (tap, 64-channel slice), and the box of tap
//                        (kh, kw) starts at (w = kw - 1, h = h0 + kh - 1): rows and columns outside the map arrive as
//                        zeros, which is the conv's padding. No im2col matrix exists in HBM.
//   B operand  [Cout, K] bf16, K ordered (kh, kw, cin), BatchNorm scale folded in.
//   epilogues  CEPI_RELU_BF16   out_bf16 = relu(acc + bias)                              conv1 / conv2 + bn + relu
//              CEPI_RES_RELU    out_f32 = relu(res_f32 + acc + bias), out_bf16 = bf16(out_f32)   conv3 + bn3 + identity + relu
//              CEPI_BIAS_F32    out_f32 = acc + bias                                     downsample conv + bn (the identity)
//              CEPI_RESB_RELU   out_bf16 = bf16(relu(res_bf16 + acc + bias))             the same block end on a bf16 residual
//              CEPI_RESB_RELU_F32  out_f32 = relu(res_bf16 + acc + bias)                 ... of the last block (fp32 features)
//   With the bf16 residual stream (programs_resnet.py, bf16 mode) a block end moves 2 + 2 bytes per element instead of
//   4 + 4 + 2: these 1x1 convs are HBM-bound (K = 64 .. 512), 1.13 of the 2.86 ms of cfg3. p.relu = 0 turns
//   CEPI_RELU_BF16 into the projection shortcut with a bf16 result.
// Tiles of an implicit conv cover tile_rows = Ht * W < 128 accumulator rows; rows beyond are never loaded or stored
// (TMA boxes of 16 rows, only the valid ones are issued).
#pragma once
#include "gemm_tma_epi.cuh"

namespace cpn {

enum ConvEpi : int { CEPI_RELU_BF16 = 0, CEPI_RES_RELU = 1, CEPI_BIAS_F32 = 2, CEPI_RESB_RELU = 3, CEPI_RESB_RELU_F32 = 4 };
__host__ __device__ constexpr bool cepi_resb(int epi) { return epi == CEPI_RESB_RELU || epi == CEPI_RESB_RELU_F32; }

struct ConvGemmParams {
    int M, N;
    int num_m_tiles, num_n_blocks, num_k_blocks;
    int tile_rows;      // rows of the output matrix per M tile: 128 (plain) or Ht * W (implicit)
    int implicit;       // 1: A comes from the 4-D activation map
    int cblocks;        // Cin / 64 (implicit)
    int ksize, pad;     // 3, 1 (implicit)
    int ht;             // image rows per tile (implicit)
    int tiles_per_img;  // H / ht (implicit)
    const float* bias;  // [N]
    int relu;           // CEPI_RELU_BF16 only: 0 = no activation (projection shortcut)
};

template <int BN, int EPI, int CG>
struct ConvGemmCfg {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = (BN / CG) * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int EW = GEMM_EPI_WARPS;
    static constexpr int NCHUNK = BN / 32;
    static constexpr int STG_BLOCK = 32 * 128;        // fp32 chunk: 32 rows x 128 B; bf16 chunk: 32 rows x 64 B
    // block 0: fp32 / bf16 output; block 1: bf16 copy (CEPI_RES_RELU), or bf16 output in its first half and the bf16
    // residual chunk in its second half (CEPI_RESB_*)
    static constexpr int STG_PER_WARP = 2 * STG_BLOCK;
    static constexpr int THREADS = 128 + 32 * EW;
    static constexpr int STAGING_BYTES = EW * STG_PER_WARP;
    static constexpr int BAR_BYTES = (2 * 8 + 4 + EW) * 8 + 16;
    static constexpr int BUDGET = 227 * 1024 - 1024 - STAGING_BYTES - BAR_BYTES;
    static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 8 ? 8 : (BUDGET / STAGE_BYTES);
    static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = TILE_BYTES + STAGING_BYTES + BAR_BYTES + 1024;
    static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                     : (2 * BN <= 256) ? 256 : 512;
    static_assert(BN % 32 == 0 && BN <= 256, "tile width: whole 32-column chunks, UMMA N <= 256");
    static_assert(B_BYTES % 1024 == 0, "operand blocks must keep the swizzle atom alignment");
    static_assert(STAGES >= 2, "not enough shared memory for a pipeline");
};

template <int BN, int EPI, int CG>
__global__ void __launch_bounds__(128 + 32 * GEMM_EPI_WARPS, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_res, const __grid_constant__ CUtensorMap tmap_out,
                 const __grid_constant__ CUtensorMap tmap_out_bf16, const ConvGemmParams p) {
    using S = ConvGemmCfg<BN, EPI, CG>;
    constexpr int STAGES = S::STAGES, EW = S::EW;
    const uint32_t rank = CG == 2 ? ptx::cluster_ctarank() : 0u;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    auto stage_a = [&](int s) { return smem + s * S::STAGE_BYTES; };
    auto stage_b = [&](int s) { return smem + s * S::STAGE_BYTES + S::A_BYTES; };
    uint8_t* staging = smem + S::TILE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + S::STAGING_BYTES);
    uint64_t* empty_bar = full_bar + 8;
    uint64_t* tmem_full_bar = empty_bar + 8;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint64_t* res_bar = tmem_empty_bar + 2;  // [epilogue warp]
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(res_bar + EW);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_m_groups = (p.num_m_tiles + CG - 1) / CG;  // a CTA pair takes two consecutive M tiles
    const int num_tiles = num_m_groups * p.num_n_blocks;
    const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;
    // bytes one TMA A box delivers: tile_rows rows of 128 B (implicit tiles may be shorter than the 128-row stage)
    const uint32_t a_tx = static_cast<uint32_t>(p.tile_rows) * 128u;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_b);
        ptx::prefetch_tmap(&tmap_out);
        if (EPI == CEPI_RES_RELU) {
            ptx::prefetch_tmap(&tmap_res);
            ptx::prefetch_tmap(&tmap_out_bf16);
        }
        if (cepi_resb(EPI)) ptx::prefetch_tmap(&tmap_res);
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
                const int m_grp = tile / p.num_n_blocks, n_blk = tile % p.num_n_blocks;
                const int m_tile = m_grp * CG + static_cast<int>(rank);
                const int n0 = n_blk * BN + static_cast<int>(rank) * (BN / CG);
                const int img = p.implicit ? m_tile / p.tiles_per_img : 0;
                const int h0 = p.implicit ? (m_tile - img * p.tiles_per_img) * p.ht : 0;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (CG == 2) {
                        if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * (a_tx + S::B_BYTES));
                    } else {
                        ptx::mbar_arrive_expect_tx(&full_bar[stage], a_tx + S::B_BYTES);
                    }
                    if (p.implicit) {
                        const int tap = kb / p.cblocks, cb = kb - tap * p.cblocks;
                        const int kh = tap / p.ksize, kw = tap - kh * p.ksize;
                        if (CG == 2) ptx::tma_load_4d_2sm(stage_a(stage), &tmap_a, cb * 64, kw - p.pad, h0 + kh - p.pad, img, &full_bar[stage]);
                        else ptx::tma_load_4d(stage_a(stage), &tmap_a, cb * 64, kw - p.pad, h0 + kh - p.pad, img, &full_bar[stage]);
                    } else {
                        if (CG == 2) ptx::tma_load_2d_2sm(stage_a(stage), &tmap_a, kb * GEMM_BK, m_tile * GEMM_BM, &full_bar[stage]);
                        else ptx::tma_load_2d(stage_a(stage), &tmap_a, kb * GEMM_BK, m_tile * GEMM_BM, &full_bar[stage]);
                    }
                    if (CG == 2) ptx::tma_load_2d_2sm(stage_b(stage), &tmap_b, kb * GEMM_BK, n0, &full_bar[stage]);
                    else ptx::tma_load_2d(stage_b(stage), &tmap_b, kb * GEMM_BK, n0, &full_bar[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (leader CTA, one elected lane)
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
                    for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k)
                        ptx::umma_bf16_if<CG>(el, tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
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
        const int q = warp & 3;   // TMEM lane quarter = rows 32 q .. 32 q + 31 of the tile
        const int grp = ew >> 2;  // this warp takes chunks grp, grp + 2, ...
        const uint32_t stg = ptx::smem_u32(staging) + ew * S::STG_PER_WARP;
        uint8_t* stg_ptr = staging + ew * S::STG_PER_WARP;
        uint64_t* rbar = res_bar + ew;
        // 16-row TMA boxes: which halves of this warp's 32 rows exist in the tile
        const bool half_ok[2] = {q * 32 < p.tile_rows, q * 32 + 16 < p.tile_rows};
        auto release_tmem = [&](int as) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) ptx::mbar_arrive_leader(&tmem_empty_bar[as]); else ptx::mbar_arrive(&tmem_empty_bar[as]); }
        };
        auto load_res = [&](int col0, int row0) {  // lane 0: residual chunk -> staging block 0
            ptx::bulk_wait_read<0>();
            const uint32_t bytes = (half_ok[0] ? 2048u : 0u) + (half_ok[1] ? 2048u : 0u);
            ptx::mbar_arrive_expect_tx(rbar, bytes);  // a warp whose rows are all outside the tile still completes the phase
            if (half_ok[0]) ptx::tma_load_2d(stg_ptr, &tmap_res, col0, row0, rbar);
            if (half_ok[1]) ptx::tma_load_2d(stg_ptr + 2048, &tmap_res, col0, row0 + 16, rbar);
        };
        constexpr int RESB_OFF = S::STG_BLOCK + 2048;  // bf16 residual chunk: 32 rows x 64 B, never read by a store
        auto load_resb = [&](int col0, int row0) {  // lane 0
            const uint32_t bytes = (half_ok[0] ? 1024u : 0u) + (half_ok[1] ? 1024u : 0u);
            ptx::mbar_arrive_expect_tx(rbar, bytes);
            if (half_ok[0]) ptx::tma_load_2d(stg_ptr + RESB_OFF, &tmap_res, col0, row0, rbar);
            if (half_ok[1]) ptx::tma_load_2d(stg_ptr + RESB_OFF + 1024, &tmap_res, col0, row0 + 16, rbar);
        };
        uint32_t g = 0;  // residual phases consumed by this warp
        int it = 0;
        for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
            const int m_grp = tile / p.num_n_blocks, n_blk = tile % p.num_n_blocks;
            const int m_tile = m_grp * CG + static_cast<int>(rank);
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            const int row0 = m_tile * p.tile_rows + q * 32;
            const int col_tile = n_blk * BN;
            if (EPI == CEPI_RES_RELU && grp < S::NCHUNK) {
                if (lane == 0) load_res(col_tile + grp * 32, row0);  // travels while the MMAs run
                __syncwarp();
            }
            if (cepi_resb(EPI) && grp < S::NCHUNK) {
                if (lane == 0) load_resb(col_tile + grp * 32, row0);
                __syncwarp();
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
                if (EPI != CEPI_RES_RELU && !cepi_resb(EPI)) {
                    if (lane == 0) ptx::bulk_wait_read<0>();  // the stores that last used the staging blocks have read them
                    __syncwarp();
                }
                ptx::tmem_ld_wait();
                if (last) release_tmem(as);
                if (cepi_resb(EPI)) {
                    // residual chunk (bf16, 64-byte swizzle) + acc + bias -> relu -> bf16 chunk or fp32 chunk
                    const uint32_t sw2 = static_cast<uint32_t>((lane >> 1) & 3);
                    const uint32_t rrow = stg + RESB_OFF + static_cast<uint32_t>(lane) * 64u;
                    ptx::mbar_wait(rbar, g & 1);
                    ++g;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint4 rs = ptx::lds_u4(rrow + ((static_cast<uint32_t>(j) ^ sw2) << 4));
                        const uint32_t rw[4] = {rs.x, rs.y, rs.z, rs.w};
                        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 8 * j));
                        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 8 * j + 4));
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            v[8 * j + 2 * e] = fmaxf(__uint_as_float(r0[8 * j + 2 * e]) + bb[2 * e] + __uint_as_float(rw[e] << 16), 0.f);
                            v[8 * j + 2 * e + 1] =
                                fmaxf(__uint_as_float(r0[8 * j + 2 * e + 1]) + bb[2 * e + 1] + __uint_as_float(rw[e] & 0xffff0000u), 0.f);
                        }
                    }
                    if (lane == 0) ptx::bulk_wait_read<0>();  // the store that last used the output block has read it
                    __syncwarp();
                    if (EPI == CEPI_RESB_RELU) {
                        const uint32_t orow = stg + S::STG_BLOCK + static_cast<uint32_t>(lane) * 64u;
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            ptx::sts_u4(orow + ((static_cast<uint32_t>(j) ^ sw2) << 4),
                                        make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                                   pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7])));
                    } else {
                        const uint32_t sw = static_cast<uint32_t>(lane & 7);
                        const uint32_t buf = stg + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            ptx::sts_u4(buf + ((static_cast<uint32_t>(j) ^ sw) << 4),
                                        make_uint4(__float_as_uint(v[4 * j]), __float_as_uint(v[4 * j + 1]), __float_as_uint(v[4 * j + 2]),
                                                   __float_as_uint(v[4 * j + 3])));
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (EPI == CEPI_RESB_RELU) {
                            if (half_ok[0]) ptx::tma_store_2d(&tmap_out, stg_ptr + S::STG_BLOCK, col0, row0);
                            if (half_ok[1]) ptx::tma_store_2d(&tmap_out, stg_ptr + S::STG_BLOCK + 1024, col0, row0 + 16);
                        } else {
                            if (half_ok[0]) ptx::tma_store_2d(&tmap_out, stg_ptr, col0, row0);
                            if (half_ok[1]) ptx::tma_store_2d(&tmap_out, stg_ptr + 2048, col0, row0 + 16);
                        }
                        ptx::bulk_commit();
                        if (!last) load_resb(col0 + 64, row0);  // this warp's next chunk (its block was read above)
                    }
                    __syncwarp();
                } else if (EPI == CEPI_RELU_BF16) {
                    // 32 columns -> relu -> bf16, 32 rows x 64 B in the 64-byte TMA swizzle
                    const float lo_clamp = p.relu ? 0.f : -INFINITY;
                    const uint32_t sw = static_cast<uint32_t>((lane >> 1) & 3);
                    const uint32_t row = stg + static_cast<uint32_t>(lane) * 64u;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 8 * j));
                        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 8 * j + 4));
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        uint32_t pk[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            pk[e] = pack_bf16x2(fmaxf(__uint_as_float(r0[8 * j + 2 * e]) + bb[2 * e], lo_clamp),
                                                fmaxf(__uint_as_float(r0[8 * j + 2 * e + 1]) + bb[2 * e + 1], lo_clamp));
                        ptx::sts_u4(row + ((static_cast<uint32_t>(j) ^ sw) << 4), make_uint4(pk[0], pk[1], pk[2], pk[3]));
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (half_ok[0]) ptx::tma_store_2d(&tmap_out, stg_ptr, col0, row0);
                        if (half_ok[1]) ptx::tma_store_2d(&tmap_out, stg_ptr + 1024, col0, row0 + 16);
                        ptx::bulk_commit();
                    }
                } else {
                    const uint32_t sw = static_cast<uint32_t>(lane & 7);
                    const uint32_t buf = stg + static_cast<uint32_t>(lane) * 128u;
                    const uint32_t sw2 = static_cast<uint32_t>((lane >> 1) & 3);
                    const uint32_t row2 = stg + S::STG_BLOCK + static_cast<uint32_t>(lane) * 64u;
                    if (EPI == CEPI_RES_RELU) {
                        ptx::mbar_wait(rbar, g & 1);
                        ++g;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint32_t a = buf + ((static_cast<uint32_t>(j) ^ sw) << 4);
                        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 4 * j));
                        float4 v = make_float4(__uint_as_float(r0[4 * j]) + b4.x, __uint_as_float(r0[4 * j + 1]) + b4.y,
                                               __uint_as_float(r0[4 * j + 2]) + b4.z, __uint_as_float(r0[4 * j + 3]) + b4.w);
                        if (EPI == CEPI_RES_RELU) {
                            const float4 rs = ptx::lds_f4(a);
                            v.x = fmaxf(v.x + rs.x, 0.f); v.y = fmaxf(v.y + rs.y, 0.f);
                            v.z = fmaxf(v.z + rs.z, 0.f); v.w = fmaxf(v.w + rs.w, 0.f);
                            // bf16 copy (the next conv's A operand): 8-byte halves of the 16-byte chunk j >> 1
                            const uint32_t a2 = row2 + ((static_cast<uint32_t>(j >> 1) ^ sw2) << 4) + static_cast<uint32_t>((j & 1) * 8);
                            ptx::sts_f2(a2, make_float2(__uint_as_float(pack_bf16x2(v.x, v.y)), __uint_as_float(pack_bf16x2(v.z, v.w))));
                        }
                        ptx::sts_u4(a, make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z),
                                                  __float_as_uint(v.w)));
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (half_ok[0]) ptx::tma_store_2d(&tmap_out, stg_ptr, col0, row0);
                        if (half_ok[1]) ptx::tma_store_2d(&tmap_out, stg_ptr + 2048, col0, row0 + 16);
                        if (EPI == CEPI_RES_RELU) {
                            if (half_ok[0]) ptx::tma_store_2d(&tmap_out_bf16, stg_ptr + S::STG_BLOCK, col0, row0);
                            if (half_ok[1]) ptx::tma_store_2d(&tmap_out_bf16, stg_ptr + S::STG_BLOCK + 1024, col0, row0 + 16);
                        }
                        ptx::bulk_commit();
                        if (EPI == CEPI_RES_RELU && !last) load_res(col0 + 64, row0);  // this warp's next chunk
                    }
                    __syncwarp();
                }
            }
            if (grp >= S::NCHUNK) release_tmem(as);
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
