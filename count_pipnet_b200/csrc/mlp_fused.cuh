// Fused CNBlock MLP (bf16 mode): x += layer_scale * (fc2(gelu(fc1(LN(y)))) with the 4C-wide hidden activations kept on
// the SM (torchvision convnext.py:53-57, 64-66; LayerNorm folded into fc1 as in gemm_tma_epi.cuh EPI_LN_GELU_BF16).
//
// One CTA (CG = 1) or CTA pair (CG = 2, tcgen05 cta_group::2) owns a tile of 128 rows per CTA:
//   X  [128, C] bf16 (un-normalised depthwise-conv output) stays in shared memory for the whole tile;
//   the hidden dimension is walked in chunks of 64 columns:
//     G1(j): acc1[j & 1] (TMEM, 64 columns)  = X * W1g[64 j .. 64 j + 63, :]^T                (K = C)
//     E1(j): 8 epilogue warps: rstd * (acc1 - mean * colsum) + b1f -> GELU -> bf16 -> H slot (shared memory, the
//            128-byte-swizzled K-major layout tcgen05 reads as the A operand)
//     G2(j): acc2 (TMEM, C columns) += H(j) * W2[:, 64 j .. 64 j + 63]^T                      (K = 64)
//   issue order G1(0) G1(1) G2(0) G1(2) G2(1) ... so the tensor pipe always has a queued GEMM while E1 runs;
//   final epilogue: x[row, :] += scale * (acc2 + b2), per-lane 128-byte row segments straight from registers.
// The weights stream through a ring of equal-sized stages (one stage = the W1 rows of a chunk over all K, or the W2
// k-block of a chunk over all C outputs); with CG = 2 each CTA stages half of every weight tile.
// TMEM: acc2 at columns [0, C), acc1 double-buffered at [C, C + 128).
#pragma once
#include "gelu_fast.cuh"
#include "gemm_tma_epi.cuh"

namespace cpn {

struct MlpParams {
    int M;
    int num_tiles;        // tiles of 128 * CG rows
    const float* stats;   // [M, nparts, 2]
    int nparts;
    float inv_c, eps;
    const float* colsum;  // [4C] row sums of bf16(W1 * gamma)
    const float* bias1;   // [4C] b1 + W1 * beta
    const float* bias2;   // [C]
    const float* scale;   // [C] layer_scale
    float* x;             // [M, C] fp32 residual stream, updated in place
    long long* ts;        // -DCPN_DEV_KNOBS builds only (CPN_MLP_TS): clock64 trace of CTA 0's first tile, [2][32][8]
    int stagger;          // cycles by which CTA groups with one tile less than the others start late (see api.cu)
    int dbg;              // -DCPN_DEV_KNOBS builds only (CPN_MLP_DBG): 1 = E1 skips its math, 2 = no MMAs, 4 = no weight loads
};

template <int C, int CG>
struct MlpCfg {
    static constexpr int KB1 = (C + 63) / 64;            // k-blocks of G1
    static constexpr int NJ = C / 16;                    // hidden chunks of 64 columns (4C / 64)
    static constexpr int NSPLIT = C > 256 ? 2 : 1;       // G2 instructions per k-step (UMMA N <= 256)
    static constexpr int N2 = C / NSPLIT;
    static constexpr int X_BLOCK = GEMM_BM * 128;        // 128 rows x 64 bf16
    static constexpr int X_BYTES = KB1 * X_BLOCK;
    static constexpr int W1_KB_BYTES = (64 / CG) * 128;  // W1 rows of a chunk staged by this CTA, one k-block
    static constexpr int W1_BYTES = KB1 * W1_KB_BYTES;
    static constexpr int W2_NS_BYTES = (N2 / CG) * 128;  // W2 rows staged by this CTA for one G2 instruction
    static constexpr int W2_BYTES = NSPLIT * W2_NS_BYTES;
    static constexpr int W_STAGE = ((W1_BYTES > W2_BYTES ? W1_BYTES : W2_BYTES) + 1023) / 1024 * 1024;
    static constexpr int NH = 3;
    static constexpr int H_SLOT = GEMM_BM * 128;
    static constexpr int BAR_BYTES = 256;
    static constexpr int CB_BYTES = 8 * 4 * 256;          // per epilogue warp: 4 chunks x (32 colsum + 32 bias1)
    static constexpr int BUDGET = 227 * 1024 - 1024 - BAR_BYTES - CB_BYTES - X_BYTES - NH * H_SLOT;
    static constexpr int NW = (BUDGET / W_STAGE) > 6 ? 6 : (BUDGET / W_STAGE);
    static constexpr int TOTAL = X_BYTES + NH * H_SLOT + NW * W_STAGE + BAR_BYTES + CB_BYTES + 1024;
    static constexpr int ACC1_COL = C;
    static constexpr int TMEM_COLS = (C + 128 <= 256) ? 256 : 512;
    static constexpr int EW = 8;
    static constexpr int THREADS = 128 + 32 * EW;
    static_assert(C % 32 == 0 && C <= 384, "acc2 (C columns) and two 64-column acc1 buffers must fit in 512 TMEM columns");
    static_assert(N2 % (16 * CG) == 0 && N2 <= 256, "UMMA N constraint");
    static_assert(W1_KB_BYTES % 1024 == 0 && W2_NS_BYTES % 1024 == 0, "operand blocks must keep the swizzle atom alignment");
    static_assert(NW >= 2, "not enough shared memory for the weight ring");
    static_assert(NJ % 2 == 0, "acc1 buffer parity is tracked per tile");
};

template <int C, int CG>
__global__ void __launch_bounds__(MlpCfg<C, CG>::THREADS, 1)
mlp_fused_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                 const __grid_constant__ CUtensorMap tmap_w2, const MlpParams p) {
    using S = MlpCfg<C, CG>;
    constexpr int NW = S::NW, NH = S::NH, NJ = S::NJ, KB1 = S::KB1, EW = S::EW;
    const uint32_t rank = CG == 2 ? ptx::cluster_ctarank() : 0u;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* xs = smem;
    uint8_t* hs = xs + S::X_BYTES;
    uint8_t* ws = hs + NH * S::H_SLOT;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ws + NW * S::W_STAGE);
    uint64_t* w_full = bars;                 // [NW]
    uint64_t* w_empty = w_full + 6;          // [NW]
    uint64_t* h_full = w_empty + 6;          // [NH]
    uint64_t* h_empty = h_full + NH;         // [NH]
    uint64_t* a1_full = h_empty + NH;        // [2]
    uint64_t* a1_empty = a1_full + 2;        // [2]
    uint64_t* x_full = a1_empty + 2;
    uint64_t* x_empty = x_full + 1;
    uint64_t* a2_full = x_empty + 1;
    uint64_t* a2_empty = a2_full + 1;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(a2_empty + 1);
    uint8_t* cbs = reinterpret_cast<uint8_t*>(bars) + S::BAR_BYTES;  // [EW][4][2][32] floats
    static_assert((12 + 2 * NH + 8) * 8 + 8 <= S::BAR_BYTES, "barrier block too small");

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_x);
        ptx::prefetch_tmap(&tmap_w1);
        ptx::prefetch_tmap(&tmap_w2);
        for (int s = 0; s < NW; ++s) {
            ptx::mbar_init(&w_full[s], 1);
            ptx::mbar_init(&w_empty[s], 1);
        }
        for (int s = 0; s < NH; ++s) {
            ptx::mbar_init(&h_full[s], CG * EW);  // both CTAs' epilogue warps report to the leader
            ptx::mbar_init(&h_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&a1_full[s], 1);
            ptx::mbar_init(&a1_empty[s], CG * EW);
        }
        ptx::mbar_init(x_full, 1);
        ptx::mbar_init(x_empty, 1);
        ptx::mbar_init(a2_full, 1);
        ptx::mbar_init(a2_empty, CG * EW);
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
    // everything above is on-chip set-up; the operands come from the previous kernel of the stream
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (one thread per CTA)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            auto load = [&](void* dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar) {
                if (CG == 2) ptx::tma_load_2d_2sm(dst, m, c0, c1, bar);
                else ptx::tma_load_2d(dst, m, c0, c1, bar);
            };
            if (p.stagger != 0 && tile0 + ((p.num_tiles - 1) / tile_step) * tile_step >= p.num_tiles) {
                // this group has one tile less than the longest-running ones: starting late costs nothing and takes
                // its residual read-modify-write bursts (final epilogue) out of phase with theirs
                const long long t0 = clock64();
                const long long d = p.stagger >= 0 ? p.stagger : (-p.stagger) / 2 + ((-p.stagger) * ((tile0 * 5) & 15)) / 16;
                while (clock64() - t0 < d) {}
            }
            for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
                const int m0 = (tile * CG + static_cast<int>(rank)) * GEMM_BM;
                ptx::mbar_wait(x_empty, (it & 1) ^ 1);  // the last G1 of the previous tile has read X
                if (rank == 0) ptx::mbar_arrive_expect_tx(x_full, CG * S::X_BYTES);
#pragma unroll
                for (int kb = 0; kb < KB1; ++kb) load(xs + kb * S::X_BLOCK, &tmap_x, kb * 64, m0, x_full);
                for (int s = 0; s <= NJ; ++s) {
                    if (s < NJ) {  // W1 rows of chunk s, all k-blocks
                        ptx::mbar_wait(&w_empty[stage], phase ^ 1);
                        if (CPN_DBG(p.dbg, 4)) { if (rank == 0) ptx::mbar_arrive(&w_full[stage]); }
                        else if (rank == 0) ptx::mbar_arrive_expect_tx(&w_full[stage], CG * S::W1_BYTES);
                        uint8_t* dst = ws + stage * S::W_STAGE;
#pragma unroll
                        for (int kb = 0; kb < KB1; ++kb)
                            if (!(CPN_DBG(p.dbg, 4))) load(dst + kb * S::W1_KB_BYTES, &tmap_w1, kb * 64, s * 64 + static_cast<int>(rank) * (64 / CG),
                                 &w_full[stage]);
                        if (++stage == NW) { stage = 0; phase ^= 1; }
                    }
                    if (s >= 1) {  // W2 k-block of chunk s - 1, all outputs
                        ptx::mbar_wait(&w_empty[stage], phase ^ 1);
                        if (CPN_DBG(p.dbg, 4)) { if (rank == 0) ptx::mbar_arrive(&w_full[stage]); }
                        else if (rank == 0) ptx::mbar_arrive_expect_tx(&w_full[stage], CG * S::W2_BYTES);
                        uint8_t* dst = ws + stage * S::W_STAGE;
#pragma unroll
                        for (int ns = 0; ns < S::NSPLIT; ++ns)
                            if (!(CPN_DBG(p.dbg, 4))) load(dst + ns * S::W2_NS_BYTES, &tmap_w2, (s - 1) * 64,
                                 ns * S::N2 + static_cast<int>(rank) * (S::N2 / CG), &w_full[stage]);
                        if (++stage == NW) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (leader CTA, one thread)
        if (rank == 0) {  // the whole warp walks the loop (uniform control flow); one elected lane issues
            const uint32_t el = ptx::elect_one();
            constexpr uint32_t idesc1 = ptx::make_idesc_bf16(GEMM_BM * CG, 64);
            constexpr uint32_t idesc2 = ptx::make_idesc_bf16(GEMM_BM * CG, S::N2);
            auto mma = [&](uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
                ptx::umma_bf16_if<CG>(el, d, a, b, idesc, acc);
            };
            auto commit = [&](uint64_t* bar) { ptx::umma_commit_if<CG>(el, bar); };
            int stage = 0;
            uint32_t phase = 0;
            uint32_t g1 = 0;  // G1 chunks issued so far: acc1 buffer g1 & 1, barrier parity (g1 >> 1) & 1
            uint32_t hslot = 0, hphase = 0;
            int it = 0;
            for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
                ptx::mbar_wait(x_full, it & 1);
                ptx::tc_fence_after();
                for (int s = 0; s <= NJ; ++s) {
                    if (s < NJ) {
                        const uint32_t buf = g1 & 1;
                        const bool trace = CPN_TRACE(p.ts) && blockIdx.x == 0 && it == (p.dbg >> 8) && s < 32 && lane == 0;
                        long long* tsp = p.ts + 256 + s * 8;
                        if (trace) tsp[0] = clock64();
                        ptx::mbar_wait(&w_full[stage], phase);
                        if (trace) tsp[1] = clock64();
                        ptx::mbar_wait(&a1_empty[buf], ((g1 >> 1) & 1) ^ 1);
                        if (trace) tsp[2] = clock64();
                        ptx::tc_fence_after();
                        const uint32_t d = tmem_base + S::ACC1_COL + buf * 64;
                        const uint32_t wbase = ptx::smem_u32(ws + stage * S::W_STAGE);
#pragma unroll
                        for (int kb = 0; kb < KB1; ++kb) {
                            const uint64_t adesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(xs + kb * S::X_BLOCK));
                            const uint64_t bdesc = ptx::make_kmajor_sw128_desc(wbase + kb * S::W1_KB_BYTES);
                            constexpr int FULL = GEMM_BK / GEMM_UMMA_K;
                            const int ksteps = (kb == KB1 - 1 && (C % 64) != 0) ? (C % 64) / GEMM_UMMA_K : FULL;
#pragma unroll
                            for (int k = 0; k < FULL; ++k)
                                if (k < ksteps && !(CPN_DBG(p.dbg, 2))) mma(d, adesc + 2 * k, bdesc + 2 * k, idesc1, (kb | k) != 0);
                        }
                        commit(&w_empty[stage]);
                        commit(&a1_full[buf]);
                        if (s == NJ - 1) commit(x_empty);
                        if (++stage == NW) { stage = 0; phase ^= 1; }
                        ++g1;
                    }
                    if (s >= 1) {
                        const int j = s - 1;
                        const bool trace = CPN_TRACE(p.ts) && blockIdx.x == 0 && it == (p.dbg >> 8) && s < 32 && lane == 0;
                        long long* tsp = p.ts + 256 + s * 8;
                        if (trace) tsp[3] = clock64();
                        ptx::mbar_wait(&w_full[stage], phase);
                        if (trace) tsp[4] = clock64();
                        ptx::mbar_wait(&h_full[hslot], hphase);
                        if (trace) tsp[5] = clock64();
                        if (j == 0) ptx::mbar_wait(a2_empty, (it & 1) ^ 1);  // the previous tile's result has been read
                        ptx::tc_fence_after();
                        const uint64_t adesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(hs + hslot * S::H_SLOT));
                        const uint32_t wbase = ptx::smem_u32(ws + stage * S::W_STAGE);
#pragma unroll
                        for (int ns = 0; ns < S::NSPLIT; ++ns) {
                            const uint64_t bdesc = ptx::make_kmajor_sw128_desc(wbase + ns * S::W2_NS_BYTES);
#pragma unroll
                            for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k)
                                if (!(CPN_DBG(p.dbg, 2))) mma(tmem_base + ns * S::N2, adesc + 2 * k, bdesc + 2 * k, idesc2, (j | k) != 0);
                        }
                        commit(&w_empty[stage]);
                        commit(&h_empty[hslot]);
                        if (j == NJ - 1) commit(a2_full);
                        if (++stage == NW) { stage = 0; phase ^= 1; }
                        if (++hslot == NH) { hslot = 0; hphase ^= 1; }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue warps
        const int ew = warp - 4;
        const int q = warp & 3;    // TMEM lane quarter = 32-row block of the tile
        const int half = ew >> 2;  // which 32 of a chunk's 64 hidden columns / every other 32-column block of the output
        const uint32_t sw = static_cast<uint32_t>(lane & 7);
        const uint32_t row_off = static_cast<uint32_t>(q * 32 + lane) * 128u;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        auto arrive = [&](uint64_t* bar) {
            if (CG == 2) ptx::mbar_arrive_leader(bar);
            else ptx::mbar_arrive(bar);
        };
        // colsum / bias1 of this warp's 32 columns of a chunk: private ring of 4 slots x (32 + 32) floats, filled by
        // cp.async two chunks ahead (the L1 left next to ~220 KB of shared memory is too small to keep them)
        const uint32_t cb_addr = ptx::smem_u32(cbs) + static_cast<uint32_t>(ew) * (4 * 256);
        auto cb_issue = [&](int jj) {
            if (jj < NJ) {
                const int col = jj * 64 + half * 32 + lane;
                const uint32_t d = cb_addr + static_cast<uint32_t>(jj & 3) * 256 + static_cast<uint32_t>(lane) * 4;
                ptx::cp_async4(d, p.colsum + col);
                ptx::cp_async4(d + 128, p.bias1 + col);
            }
            ptx::cp_async_commit();
        };
        auto row_of = [&](int tile) { return (tile * CG + static_cast<int>(rank)) * GEMM_BM + q * 32 + lane; };
        float2 part[C / 32];  // partial LayerNorm sums of this lane's row, fetched one tile ahead
        auto load_stats = [&](int tile) {
            const int r = min(row_of(tile), p.M - 1);
            const float2* sp = reinterpret_cast<const float2*>(p.stats) + static_cast<size_t>(r) * p.nparts;
#pragma unroll
            for (int i = 0; i < C / 32; ++i) part[i] = __ldg(sp + i);
        };
        if (tile0 < p.num_tiles) load_stats(tile0);
        uint32_t g1 = 0, hslot = 0, hphase = 0;
        int it = 0;
        for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
            const int row = row_of(tile);
            const bool ttrace = CPN_TRACE(p.ts) && blockIdx.x == 0 && warp == 4 && lane == 0 && it < 8;
            if (ttrace) p.ts[480 + it * 4] = clock64();
            cb_issue(0);
            cb_issue(1);
            float ln_rs, ln_nm;
            {
                float s1 = 0.f, s2 = 0.f;
#pragma unroll
                for (int i = 0; i < C / 32; ++i) {
                    s1 += part[i].x;
                    s2 += part[i].y;
                }
                const float mean = s1 * p.inv_c;
                const float var = fmaxf(s2 * p.inv_c - mean * mean, 0.f);
                ln_rs = rsqrtf(var + p.eps);
                ln_nm = -mean * ln_rs;
            }
            const float2 ln_rs2 = make_float2(ln_rs, ln_rs), ln_nm2 = make_float2(ln_nm, ln_nm);
#pragma unroll 1
            for (int j = 0; j < NJ; ++j, ++g1) {
                const uint32_t buf = g1 & 1;
                uint32_t r0[32];
                const bool trace = CPN_TRACE(p.ts) && blockIdx.x == 0 && warp == 4 && lane == 0 && it == (p.dbg >> 8) && j < 32;
                long long* tsp = p.ts + j * 8;
                if (trace) tsp[0] = clock64();
                cb_issue(j + 2);
                ptx::mbar_wait(&a1_full[buf], (g1 >> 1) & 1);
                if (trace) tsp[1] = clock64();
                ptx::tc_fence_after();
                ptx::tmem_ld32(lane_addr + S::ACC1_COL + buf * 64 + half * 32, r0);
                ptx::tmem_ld_wait();
                if (trace) tsp[2] = clock64();
                ptx::tc_fence_before();
                ptx::cp_async_wait<2>();  // this chunk's coefficients have landed (this lane's part)
                __syncwarp();             // ... and every other lane's
                if (lane == 0) arrive(&a1_empty[buf]);
                if (trace) tsp[3] = clock64();
                ptx::mbar_wait(&h_empty[hslot], hphase ^ 1);  // the G2 that last read this slot has completed
                if (trace) tsp[4] = clock64();
                const uint32_t dst = ptx::smem_u32(hs + hslot * S::H_SLOT) + row_off;
                const uint32_t cb = cb_addr + static_cast<uint32_t>(j & 3) * 256;
                if (CPN_DBG(p.dbg, 1)) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        ptx::sts_u4(dst + ((static_cast<uint32_t>(4 * half + i) ^ sw) << 4),
                                    make_uint4(r0[8 * i], r0[8 * i + 2], r0[8 * i + 4], r0[8 * i + 6]));
                } else
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t* src = r0 + 8 * i;
                    const float4 s0 = ptx::lds_f4(cb + 32 * i);
                    const float4 s1 = ptx::lds_f4(cb + 32 * i + 16);
                    const float4 b0 = ptx::lds_f4(cb + 128 + 32 * i);
                    const float4 b1 = ptx::lds_f4(cb + 128 + 32 * i + 16);
                    float2 v0 = make_float2(__uint_as_float(src[0]), __uint_as_float(src[1]));
                    float2 v1 = make_float2(__uint_as_float(src[2]), __uint_as_float(src[3]));
                    float2 v2 = make_float2(__uint_as_float(src[4]), __uint_as_float(src[5]));
                    float2 v3 = make_float2(__uint_as_float(src[6]), __uint_as_float(src[7]));
                    v0 = ffma2(v0, ln_rs2, ffma2(ln_nm2, make_float2(s0.x, s0.y), make_float2(b0.x, b0.y)));
                    v1 = ffma2(v1, ln_rs2, ffma2(ln_nm2, make_float2(s0.z, s0.w), make_float2(b0.z, b0.w)));
                    v2 = ffma2(v2, ln_rs2, ffma2(ln_nm2, make_float2(s1.x, s1.y), make_float2(b1.x, b1.y)));
                    v3 = ffma2(v3, ln_rs2, ffma2(ln_nm2, make_float2(s1.z, s1.w), make_float2(b1.z, b1.w)));
                    const float2 y0 = gelu_sat2(v0);
                    const float2 y1 = gelu_sat2(v1);
                    const float2 y2 = gelu_sat2(v2);
                    const float2 y3 = gelu_sat2(v3);
                    ptx::sts_u4(dst + ((static_cast<uint32_t>(4 * half + i) ^ sw) << 4),
                                make_uint4(pack2(y0.x, y0.y), pack2(y1.x, y1.y), pack2(y2.x, y2.y), pack2(y3.x, y3.y)));
                }
                if (trace) tsp[5] = clock64();
                ptx::fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's shared-memory reads
                if (trace) tsp[6] = clock64();
                __syncwarp();
                if (lane == 0) arrive(&h_full[hslot]);
                if (trace) tsp[7] = clock64();
                if (++hslot == NH) { hslot = 0; hphase ^= 1; }
            }
            // ---- x[row, :] += scale * (acc2 + b2): this warp takes every other 32-column block; the residual of the
            // next block and the next tile's LayerNorm sums travel while the current block is finished
            if (ttrace) p.ts[480 + it * 4 + 1] = clock64();
            // ---- x[row, :] += scale * (acc2 + b2). This warp takes every other 32-column block of its 32 rows. The
            // accumulator (lane = row) is transposed through a 4 KB swizzled staging block (this warp's slice of the
            // idle H ring) so that the residual loads and the stores are coalesced: lane l handles the 16-byte
            // column group l & 7 of rows 4 i + (l >> 3). The residual of the next block travels one block ahead.
            constexpr int NBLK = C / 32;
            constexpr int NBMAX = (NBLK + 1) / 2;
            const uint32_t stg = ptx::smem_u32(hs) + static_cast<uint32_t>(ew) * 4096;
            const int row_w = (tile * CG + static_cast<int>(rank)) * GEMM_BM + q * 32;  // first row of this warp
            const int rsub = lane >> 3, cgrp = lane & 7;
            float* xbase = p.x + static_cast<size_t>(row_w) * C + cgrp * 4;
            float4 res[2][8];
            auto load_res = [&](int c, float4 (&dst)[8]) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = 4 * i + rsub;
                    dst[i] = row_w + r < p.M ? *reinterpret_cast<const float4*>(xbase + static_cast<size_t>(r) * C + c * 32)
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            };
            if (half < NBLK) load_res(half, res[0]);
            if (tile + tile_step < p.num_tiles) load_stats(tile + tile_step);
            ptx::mbar_wait(a2_full, it & 1);
            if (ttrace) p.ts[480 + it * 4 + 2] = clock64();
            ptx::tc_fence_after();
#pragma unroll
            for (int b = 0; b < NBMAX; ++b) {
                const int c = half + 2 * b;
                if (c < NBLK) {
                    if (c + 2 < NBLK) load_res(c + 2, res[(b + 1) & 1]);
                    const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias2 + c * 32 + cgrp * 4));
                    const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + c * 32 + cgrp * 4));
                    uint32_t r0[32];
                    ptx::tmem_ld32(lane_addr + c * 32, r0);
                    ptx::tmem_ld_wait();
                    if (c + 2 >= NBLK) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) arrive(a2_empty);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        ptx::sts_u4(stg + static_cast<uint32_t>(lane) * 128 + ((static_cast<uint32_t>(i) ^ sw) << 4),
                                    make_uint4(r0[4 * i], r0[4 * i + 1], r0[4 * i + 2], r0[4 * i + 3]));
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int r = 4 * i + rsub;
                        const float4 a4 = ptx::lds_f4(stg + static_cast<uint32_t>(r) * 128 +
                                                      ((static_cast<uint32_t>(cgrp) ^ static_cast<uint32_t>(r & 7)) << 4));
                        const float4 rr = res[b & 1][i];
                        float4 v;
                        v.x = fmaf(a4.x + b4.x, s4.x, rr.x);
                        v.y = fmaf(a4.y + b4.y, s4.y, rr.y);
                        v.z = fmaf(a4.z + b4.z, s4.z, rr.z);
                        v.w = fmaf(a4.w + b4.w, s4.w, rr.w);
                        if (row_w + r < p.M) *reinterpret_cast<float4*>(xbase + static_cast<size_t>(r) * C + c * 32) = v;
                    }
                    __syncwarp();  // the staging block is rewritten by the next block
                }
            }
            if (half >= NBLK) {  // C = 32: the second column group has no block
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) arrive(a2_empty);
            }
            // the staging blocks alias H slots that other epilogue warps write in the next tile's first chunks
            asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
            if (ttrace) p.ts[480 + it * 4 + 3] = clock64();
        }
        ptx::cp_async_wait<0>();
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
