// Persistent warp-specialised bf16 GEMM for sm_100a:  C[M,N] = A[M,K] * B[N,K]^T  (+ fused epilogue).
//
//   * A (activations, row = patch token) and B (weights, row = output channel) are K-major bf16 in HBM,
//     staged into 128B-swizzled shared memory by TMA (one elected producer thread).
//   * tcgen05.mma (cta_group::1, M=128, N=BN, K=16) issued by one thread, fp32 accumulators in TMEM,
//     two accumulator stages so the epilogue of tile i overlaps the main loop of tile i+1.
//   * 8 epilogue warps read TMEM with tcgen05.ld (each warp owns 32 lanes x half the columns), apply
//     bias / GELU / ReLU / layer-scale / residual and store bf16 and/or fp32 rows.
//
// Replaces, on the reference's hot path: nn.Linear C->4C + GELU and nn.Linear 4C->C * layer_scale + residual of
// torchvision CNBlock (torchvision/models/convnext.py:51-67), the 2x2 downsample convs (:146-150) and the
// 1x1/3x3/7x7 convs + folded BatchNorm + ReLU of features/resnet_features.py:97-117,199-210 (via im2col rows).
#pragma once
#include "ptx.cuh"

namespace cpn {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;       // 64 bf16 = 128 B = one swizzle row
constexpr int GEMM_UMMA_K = 16;
constexpr int GEMM_THREADS = 384; // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4..11 epilogue
constexpr int GEMM_EPI_WARPS = 8;

enum GemmAct : int { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };

struct GemmParams {
    int M, N, K;
    int num_m_blocks, num_n_blocks, num_k_blocks;
    const float* bias;              // [N] or null
    const float* scale;             // [N] layer-scale or null (applied after activation)
    const float* res_f32;           // [M,N] fp32 residual or null (added after scale)
    const __nv_bfloat16* res_bf16;  // [M,N] bf16 residual or null
    float* out_f32;                 // [M,N] or null
    __nv_bfloat16* out_bf16;        // [M,N] or null
    __nv_bfloat16* out_bf16_lo;     // [M,N] or null: low part of the hi/lo split (fp32 mode)
    int relu_after_res;             // ResNet block end: relu(acc + bias + residual)
};

template <int BN, bool SPLIT>
struct GemmSmem {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = BN * GEMM_BK * 2;
    // fp32 mode (SPLIT) stages the hi and lo bf16 planes of both operands: [A_hi | B_hi | A_lo | B_lo]
    static constexpr int STAGE_BYTES = (A_BYTES + B_BYTES) * (SPLIT ? 2 : 1);
    // Fill what a 227 KB CTA offers, at most 8 stages.
    static constexpr int STAGES = ((220 * 1024) / STAGE_BYTES) > 8 ? 8 : ((220 * 1024) / STAGE_BYTES);
    static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
    static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16;
    static constexpr int TOTAL = TILE_BYTES + BAR_BYTES + 1024;  // + slack for the 1024-B alignment
    static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                     : (2 * BN <= 256) ? 256 : 512;
};

__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

template <int BN, int ACT, bool SPLIT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ CUtensorMap tmap_a_lo, const __grid_constant__ CUtensorMap tmap_b_lo,
                         const GemmParams p) {
    using S = GemmSmem<BN, SPLIT>;
    constexpr int STAGES = S::STAGES;
    static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N constraint for M=128");
    static_assert((BN / 2) % 16 == 0 || BN == 16, "epilogue splits the columns in two halves of 16-col chunks");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    auto stage_a = [&](int s) { return smem + s * S::STAGE_BYTES; };
    auto stage_b = [&](int s) { return smem + s * S::STAGE_BYTES + S::A_BYTES; };
    auto stage_a_lo = [&](int s) { return smem + s * S::STAGE_BYTES + S::A_BYTES + S::B_BYTES; };
    auto stage_b_lo = [&](int s) { return smem + s * S::STAGE_BYTES + 2 * S::A_BYTES + S::B_BYTES; };
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::TILE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_b);
        if (SPLIT) {
            ptx::prefetch_tmap(&tmap_a_lo);
            ptx::prefetch_tmap(&tmap_b_lo);
        }
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tmem_full_bar[s], 1);
            ptx::mbar_init(&tmem_empty_bar[s], GEMM_EPI_WARPS);
        }
        ptx::fence_mbar_init();
    } else if (warp == 2) {
        ptx::tmem_alloc(tmem_holder, S::TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m_blk = tile / p.num_n_blocks;
                const int n_blk = tile % p.num_n_blocks;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                    ptx::tma_load_2d(stage_a(stage), &tmap_a, kb * GEMM_BK, m_blk * GEMM_BM, &full_bar[stage]);
                    ptx::tma_load_2d(stage_b(stage), &tmap_b, kb * GEMM_BK, n_blk * BN, &full_bar[stage]);
                    if (SPLIT) {
                        ptx::tma_load_2d(stage_a_lo(stage), &tmap_a_lo, kb * GEMM_BK, m_blk * GEMM_BM, &full_bar[stage]);
                        ptx::tma_load_2d(stage_b_lo(stage), &tmap_b_lo, kb * GEMM_BK, n_blk * BN, &full_bar[stage]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (single thread)
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(GEMM_BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                ptx::mbar_wait(&tmem_empty_bar[as], aphase ^ 1);  // epilogue has drained this accumulator
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + as * BN;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_a(stage)));
                    const uint64_t bdesc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_b(stage)));
                    if (SPLIT) {
                        // x*w ~= x_lo*w_hi + x_hi*w_lo + x_hi*w_hi, small terms first
                        const uint64_t adesc_lo = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_a_lo(stage)));
                        const uint64_t bdesc_lo = ptx::make_kmajor_sw128_desc(ptx::smem_u32(stage_b_lo(stage)));
#pragma unroll
                        for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k) {
                            ptx::umma_bf16(tmem_d, adesc_lo + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                            ptx::umma_bf16(tmem_d, adesc + 2 * k, bdesc_lo + 2 * k, idesc, 1);
                            ptx::umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, 1);
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < GEMM_BK / GEMM_UMMA_K; ++k) {
                            // advance 32 B (= 16 bf16) inside the 128-B swizzle row: +2 in 16-byte units
                            ptx::umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                        }
                    }
                    ptx::umma_commit(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
                    if (kb == p.num_k_blocks - 1) ptx::umma_commit(&tmem_full_bar[as]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue (8 warps)
        const int ew = warp - 4;
        const int q = warp & 3;            // TMEM lane quarter this warp may access
        const int half = ew >> 2;          // which half of the BN columns
        constexpr int COLS_PER_WARP = BN / 2;
        constexpr int CHUNKS = COLS_PER_WARP / 16;
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m_blk = tile / p.num_n_blocks;
            const int n_blk = tile % p.num_n_blocks;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            ptx::mbar_wait(&tmem_full_bar[as], aphase);
            ptx::tc_fence_after();
            const int row = m_blk * GEMM_BM + q * 32 + lane;
            const bool row_ok = row < p.M;
            const uint32_t taddr_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + half * COLS_PER_WARP;
            const size_t row_off = static_cast<size_t>(row) * p.N;
#pragma unroll 1
            for (int c = 0; c < CHUNKS; ++c) {
                uint32_t r[16];
                ptx::tmem_ld16(taddr_row + c * 16, r);
                ptx::tmem_ld_wait();
                const int col0 = n_blk * BN + half * COLS_PER_WARP + c * 16;
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                if (p.bias != nullptr) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                        v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
                    }
                }
                if (ACT == ACT_GELU) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = gelu_erf(v[j]);
                } else if (ACT == ACT_RELU) {
                    if (!p.relu_after_res) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.0f);
                    }
                }
                if (p.scale != nullptr) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + col0 + j));
                        v[j] *= s4.x; v[j + 1] *= s4.y; v[j + 2] *= s4.z; v[j + 3] *= s4.w;
                    }
                }
                if (row_ok) {
                    if (p.res_f32 != nullptr) {
                        const float4* rp = reinterpret_cast<const float4*>(p.res_f32 + row_off + col0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float4 r4 = rp[j];
                            v[4 * j] += r4.x; v[4 * j + 1] += r4.y; v[4 * j + 2] += r4.z; v[4 * j + 3] += r4.w;
                        }
                    }
                    if (p.res_bf16 != nullptr) {
                        const uint4* rp = reinterpret_cast<const uint4*>(p.res_bf16 + row_off + col0);
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const uint4 u = rp[j];
                            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                v[8 * j + 2 * t] += __uint_as_float(w[t] << 16);
                                v[8 * j + 2 * t + 1] += __uint_as_float(w[t] & 0xFFFF0000u);
                            }
                        }
                    }
                    if (ACT == ACT_RELU && p.relu_after_res) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.0f);
                    }
                    if (p.out_f32 != nullptr) {
                        float4* op = reinterpret_cast<float4*>(p.out_f32 + row_off + col0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) op[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (p.out_bf16 != nullptr) {
                        uint32_t h[8];
                        uint32_t l[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * j]);
                            const __nv_bfloat16 h1 = __float2bfloat16_rn(v[2 * j + 1]);
                            h[j] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) |
                                   (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
                            if (p.out_bf16_lo != nullptr) {
                                const __nv_bfloat16 l0 = __float2bfloat16_rn(v[2 * j] - __bfloat162float(h0));
                                const __nv_bfloat16 l1 = __float2bfloat16_rn(v[2 * j + 1] - __bfloat162float(h1));
                                l[j] = static_cast<uint32_t>(__bfloat16_as_ushort(l0)) |
                                       (static_cast<uint32_t>(__bfloat16_as_ushort(l1)) << 16);
                            }
                        }
                        uint4* op = reinterpret_cast<uint4*>(p.out_bf16 + row_off + col0);
                        op[0] = make_uint4(h[0], h[1], h[2], h[3]);
                        op[1] = make_uint4(h[4], h[5], h[6], h[7]);
                        if (p.out_bf16_lo != nullptr) {
                            uint4* lp = reinterpret_cast<uint4*>(p.out_bf16_lo + row_off + col0);
                            lp[0] = make_uint4(l[0], l[1], l[2], l[3]);
                            lp[1] = make_uint4(l[4], l[5], l[6], l[7]);
                        }
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty_bar[as]);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, S::TMEM_COLS);
    }
}

}  // namespace cpn
