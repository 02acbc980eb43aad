// Prototype head of PIP-Net / Count-PIP-Net on channels-last features.
//   addon_conv1x1_kernel   optional 1x1 add-on conv C -> P (+bias)                     pipnet/pipnet.py:102, count_pipnet.py:378
//   softmax_pool_kernel    per-patch softmax over P (or hard Gumbel-softmax), then the spatial max (PIP-Net,
//                          pipnet/pipnet.py:96-108) or spatial sum (Count-PIP-Net, count_pipnet.py:88) as per-chunk partials;
//                          the patch x prototype map is written (NCHW) only when the caller passes a buffer
//   head_tail_kernel       chunk reduction, inference threshold (pipnet.py:36) or round/clamp (count_pipnet.py:90-97),
//                          count encoding (count_pipnet_utils.py:141-185, 371-385, 429-439, 461-462, 498-519) and the
//                          ReLU-weight classifier (pipnet.py:71)
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

enum PoolMode : int { POOL_MAX = 0, POOL_SUM = 1 };

// ------------------------------------------------------------------------------------------------ add-on conv
// Warp per patch token, lanes split the C input channels, weights [P][C] staged in shared memory.
__global__ void __launch_bounds__(256)
addon_conv1x1_kernel(const float* __restrict__ feat, const float* __restrict__ w, const float* __restrict__ bias,
                     float* __restrict__ z, long long T, int C, int P) {
    extern __shared__ float s_w[];  // [P][C]
    for (int i = threadIdx.x; i < P * C; i += blockDim.x) s_w[i] = w[i];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (long long t = static_cast<long long>(blockIdx.x) * wpb + warp; t < T; t += static_cast<long long>(gridDim.x) * wpb) {
        const float* f = feat + t * C;
        for (int p = 0; p < P; ++p) {
            float a = 0.f;
            for (int c = lane; c < C; c += 32) a = fmaf(__ldg(f + c), s_w[p * C + c], a);
            a = warp_sum(a);
            if (lane == 0) z[t * P + p] = a + (bias != nullptr ? bias[p] : 0.f);
        }
    }
}

// ------------------------------------------------------------------------------------------------ fused add-on head
// Add-on 1x1 conv (C -> P, P small) + softmax / hard Gumbel-softmax over P + spatial max or sum, in ONE kernel: the patch
// logits z [B*HW, P] never exist in HBM (pipnet/pipnet.py:102-108, count_pipnet.py:378-381 + 88). grid = (chunks, B); a warp
// takes 32 tokens at a time: the feature rows are staged through a padded shared-memory tile in 64-channel slices (coalesced
// 128-byte loads, conflict-free column reads), then LANE = TOKEN: P dot products in registers with the weights broadcast
// from shared memory, the softmax in registers, and the per-prototype pooling as a warp reduction over the 32 tokens.
// gumbel / map are [B, P, HW] (the reference's NCHW layout): coalesced across the tokens of a warp.
constexpr int ADDON_WARPS = 4;
template <int P>
__global__ void __launch_bounds__(32 * ADDON_WARPS)
addon_head_kernel(const float* __restrict__ feat, const float* __restrict__ w, const float* __restrict__ bias,
                  const float* __restrict__ gumbel, float tau, int hard, int mode, float* __restrict__ map,
                  float* __restrict__ partial, int* __restrict__ partial_arg, int HW, int C, int chunks) {
    constexpr int CK = 64, TS = CK + 1, WARPS = ADDON_WARPS;
    extern __shared__ float s_head[];
    float* s_w = s_head;                          // [C][P]: w transposed, so that a token's P taps of channel c are contiguous
    float* s_tile = s_w + static_cast<size_t>(C) * P;   // [WARPS][32][TS]
    float* s_pool = s_tile + WARPS * 32 * TS;     // [WARPS][P]
    int* s_arg = reinterpret_cast<int*>(s_pool + WARPS * P);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.y, chunk = blockIdx.x;
    for (int i = threadIdx.x; i < C * P; i += blockDim.x) {
        const int pp = i / C, c = i - pp * C;
        s_w[c * P + pp] = w[i];
    }
    __syncthreads();
    const int per_chunk = (HW + chunks - 1) / chunks;
    const int t0 = chunk * per_chunk, t1 = min(HW, t0 + per_chunk);
    const bool want_arg = partial_arg != nullptr && mode == POOL_MAX;
    float pool[P];
    int parg[P];
#pragma unroll
    for (int pp = 0; pp < P; ++pp) { pool[pp] = 0.f; parg[pp] = 0x7fffffff; }
    float* tile = s_tile + warp * 32 * TS;
    for (int tb = t0 + warp * 32; tb < t1; tb += WARPS * 32) {
        const int t = tb + lane;
        const bool live = t < t1;
        float acc[P];
#pragma unroll
        for (int pp = 0; pp < P; ++pp) acc[pp] = 0.f;
        for (int c0 = 0; c0 < C; c0 += CK) {
            __syncwarp();
            // stage rows tb .. tb + 31, channels c0 .. c0 + 63: two 128-byte loads per token, 16 loads in flight
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 8) {
                float v[8][2];
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) {
                    const int tr = min(tb + r0 + rr, t1 - 1);
                    const float* f = feat + (static_cast<size_t>(b) * HW + tr) * C + c0;
                    v[rr][0] = __ldg(f + lane);
                    v[rr][1] = __ldg(f + 32 + lane);
                }
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) {
                    tile[(r0 + rr) * TS + lane] = v[rr][0];
                    tile[(r0 + rr) * TS + 32 + lane] = v[rr][1];
                }
            }
            __syncwarp();
            const float* row = tile + lane * TS;
#pragma unroll 4
            for (int c = 0; c < CK; ++c) {
                const float x = row[c];
                const float4* wv = reinterpret_cast<const float4*>(s_w + static_cast<size_t>(c0 + c) * P);
#pragma unroll
                for (int q = 0; q < P / 4; ++q) {
                    const float4 w4 = wv[q];
                    acc[4 * q] = fmaf(x, w4.x, acc[4 * q]);
                    acc[4 * q + 1] = fmaf(x, w4.y, acc[4 * q + 1]);
                    acc[4 * q + 2] = fmaf(x, w4.z, acc[4 * q + 2]);
                    acc[4 * q + 3] = fmaf(x, w4.w, acc[4 * q + 3]);
                }
            }
        }
        // logits (+ Gumbel noise), softmax over the P prototypes of this token
        const size_t gbase = static_cast<size_t>(b) * P * HW + (live ? t : t1 - 1);
        float m = -INFINITY;
        int am = 0;
#pragma unroll
        for (int pp = 0; pp < P; ++pp) {
            float v = acc[pp] + (bias != nullptr ? __ldg(bias + pp) : 0.f);
            if (gumbel != nullptr) v = (v + __ldg(gumbel + gbase + static_cast<size_t>(pp) * HW)) / tau;
            acc[pp] = v;
            if (v > m) { m = v; am = pp; }  // first index on ties, as torch.max
        }
        float ssum = 0.f;
#pragma unroll
        for (int pp = 0; pp < P; ++pp) {
            acc[pp] = expf(acc[pp] - m);
            ssum += acc[pp];
        }
        const float inv_s = 1.0f / ssum;
#pragma unroll
        for (int pp = 0; pp < P; ++pp) {
            float y = acc[pp] * inv_s;
            if (hard) y = ((pp == am ? 1.0f : 0.0f) - y) + y;  // y_hard - y_soft + y_soft (count_pipnet_utils.py:33-38)
            if (live && map != nullptr) map[gbase + static_cast<size_t>(pp) * HW] = y;
            if (!live) y = 0.f;  // probabilities are >= 0: neutral for max and sum
            // pooling over the 32 tokens of the warp (token order; strict comparisons keep the lowest token on ties)
            if (mode == POOL_MAX) {
                float bv = y;
                int bt = live ? t : 0x7fffffff;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                    const int ot = __shfl_xor_sync(0xffffffffu, bt, o);
                    if (ov > bv || (ov == bv && ot < bt)) { bv = ov; bt = ot; }
                }
                if (bv > pool[pp]) { pool[pp] = bv; parg[pp] = bt; }
            } else {
                pool[pp] += warp_sum(y);
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int pp = 0; pp < P; ++pp) {
            s_pool[warp * P + pp] = pool[pp];
            s_arg[warp * P + pp] = parg[pp];
        }
    }
    __syncthreads();
    if (threadIdx.x < P) {
        const int pp = threadIdx.x;
        float a = s_pool[pp];
        int ai = s_arg[pp];
        for (int wv = 1; wv < WARPS; ++wv) {
            const float o = s_pool[wv * P + pp];
            const int oi = s_arg[wv * P + pp];
            if (mode == POOL_MAX) {
                if (o > a || (o == a && oi < ai)) { a = o; ai = oi; }
            } else {
                a += o;
            }
        }
        partial[(static_cast<size_t>(b) * chunks + chunk) * P + pp] = a;
        if (want_arg) partial_arg[(static_cast<size_t>(b) * chunks + chunk) * P + pp] = ai;
    }
}

// ------------------------------------------------------------------------------------------------ softmax + pool

// grid = (chunks, B); each CTA owns a contiguous range of one image's patch tokens, one warp per token at a time.
// z:       [B, HW, P] fp32 logits (token-major)
// gumbel:  [B, P, H*W] fp32 noise in the reference's NCHW layout, or null
// map:     [B, P, H*W] fp32 output (NCHW like the reference's proto_features), or null
// partial: [B, chunks, P]
__global__ void __launch_bounds__(256)
softmax_pool_kernel(const float* __restrict__ z, const float* __restrict__ gumbel, float tau, int hard, int mode,
                    float* __restrict__ map, float* __restrict__ partial, int* __restrict__ partial_arg, int HW, int P,
                    int chunks) {
    extern __shared__ float s_pool[];  // [warps][P] (+ [warps][P] arg-max tokens when partial_arg is given)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const int b = blockIdx.y;
    const int chunk = blockIdx.x;
    const int per_chunk = (HW + chunks - 1) / chunks;
    const int t0 = chunk * per_chunk;
    const int t1 = min(HW, t0 + per_chunk);
    float* my_pool = s_pool + warp * P;
    int* s_arg = reinterpret_cast<int*>(s_pool + wpb * P);
    int* my_arg = s_arg + warp * P;
    const bool want_arg = partial_arg != nullptr && mode == POOL_MAX;
    for (int p = lane; p < P; p += 32) {
        my_pool[p] = 0.f;  // probabilities are >= 0, so 0 is neutral for max and sum
        if (want_arg) my_arg[p] = 0x7fffffff;
    }
    __syncwarp();
    for (int t = t0 + warp; t < t1; t += wpb) {
        const float* zr = z + (static_cast<size_t>(b) * HW + t) * P;
        const float* gr = gumbel != nullptr ? gumbel + static_cast<size_t>(b) * P * HW + t : nullptr;
        // pass 1: row max (and arg-max, first index on ties)
        float m = -INFINITY;
        int am = 0x7fffffff;
        for (int p = lane; p < P; p += 32) {
            float v = __ldg(zr + p);
            if (gr != nullptr) v = (v + __ldg(gr + static_cast<size_t>(p) * HW)) / tau;
            if (v > m) { m = v; am = p; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, m, o);
            const int oa = __shfl_xor_sync(0xffffffffu, am, o);
            if (om > m || (om == m && oa < am)) { m = om; am = oa; }
        }
        // pass 2: sum of exponentials
        float s = 0.f;
        for (int p = lane; p < P; p += 32) {
            float v = __ldg(zr + p);
            if (gr != nullptr) v = (v + __ldg(gr + static_cast<size_t>(p) * HW)) / tau;
            s += expf(v - m);
        }
        s = warp_sum(s);
        const float inv_s = 1.0f / s;
        // pass 3: probabilities -> pool (+ optional map)
        for (int p = lane; p < P; p += 32) {
            float v = __ldg(zr + p);
            if (gr != nullptr) v = (v + __ldg(gr + static_cast<size_t>(p) * HW)) / tau;
            float y = expf(v - m) * inv_s;
            if (hard) y = ((p == am ? 1.0f : 0.0f) - y) + y;  // y_hard - y_soft + y_soft (straight-through form)
            if (mode == POOL_MAX) {
                if (y > my_pool[p]) {  // strict: the first (lowest) token keeps a tie, as torch.max does
                    my_pool[p] = y;
                    if (want_arg) my_arg[p] = t;
                }
            } else {
                my_pool[p] += y;
            }
            if (map != nullptr) map[(static_cast<size_t>(b) * P + p) * HW + t] = y;
        }
    }
    __syncthreads();
    float* out = partial + (static_cast<size_t>(b) * chunks + chunk) * P;
    for (int p = threadIdx.x; p < P; p += blockDim.x) {
        float a = s_pool[p];
        int ai = want_arg ? s_arg[p] : 0;
        for (int wv = 1; wv < wpb; ++wv) {
            const float o = s_pool[wv * P + p];
            if (mode == POOL_MAX) {
                const int oi = want_arg ? s_arg[wv * P + p] : 0;
                if (o > a || (o == a && oi < ai)) { a = o; ai = oi; }
            } else {
                a += o;
            }
        }
        out[p] = a;
        if (want_arg) partial_arg[(static_cast<size_t>(b) * chunks + chunk) * P + p] = ai;
    }
}

// Per (image, prototype): token index (h * W + w) of the spatial maximum, from the per-chunk maxima and arg-maxima
// (chunks are in token order, so the first chunk keeps a tie; 0x7fffffff marks "no token seen": all-zero rows -> 0).
__global__ void pool_argmax_reduce_kernel(const float* __restrict__ partial, const int* __restrict__ partial_arg,
                                          int chunks, int* __restrict__ arg_out, int P) {
    const int b = blockIdx.y;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const size_t base = static_cast<size_t>(b) * chunks * P + p;
    float a = partial[base];
    int ai = partial_arg[base];
    for (int c = 1; c < chunks; ++c) {
        const float o = partial[base + static_cast<size_t>(c) * P];
        const int oi = partial_arg[base + static_cast<size_t>(c) * P];
        if (o > a || (o == a && oi < ai)) { a = o; ai = oi; }
    }
    arg_out[static_cast<size_t>(b) * P + p] = ai == 0x7fffffff ? 0 : ai;
}

// Running top-k images per prototype (util/vis_pipnet.py:226-271: an image enters prototype p's list while the list is
// shorter than k or when its score is strictly greater than the last entry's; the list stays sorted by descending
// score, equal scores in arrival order). One thread per prototype walks the batch in image order, so the result is
// what the reference's one-image-at-a-time loop builds. Lists are [k, P] (coalesced across the threads of a warp).
__global__ void topk_update_kernel(const float* __restrict__ pooled, const int* __restrict__ argmax, int B, int P, int k,
                                   long long image_base, float* __restrict__ val, long long* __restrict__ img,
                                   int* __restrict__ tok, int* __restrict__ cnt) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int n = cnt[p];
    for (int b = 0; b < B; ++b) {
        const float s = pooled[static_cast<size_t>(b) * P + p];
        if (n == k && !(s > val[static_cast<size_t>(k - 1) * P + p])) continue;
        int i = n < k ? n : k - 1;  // slot that opens up (append, or the dropped last entry)
        while (i > 0 && val[static_cast<size_t>(i - 1) * P + p] < s) {
            val[static_cast<size_t>(i) * P + p] = val[static_cast<size_t>(i - 1) * P + p];
            img[static_cast<size_t>(i) * P + p] = img[static_cast<size_t>(i - 1) * P + p];
            tok[static_cast<size_t>(i) * P + p] = tok[static_cast<size_t>(i - 1) * P + p];
            --i;
        }
        val[static_cast<size_t>(i) * P + p] = s;
        img[static_cast<size_t>(i) * P + p] = image_base + b;
        tok[static_cast<size_t>(i) * P + p] = argmax[static_cast<size_t>(b) * P + p];
        if (n < k) ++n;
    }
    cnt[p] = n;
}

// Plain softmax WITH the patch x prototype map written (cfg5: the caller asks for the map for visualisation). The map is
// [B, P, HW] (the reference's NCHW proto_features), i.e. prototype-major, while a token's probabilities are produced
// token-major: writing them straight out is one 4-byte store per 128-byte line (the generic kernel: 0.46 ms at cfg5).
// Here the CTA stages 32 tokens x P probabilities in shared memory (float4 slots XOR-swizzled by token & 7, conflict-free
// both ways) and writes every prototype's 32 tokens as one 128-byte segment. Pooling as in the register kernel.
template <int NV>
__global__ void __launch_bounds__(256)
softmax_pool_map_kernel(const float* __restrict__ z, int mode, float* __restrict__ map, float* __restrict__ partial, int HW,
                        int chunks) {
    constexpr int P = NV * 128, F = P / 4, WARPS = 8, TT = 32;
    extern __shared__ float s_map[];         // [TT][P] swizzled tile, then [WARPS][P] pooling partials
    float4* tile = reinterpret_cast<float4*>(s_map);
    float* s_pool = s_map + TT * P;
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int per_chunk = (HW + chunks - 1) / chunks;
    const int t0 = chunk * per_chunk, t1 = min(HW, t0 + per_chunk);
    constexpr float LOG2E = 1.4426950408889634f;
    float4 pool[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) pool[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* zb = reinterpret_cast<const float4*>(z + static_cast<size_t>(b) * HW * P) + lane;
    float* mb = map + static_cast<size_t>(b) * P * HW;
    for (int tb = t0; tb < t1; tb += TT) {
        // ---- each warp: 4 tokens of the tile, probabilities to the pool and to the shared tile
#pragma unroll 1
        for (int u = 0; u < TT / WARPS; ++u) {
            const int tk = warp * (TT / WARPS) + u, t = tb + tk;
            if (t >= t1) break;  // warp-uniform
            float4 v[NV];
            const float4* zr = zb + static_cast<size_t>(t) * F;
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = __ldg(zr + 32 * i);
            float m = -INFINITY;
#pragma unroll
            for (int i = 0; i < NV; ++i) m = fmaxf(fmaxf(fmaxf(v[i].x, v[i].y), fmaxf(v[i].z, v[i].w)), m);
            m = warp_max(m);
            const float ms = m * LOG2E;
            float ssum = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                v[i].x = exp2f(fmaf(v[i].x, LOG2E, -ms));
                v[i].y = exp2f(fmaf(v[i].y, LOG2E, -ms));
                v[i].z = exp2f(fmaf(v[i].z, LOG2E, -ms));
                v[i].w = exp2f(fmaf(v[i].w, LOG2E, -ms));
                ssum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
            }
            ssum = warp_sum(ssum);
            const float inv_s = 1.0f / ssum;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const float4 y = make_float4(v[i].x * inv_s, v[i].y * inv_s, v[i].z * inv_s, v[i].w * inv_s);
                if (mode == POOL_MAX) {
                    pool[i].x = fmaxf(pool[i].x, y.x); pool[i].y = fmaxf(pool[i].y, y.y);
                    pool[i].z = fmaxf(pool[i].z, y.z); pool[i].w = fmaxf(pool[i].w, y.w);
                } else {
                    pool[i].x += y.x; pool[i].y += y.y; pool[i].z += y.z; pool[i].w += y.w;
                }
                tile[tk * F + ((lane + 32 * i) ^ (tk & 7))] = y;
            }
        }
        __syncthreads();
        // ---- every prototype's 32 tokens as one 128-byte segment: lane = token, a warp takes every 8th float4 column
        {
            const int t = tb + lane;
            if (t < t1) {
                for (int f = warp; f < F; f += WARPS) {
                    const float4 y = tile[lane * F + (f ^ (lane & 7))];
                    float* o = mb + static_cast<size_t>(4 * f) * HW + t;
                    o[0] = y.x;
                    o[HW] = y.y;
                    o[2 * static_cast<size_t>(HW)] = y.z;
                    o[3 * static_cast<size_t>(HW)] = y.w;
                }
            }
        }
        __syncthreads();
    }
    float4* mine = reinterpret_cast<float4*>(s_pool + warp * P) + lane;
#pragma unroll
    for (int i = 0; i < NV; ++i) mine[32 * i] = pool[i];
    __syncthreads();
    float* out = partial + (static_cast<size_t>(b) * chunks + chunk) * P;
    for (int pp = threadIdx.x; pp < P; pp += blockDim.x) {
        float a = s_pool[pp];
        for (int wv = 1; wv < WARPS; ++wv) {
            const float o = s_pool[wv * P + pp];
            a = mode == POOL_MAX ? fmaxf(a, o) : a + o;
        }
        out[pp] = a;
    }
}

// Register-resident variant for the common case (plain softmax, no map, P a multiple of 128 up to 1024): each lane
// owns the same 4*NV prototypes of every token (NV float4 per lane, loaded once, coalesced 512 B per warp and
// instruction), so the row max / sum are two warp reductions, the probabilities never leave registers and the running
// spatial max / sum stays in registers too; two tokens are in flight per warp. Same grid and partial layout as above.
// exp() is ex2.approx of a pre-scaled argument: relative error 2^-22, well inside the fp32-mode parity bar (1e-4).
// U = tokens in flight per warp (2; 1 for P > 1024, where a token alone is 16 independent 512-byte loads and 64 registers).
template <int NV, bool ARG, int U = 2>
__global__ void __launch_bounds__(256)
softmax_pool_reg_kernel(const float* __restrict__ z, int mode, float* __restrict__ partial,
                        int* __restrict__ partial_arg, int HW, int chunks) {
    constexpr int P = NV * 128;
    extern __shared__ float s_pool[];  // [warps][P]
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const int b = blockIdx.y;
    const int chunk = blockIdx.x;
    const int per_chunk = (HW + chunks - 1) / chunks;
    const int t0 = chunk * per_chunk;
    const int t1 = min(HW, t0 + per_chunk);
    constexpr float LOG2E = 1.4426950408889634f;
    float4 pool[NV];
    int4 parg[ARG ? NV : 1];  // token of the running maximum (max pooling with ARG)
#pragma unroll
    for (int i = 0; i < NV; ++i) pool[i] = make_float4(0.f, 0.f, 0.f, 0.f);  // probabilities are >= 0
#pragma unroll
    for (int i = 0; i < (ARG ? NV : 1); ++i) parg[i] = make_int4(0x7fffffff, 0x7fffffff, 0x7fffffff, 0x7fffffff);
    const float4* zb = reinterpret_cast<const float4*>(z + static_cast<size_t>(b) * HW * P) + lane;
    for (int t = t0 + U * warp; t < t1; t += U * wpb) {
        const bool two = t + 1 < t1;
        float4 v[U][NV];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const float4* zr = zb + static_cast<size_t>(u == 1 && two ? t + 1 : t) * (P / 4);
#pragma unroll
            for (int i = 0; i < NV; ++i) v[u][i] = __ldg(zr + 32 * i);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float m = -INFINITY;
#pragma unroll
            for (int i = 0; i < NV; ++i) m = fmaxf(fmaxf(fmaxf(v[u][i].x, v[u][i].y), fmaxf(v[u][i].z, v[u][i].w)), m);
            m = warp_max(m);
            const float ms = m * LOG2E;
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                v[u][i].x = exp2f(fmaf(v[u][i].x, LOG2E, -ms));
                v[u][i].y = exp2f(fmaf(v[u][i].y, LOG2E, -ms));
                v[u][i].z = exp2f(fmaf(v[u][i].z, LOG2E, -ms));
                v[u][i].w = exp2f(fmaf(v[u][i].w, LOG2E, -ms));
                s += (v[u][i].x + v[u][i].y) + (v[u][i].z + v[u][i].w);
            }
            s = warp_sum(s);
            const float inv_s = (u == 1 && !two) ? 0.f : 1.0f / s;  // the duplicate of an odd tail contributes nothing
            if (mode == POOL_MAX && ARG) {
                const int tcur = t + u;  // strict comparisons: the lowest token keeps a tie (tokens come in order)
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    const float yx = v[u][i].x * inv_s, yy = v[u][i].y * inv_s, yz = v[u][i].z * inv_s, yw = v[u][i].w * inv_s;
                    if (yx > pool[i].x) { pool[i].x = yx; parg[ARG ? i : 0].x = tcur; }
                    if (yy > pool[i].y) { pool[i].y = yy; parg[ARG ? i : 0].y = tcur; }
                    if (yz > pool[i].z) { pool[i].z = yz; parg[ARG ? i : 0].z = tcur; }
                    if (yw > pool[i].w) { pool[i].w = yw; parg[ARG ? i : 0].w = tcur; }
                }
            } else if (mode == POOL_MAX) {
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    pool[i].x = fmaxf(pool[i].x, v[u][i].x * inv_s);
                    pool[i].y = fmaxf(pool[i].y, v[u][i].y * inv_s);
                    pool[i].z = fmaxf(pool[i].z, v[u][i].z * inv_s);
                    pool[i].w = fmaxf(pool[i].w, v[u][i].w * inv_s);
                }
            } else {
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    pool[i].x = fmaf(v[u][i].x, inv_s, pool[i].x);
                    pool[i].y = fmaf(v[u][i].y, inv_s, pool[i].y);
                    pool[i].z = fmaf(v[u][i].z, inv_s, pool[i].z);
                    pool[i].w = fmaf(v[u][i].w, inv_s, pool[i].w);
                }
            }
        }
    }
    float4* mine = reinterpret_cast<float4*>(s_pool + warp * P) + lane;
    int* s_arg = reinterpret_cast<int*>(s_pool + wpb * P);
#pragma unroll
    for (int i = 0; i < NV; ++i) mine[32 * i] = pool[i];
    if (ARG) {
        int4* marg = reinterpret_cast<int4*>(s_arg + warp * P) + lane;
#pragma unroll
        for (int i = 0; i < (ARG ? NV : 1); ++i) marg[32 * i] = parg[i];
    }
    __syncthreads();
    float* out = partial + (static_cast<size_t>(b) * chunks + chunk) * P;
    const bool want_arg = ARG && mode == POOL_MAX;
    for (int p = threadIdx.x; p < P; p += blockDim.x) {
        float a = s_pool[p];
        int ai = want_arg ? s_arg[p] : 0;
        for (int wv = 1; wv < wpb; ++wv) {
            const float o = s_pool[wv * P + p];
            if (mode == POOL_MAX) {
                const int oi = want_arg ? s_arg[wv * P + p] : 0;
                if (o > a || (o == a && oi < ai)) { a = o; ai = oi; }
            } else {
                a += o;
            }
        }
        out[p] = a;
        if (want_arg) partial_arg[(static_cast<size_t>(b) * chunks + chunk) * P + p] = ai;
    }
}

// ------------------------------------------------------------------------------------------------ tail
enum Encoding : int { ENC_NONE = 0, ENC_ONEHOT = 1, ENC_LINEAR = 2, ENC_LINEAR_FULL = 3, ENC_BILINEAR = 4, ENC_IDENTITY = 5 };

struct TailParams {
    const float* partial;  // [B, chunks, P]
    int chunks;
    int mode;        // PoolMode
    int inference;   // forward(xs, inference)
    int use_ste;     // CountPIPNet._use_ste
    int max_count;
    int encoding;    // Encoding
    const float* enc_w0;  // linear: [E,1]; linear_full: [D,P]; bilinear: embed [D,P]
    const float* enc_w1;  // bilinear: W [D,D]
    const float* enc_w2;  // bilinear: V [D,D]
    const float* cls_w;   // [K, D]
    const float* cls_b;   // [K] or null
    float* pooled_out;    // [B, P]
    float* logits_out;    // [B, K]
    int P, D, K;
};

// gridDim.x = images, gridDim.y = class slices: every CTA of an image repeats the (cheap) chunk reduction and encoding
// and computes the classes k = blockIdx.y, blockIdx.y + gridDim.y, ... Shared memory: pooled[P] | inter[D] | tmp[D]
__global__ void __launch_bounds__(256)
head_tail_kernel(const TailParams p) {
    extern __shared__ float s_buf[];
    ptx::pdl_wait();
    ptx::pdl_launch_dependents();
    float* s_pooled = s_buf;
    float* s_inter = s_buf + p.P;
    float* s_tmp = s_inter + p.D;
    const int b = blockIdx.x;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31, wpb = blockDim.x >> 5;
    const int M = p.max_count;

    for (int i = tid; i < p.P; i += blockDim.x) {
        const float* pp = p.partial + static_cast<size_t>(b) * p.chunks * p.P + i;
        float a = pp[0];
        for (int c = 1; c < p.chunks; ++c) {
            const float o = pp[static_cast<size_t>(c) * p.P];
            a = (p.mode == POOL_MAX) ? fmaxf(a, o) : a + o;
        }
        float ret;
        if (p.mode == POOL_MAX) {
            if (p.inference) a = (a < 0.1f) ? 0.f : a;
            ret = a;
        } else {
            const float counts = a;
            if (p.use_ste || p.inference) a = rintf(a);  // torch.round: half to even
            a = fminf(fmaxf(a, 0.f), static_cast<float>(M));
            ret = p.inference ? a : counts;
        }
        s_pooled[i] = a;
        if (blockIdx.y == 0) p.pooled_out[static_cast<size_t>(b) * p.P + i] = ret;
    }
    __syncthreads();

    // count encoding -> s_inter[D]
    if (p.mode == POOL_MAX || p.encoding == ENC_IDENTITY || p.encoding == ENC_NONE) {
        for (int i = tid; i < p.D; i += blockDim.x) s_inter[i] = s_pooled[i];
    } else if (p.encoding == ENC_ONEHOT) {
        for (int i = tid; i < p.D; i += blockDim.x) {
            const int proto = i / M, bin = i % M;
            float x = s_pooled[proto];
            if (p.use_ste) x = rintf(x);
            float e = 0.f;
            if (x > 0.1f) {
                long long idx = static_cast<long long>(x) - 1;
                idx = idx < 0 ? 0 : (idx > M - 1 ? M - 1 : idx);
                e = (bin == static_cast<int>(idx)) ? 1.f : 0.f;
            }
            s_inter[i] = e;
        }
    } else if (p.encoding == ENC_LINEAR) {
        const int E = p.D / p.P;
        for (int i = tid; i < p.D; i += blockDim.x) s_inter[i] = s_pooled[i / E] * p.enc_w0[i % E];
    } else if (p.encoding == ENC_LINEAR_FULL || p.encoding == ENC_BILINEAR) {
        float* dst = (p.encoding == ENC_BILINEAR) ? s_tmp : s_inter;
        for (int i = warp; i < p.D; i += wpb) {
            float a = 0.f;
            for (int j = lane; j < p.P; j += 32) a = fmaf(s_pooled[j], p.enc_w0[static_cast<size_t>(i) * p.P + j], a);
            a = warp_sum(a);
            if (lane == 0) dst[i] = a;
        }
        if (p.encoding == ENC_BILINEAR) {
            __syncthreads();
            for (int i = warp; i < p.D; i += wpb) {
                float a = 0.f, c = 0.f;
                for (int j = lane; j < p.D; j += 32) {
                    const float e = s_tmp[j];
                    a = fmaf(e, p.enc_w1[static_cast<size_t>(i) * p.D + j], a);
                    c = fmaf(e, p.enc_w2[static_cast<size_t>(i) * p.D + j], c);
                }
                a = warp_sum(a);
                c = warp_sum(c);
                if (lane == 0) s_inter[i] = a * c;
            }
        }
    }
    __syncthreads();

    // non-negative classifier: logits[k] = sum_d relu(W[k,d]) * inter[d] + bias[k]
    for (int k = blockIdx.y + warp * gridDim.y; k < p.K; k += wpb * gridDim.y) {
        const float* wr = p.cls_w + static_cast<size_t>(k) * p.D;
        float a = 0.f;
        for (int d = lane; d < p.D; d += 32) a = fmaf(s_inter[d], fmaxf(__ldg(wr + d), 0.f), a);
        a = warp_sum(a);
        if (lane == 0) p.logits_out[static_cast<size_t>(b) * p.K + k] = a + (p.cls_b != nullptr ? p.cls_b[k] : 0.f);
    }
}

}  // namespace cpn
