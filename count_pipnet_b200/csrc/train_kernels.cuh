// Training-side kernels of the prototype head: the PIP-Net losses with their gradients in
// one pass over the prototype map, and softmax + spatial pooling forward / backward on the NCHW map the composed
// (autograd) backbone hands over. HBM-bound fp32 work: coalesced along the spatial axis, fixed-order reductions (two
// runs give the same bits), no atomics.
//
//   reference                                            kernel
//   pipnet/train.py:253-265 align_loss (both directions)  pipnet_loss_kernel, map CTAs
//   pipnet/train.py:198-201 tanh loss,
//     :216-222 log1p(out^m) + log_softmax + NLLLoss(mean),
//     :233-235 accuracy, :212-229 weighting of the terms   pipnet_loss_kernel, terms CTAs + the last CTA
//   pipnet/pipnet.py:96,105 Softmax(dim=1) + max pool,
//     pipnet/count_pipnet.py:88 spatial sum                softmax_pool_nchw_kernel, pool_fold_nchw_kernel
//   their autograd backward                                softmax_pool_nchw_bwd_kernel
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#include "device_utils.cuh"

namespace cpn {

constexpr int TRAIN_TILE = 32;      // spatial positions per CTA: one 128-byte row segment per channel
constexpr int TRAIN_THREADS = 512;  // at most; two CTAs per SM, so that one CTA's reduction phases overlap the other's loads
                                    // and the second pass of a CTA (32 positions x P channels) is still in L2

// VEC = 4 when H*W % 4 == 0 and the maps are 16-byte aligned: a lane owns 4 neighbouring positions (one LDG.128 per
// channel), 8 lanes cover the tile and a warp reads four channel rows per instruction. VEC = 1 otherwise (lane = position).
template <int VEC>
__device__ __forceinline__ void ldv(const float* __restrict__ p, float (&o)[VEC]) {
    if constexpr (VEC == 4) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        o[0] = t.x, o[1] = t.y, o[2] = t.z, o[3] = t.w;
    } else {
        o[0] = *p;
    }
}
template <int VEC>
__device__ __forceinline__ void stv_cs(float* __restrict__ p, const float (&o)[VEC]) {  // streaming store
    if constexpr (VEC == 4) {
        __stcs(reinterpret_cast<float4*>(p), make_float4(o[0], o[1], o[2], o[3]));
    } else {
        __stcs(p, o[0]);
    }
}
template <int VEC>
__device__ __forceinline__ void stv(float* __restrict__ p, const float (&o)[VEC]) {
    if constexpr (VEC == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
        *p = o[0];
    }
}

// Reduction over the channel groups of a CTA for each of the tile's 32 positions, in a fixed order: lanes of a warp that
// share positions (butterfly), then the warps in index order. Every thread passes its VEC partials for positions
// hwl * VEC + k; afterwards s_col[0..32) holds the totals (read them before the next call).
template <int VEC, bool MAX>
__device__ __forceinline__ void tile_reduce(float (&acc)[VEC], float (*s_red)[TRAIN_TILE + 1], float* s_col) {
    constexpr int L = 32 / VEC;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
    for (int o = L; o < 32; o <<= 1) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const float other = __shfl_xor_sync(0xffffffffu, acc[k], o);
            acc[k] = MAX ? fmaxf(acc[k], other) : acc[k] + other;
        }
    }
    if (lane < L) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) s_red[warp][lane * VEC + k] = acc[k];
    }
    __syncthreads();
    if (tid < TRAIN_TILE) {
        float r = s_red[0][tid];
        for (int w = 1; w < nwarps; ++w) r = MAX ? fmaxf(r, s_red[w][tid]) : r + s_red[w][tid];
        s_col[tid] = r;
    }
    __syncthreads();
}

// ---- tanh loss, classification loss, weighting: extra CTAs of the map kernel -------------------------------------------
// The first n_terms = 2 * ceil(P / 128) (tanh) + (pretrain ? 0 : 2B) (classification) CTAs of pipnet_loss_kernel do these
// terms on their first 128 threads while the other CTAs stream the map; every CTA takes a ticket after publishing its
// partials and the one that takes the last sums every partial array in a fixed order and writes result[8] (so the values
// do not depend on which CTA that is). The ticket is zeroed by a 4-byte memset node ahead of the kernel.
__device__ __forceinline__ void bar128() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
struct ValIdx {
    float v;
    int i;
};
__device__ __forceinline__ ValIdx better(ValIdx a, ValIdx b) {  // larger value, then lower index
    return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ ValIdx warp_argmax(ValIdx a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ValIdx b;
        b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
        b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
        a = better(a, b);
    }
    return a;
}

struct LossTermsArgs {
    const float* pooled;      // [2B, P]
    const float* out;         // [2B, K]
    const long long* ys;      // [B]
    const float* mult;        // [1]
    float* grad_pooled;       // [2B, P] or null
    float* grad_out;          // [2B, K] or null
    const float* align_partial;  // [n_align], written by the map CTAs
    float* tanh_partial;      // [2 * tanh_blocks]
    float* row_loss;          // [2B]
    float* row_gmult;         // [2B]
    float* row_correct;       // [2B]
    unsigned* ticket;
    float* result;            // [8]
    int B, P, K, n_align, tanh_blocks;
    int use_log1p, pretrain, finetune;
    float coeff, eps, tanh_scale, class_scale;  // tanh_scale = -tanh_w / (2P), class_scale = class_w / 2B
    float n_patches, align_w, tanh_w, class_w;
};

// pooled [2B, P]; for each view: t_p = tanh(sum_b coeff * pooled[b, p]), loss term = log(t_p + eps) (train.py:198-201);
// d loss / d pooled[b, p] = tanh_scale * coeff * (1 - t^2) / (t + eps). One CTA = 128 prototypes of one view.
__device__ __forceinline__ void tanh_terms(const LossTermsArgs& a, int view, int pblock, float* s_f) {
    const int p = pblock * 128 + threadIdx.x;
    const float* src = a.pooled + static_cast<size_t>(view) * a.B * a.P + p;
    float l = 0.f;
    if (p < a.P) {
        float s = 0.f;
        int b = 0;
        for (; b + 8 <= a.B; b += 8) {  // eight loads in flight, summed in image order
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = src[static_cast<size_t>(b + j) * a.P];
#pragma unroll
            for (int j = 0; j < 8; ++j) s = __fadd_rn(s, __fmul_rn(a.coeff, v[j]));
        }
        for (; b < a.B; ++b) s = __fadd_rn(s, __fmul_rn(a.coeff, src[static_cast<size_t>(b) * a.P]));
        const float t = tanhf(s);
        l = logf(t + a.eps);
        if (a.grad_pooled != nullptr) {
            const float g = a.tanh_scale * a.coeff * (1.f - t * t) / (t + a.eps);
            float* dst = a.grad_pooled + static_cast<size_t>(view) * a.B * a.P + p;
            for (b = 0; b < a.B; ++b) dst[static_cast<size_t>(b) * a.P] = g;
        }
    }
    l = warp_sum(l);
    if ((threadIdx.x & 31) == 0) s_f[threadIdx.x >> 5] = l;
    bar128();
    if (threadIdx.x == 0) a.tanh_partial[view * a.tanh_blocks + pblock] = (s_f[0] + s_f[1]) + (s_f[2] + s_f[3]);
}

// out [2B, K], ys [B] (labels of the pairs: train.py:169 ys = cat(ys1, ys1)), mult = the classifier's
// normalization_multiplier. v = use_log1p ? log1p(out^mult) : out; loss_i = -log_softmax(v)[y] (train.py:216-222).
// One CTA per row: grad_out (scaled by class_scale) and per row the loss, the multiplier's gradient, whether
// arg-max(out) == y (train.py:233-235; first index on ties, as torch.argmax on CUDA).
__device__ __forceinline__ void class_terms(const LossTermsArgs& a, int i, float* s_f, ValIdx* s_vi) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, K = a.K;
    const float* o = a.out + static_cast<size_t>(i) * K;
    const int y = static_cast<int>(a.ys[i % a.B]);
    const bool use_log1p = a.use_log1p != 0;
    const float m = use_log1p ? a.mult[0] : 1.f;
    auto value = [&](float x) { return use_log1p ? log1pf(powf(x, m)) : x; };

    float vmax = -INFINITY;
    ValIdx best{-INFINITY, 0x7fffffff};
    for (int k = tid; k < K; k += 128) {
        const float x = o[k];
        vmax = fmaxf(vmax, value(x));
        best = better(best, ValIdx{x, k});
    }
    vmax = warp_max(vmax);
    best = warp_argmax(best);
    if (lane == 0) {
        s_f[warp] = vmax;
        s_vi[warp] = best;
    }
    bar128();
    vmax = fmaxf(fmaxf(s_f[0], s_f[1]), fmaxf(s_f[2], s_f[3]));
    best = better(better(s_vi[0], s_vi[1]), better(s_vi[2], s_vi[3]));
    bar128();

    float se = 0.f;
    for (int k = tid; k < K; k += 128) se += expf(value(o[k]) - vmax);
    se = warp_sum(se);
    if (lane == 0) s_f[warp] = se;
    bar128();
    se = (s_f[0] + s_f[1]) + (s_f[2] + s_f[3]);
    const float lse = vmax + logf(se);
    bar128();

    float gm = 0.f;
    if (a.grad_out != nullptr) {
        float* g = a.grad_out + static_cast<size_t>(i) * K;
        for (int k = tid; k < K; k += 128) {
            const float x = o[k];
            const float gv = (expf(value(x) - lse) - (k == y ? 1.f : 0.f)) * a.class_scale;
            if (use_log1p) {
                // d log1p(x^m) / dx = m x^(m-1) / (1 + x^m); d / dm = x^m log(x) / (1 + x^m), 0 at x == 0 (torch's pow
                // backward: zero where the base is 0 and the exponent >= 0)
                const float pw = powf(x, m);
                const float q = gv / (1.f + pw);
                g[k] = m == 0.f ? 0.f : q * (m * powf(x, m - 1.f));
                gm += (x == 0.f && m >= 0.f) ? 0.f : q * (pw * logf(x));
            } else {
                g[k] = gv;
            }
        }
    }
    gm = warp_sum(gm);
    if (lane == 0) s_f[warp] = gm;
    bar128();
    if (tid == 0) {
        a.row_loss[i] = (y >= 0 && y < K) ? lse - value(o[y]) : 0.f;
        a.row_gmult[i] = (s_f[0] + s_f[1]) + (s_f[2] + s_f[3]);
        a.row_correct[i] = best.i == y ? 1.f : 0.f;
    }
}

// Fixed-order sum of v[0..n) by a CTA of 128 threads; v was written by other CTAs of this or an earlier launch (read
// through L2).
__device__ __forceinline__ float block_sum_128(const float* v, int n, float* s_f) {
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += 128) acc += __ldcg(v + i);
    acc = warp_sum(acc);
    bar128();  // s_f may still be read from the previous call
    if ((threadIdx.x & 31) == 0) s_f[threadIdx.x >> 5] = acc;
    bar128();
    return (s_f[0] + s_f[1]) + (s_f[2] + s_f[3]);
}

// train.py:212-229's weighting. result[0] align, [1] tanh, [2] class, [3] loss, [4] number of correct predictions,
// [5] d loss / d multiplier, [6..7] 0.
__device__ __forceinline__ void finalize_terms(const LossTermsArgs& a, float* s_f) {
    const float al = block_sum_128(a.align_partial, a.n_align, s_f) / a.n_patches;
    const float t1 = block_sum_128(a.tanh_partial, a.tanh_blocks, s_f) / static_cast<float>(a.P);
    const float t2 = block_sum_128(a.tanh_partial + a.tanh_blocks, a.tanh_blocks, s_f) / static_cast<float>(a.P);
    float cl = 0.f, gm = 0.f, correct = 0.f;
    if (!a.pretrain) {
        cl = block_sum_128(a.row_loss, 2 * a.B, s_f) / static_cast<float>(2 * a.B);
        gm = block_sum_128(a.row_gmult, 2 * a.B, s_f);
        correct = block_sum_128(a.row_correct, 2 * a.B, s_f);
    }
    if (threadIdx.x == 0) {
        const float th = -(t1 + t2) / 2.f;
        float loss = 0.f;
        if (!a.finetune) loss = a.align_w * al + a.tanh_w * th;
        if (!a.pretrain) loss = a.finetune ? a.class_w * cl : loss + a.class_w * cl;
        a.result[0] = al;
        a.result[1] = th;
        a.result[2] = cl;
        a.result[3] = loss;
        a.result[4] = correct;
        a.result[5] = gm;
        a.result[6] = 0.f;
        a.result[7] = 0.f;
    }
}

// The CTAs that do these terms, and the CTA that takes the last ticket, run them on their first 128 threads.
__device__ __forceinline__ void loss_terms_cta(const LossTermsArgs& a, int idx, float* s_f, ValIdx* s_vi) {
    const int n_tanh = 2 * a.tanh_blocks;
    if (idx < n_tanh)
        tanh_terms(a, idx / a.tanh_blocks, idx % a.tanh_blocks, s_f);
    else
        class_terms(a, idx - n_tanh, s_f, s_vi);
}

// ---- align loss -----------------------------------------------------------------------------------------------------
// pf [2B, P, HW]: view 1 = images [0, B), view 2 = images [B, 2B). For patch n = (b, hw): d_n = sum_c pf1 * pf2,
// loss = -mean_n log(d_n + eps) (the two directions of train.py:173 have the same value); d loss / d pf1 = grad_scale *
// pf2 / (d_n + eps) and the mirror image for pf2, grad_scale = -align_weight / (2 N).
// grid = n_terms + ceil(HW / 32) * B CTAs (the small-terms CTAs first, so that they start with the first wave), block =
// (32 / VEC) * channel groups, at least 128. partial [B * tiles] = per-CTA sum of -log(d + eps).
template <int VEC>
__global__ void __launch_bounds__(TRAIN_THREADS, 2)
    pipnet_loss_kernel(const float* __restrict__ pf, float* __restrict__ grad, float* __restrict__ partial, int n_terms,
                       int tiles, int B, int P, int HW, float eps, float grad_scale, const LossTermsArgs terms) {
    __shared__ float s_f[4];
    __shared__ ValIdx s_vi[4];
    __shared__ int s_last;
    if (static_cast<int>(blockIdx.x) < n_terms) {
        if (threadIdx.x >= 128) return;
        loss_terms_cta(terms, blockIdx.x, s_f, s_vi);
        if (threadIdx.x == 0) {  // both roles end with thread 0 writing this CTA's partials: publish them, take a ticket
            __threadfence();
            s_last = atomicAdd(terms.ticket, 1u) == gridDim.x - 1;
        }
        bar128();
        if (s_last) {
            __threadfence();
            finalize_terms(terms, s_f);
        }
        return;
    }
    constexpr int L = 32 / VEC;
    constexpr int U = VEC == 4 ? 2 : 4;  // channels in flight per thread and view
    __shared__ float s_red[TRAIN_THREADS / 32][TRAIN_TILE + 1];
    __shared__ float s_col[TRAIN_TILE];
    const int tid = threadIdx.x, hwl = tid % L, cg = tid / L, CG = blockDim.x / L;
    const int map_cta = blockIdx.x - n_terms, tile = map_cta % tiles, b = map_cta / tiles;
    const int hw0 = tile * TRAIN_TILE + hwl * VEC;
    const bool ok = hw0 < HW;
    const size_t img = static_cast<size_t>(P) * HW;
    const size_t off1 = static_cast<size_t>(b) * img + hw0, off2 = off1 + static_cast<size_t>(B) * img;
    const float* p1 = pf + off1;
    const float* p2 = pf + off2;
    float acc[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) acc[k] = 0.f;
    if (ok) {
        int c = cg;
        for (; c + (U - 1) * CG < P; c += U * CG) {
            float a[U][VEC], v[U][VEC];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                ldv<VEC>(p1 + static_cast<size_t>(c + u * CG) * HW, a[u]);
                ldv<VEC>(p2 + static_cast<size_t>(c + u * CG) * HW, v[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int k = 0; k < VEC; ++k) acc[k] = fmaf(a[u][k], v[u][k], acc[k]);
        }
        for (; c < P; c += CG) {
            float a[VEC], v[VEC];
            ldv<VEC>(p1 + static_cast<size_t>(c) * HW, a);
            ldv<VEC>(p2 + static_cast<size_t>(c) * HW, v);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = fmaf(a[k], v[k], acc[k]);
        }
    }
    tile_reduce<VEC, false>(acc, s_red, s_col);
    if (tid < TRAIN_TILE) {  // warp 0: lane = position of the tile
        const bool valid = tile * TRAIN_TILE + tid < HW;
        const float d = s_col[tid] + eps;
        s_col[tid] = valid ? grad_scale / d : 0.f;
        const float l = warp_sum(valid ? -logf(d) : 0.f);
        if (tid == 0) {
            partial[static_cast<size_t>(b) * tiles + tile] = l;
            __threadfence();
            s_last = atomicAdd(terms.ticket, 1u) == gridDim.x - 1;
        }
    }
    __syncthreads();
    if (s_last && tid < 128) {  // no barrier of the whole CTA follows: the other threads go on to the gradient pass
        __threadfence();
        finalize_terms(terms, s_f);
    }
    if (grad == nullptr || !ok) return;
    float kf[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) kf[k] = s_col[hwl * VEC + k];
    float* g1 = grad + off1;
    float* g2 = grad + off2;
    // second pass from the LAST channel row down: what the first pass read most recently is what L2 still holds
    if (cg >= P) return;
    int c = cg + (P - 1 - cg) / CG * CG;  // this thread's last row
    for (; c - (U - 1) * CG >= cg; c -= U * CG) {
        float a[U][VEC], v[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            ldv<VEC>(p1 + static_cast<size_t>(c - u * CG) * HW, a[u]);
            ldv<VEC>(p2 + static_cast<size_t>(c - u * CG) * HW, v[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                const float t = a[u][k];
                a[u][k] = kf[k] * v[u][k];
                v[u][k] = kf[k] * t;
            }
            stv_cs<VEC>(g1 + static_cast<size_t>(c - u * CG) * HW, a[u]);
            stv_cs<VEC>(g2 + static_cast<size_t>(c - u * CG) * HW, v[u]);
        }
    }
    for (; c >= cg; c -= CG) {
        float a[VEC], v[VEC];
        ldv<VEC>(p1 + static_cast<size_t>(c) * HW, a);
        ldv<VEC>(p2 + static_cast<size_t>(c) * HW, v);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const float t = a[k];
            a[k] = kf[k] * v[k];
            v[k] = kf[k] * t;
        }
        stv_cs<VEC>(g1 + static_cast<size_t>(c) * HW, a);
        stv_cs<VEC>(g2 + static_cast<size_t>(c) * HW, v);
    }
}

// buf *= *scale unless *scale == 1 (the usual loss.backward() case costs one tiny launch and no pass over memory).
__global__ void scale_unless_one_kernel(float* __restrict__ buf, long long n, const float* __restrict__ scale) {
    const float s = scale[0];
    if (s == 1.f) return;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) buf[i] *= s;
}

// ---- softmax over prototypes + spatial pooling on an NCHW map, forward ------------------------------------------------
// 2^x as one MUFU.EX2: exp(x - m) = 2^((x - m) * log2e); the subtraction first keeps the error independent of the size of
// the logits (relative error < 1e-6 for the entries that matter, x - m > -10).
constexpr float LOG2E = 1.4426950408889634f;
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// z [N, P, HW] -> pf [N, P, HW] = softmax over P per position; per-CTA pooling partials over its 32 positions:
// part_val [N, tiles, P] (max or sum), part_idx [N, tiles, P] (position of the maximum, lowest on ties; max mode only).
// grid = (tiles = ceil(HW / 32), N), block = (32 / VEC) * channel groups.
constexpr int STAGE_ROWS = 512;  // 64 KB of staging rows per CTA when the whole tile does not fit
// CACHE: the CTA's [P, 32] tile of z is kept in dynamic shared memory (P * 128 bytes, two CTAs per SM up to P = 768), so
// z is read from global memory once instead of three times; every thread reads back only what it wrote itself (a warp
// writes 512 contiguous bytes: no bank conflicts, no barrier). Without it passes 2 and 3 re-read the tile from L2 and the
// probabilities are staged STAGE_ROWS channel rows at a time for the pooling phase.
template <int VEC, bool CACHE>
__global__ void __launch_bounds__(TRAIN_THREADS, 2)
    softmax_pool_nchw_kernel(const float* __restrict__ z, float* __restrict__ pf, float* __restrict__ part_val,
                             int* __restrict__ part_idx, int mode_sum, int P, int HW) {
    constexpr int L = 32 / VEC;
    constexpr int U = VEC == 4 ? 4 : 8;  // loads in flight per thread in the first pass
    extern __shared__ __align__(16) float s_tile[];  // [P][32] when CACHE, else [min(P, STAGE_ROWS)][32] staging rows
    __shared__ float s_red[TRAIN_THREADS / 32][TRAIN_TILE + 1];
    __shared__ float s_col[TRAIN_TILE];
    const int tid = threadIdx.x, hwl = tid % L, cg = tid / L, CG = blockDim.x / L;
    const int n = blockIdx.y, hw0 = blockIdx.x * TRAIN_TILE + hwl * VEC;
    const bool ok = hw0 < HW;
    const size_t base = static_cast<size_t>(n) * P * HW + hw0;
    const float* zp = z + base;
    // this thread's slot in row c of the tile: 16-byte groups XOR-swizzled with the row number, so that the pooling phase
    // (thread = channel row, all 32 positions) reads rows 128 bytes apart without bank conflicts
    auto slot = [&](int c) { return s_tile + c * TRAIN_TILE + ((((hwl * VEC) >> 2) ^ (c & 7)) << 2) + ((hwl * VEC) & 3); };
    auto reload = [&](int c, float (&x)[VEC]) {
        if constexpr (CACHE) {
            ldv<VEC>(slot(c), x);
        } else {
            ldv<VEC>(zp + static_cast<size_t>(c) * HW, x);
        }
    };

    float m[VEC], s[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) m[k] = -INFINITY;
    if constexpr (CACHE) {
        // the whole tile in flight at once: every thread copies its own slots global -> shared (LDGSTS, no registers held),
        // waits for its own copies only, then takes the maximum from shared memory
        if (ok) {
            for (int c = cg; c < P; c += CG) {
                if constexpr (VEC == 4)
                    ptx::cp_async16(ptx::smem_u32(slot(c)), zp + static_cast<size_t>(c) * HW);
                else
                    ptx::cp_async4(ptx::smem_u32(slot(c)), zp + static_cast<size_t>(c) * HW);
            }
        }
        ptx::cp_async_commit();
        ptx::cp_async_wait<0>();
        if (ok) {
            for (int c = cg; c < P; c += CG) {
                float x[VEC];
                ldv<VEC>(slot(c), x);
#pragma unroll
                for (int k = 0; k < VEC; ++k) m[k] = fmaxf(m[k], x[k]);
            }
        }
    }
    if constexpr (CACHE) {
        tile_reduce<VEC, true>(m, s_red, s_col);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            m[k] = s_col[hwl * VEC + k];
            s[k] = 0.f;
        }
        if (ok) {
            for (int c = cg; c < P; c += CG) {
                float x[VEC];
                ldv<VEC>(slot(c), x);
#pragma unroll
                for (int k = 0; k < VEC; ++k) s[k] += ex2_approx((x[k] - m[k]) * LOG2E);
            }
        }
    } else {
        // one pass over global memory for both the maximum and the sum: running maximum per thread, the sum rescaled when
        // it moves (U loads in flight, U + 1 exponentials per U values); then the threads' (m, s) pairs are merged
#pragma unroll
        for (int k = 0; k < VEC; ++k) s[k] = 0.f;
        if (ok) {
            int c = cg;
            for (; c + (U - 1) * CG < P; c += U * CG) {
                float x[U][VEC];
#pragma unroll
                for (int u = 0; u < U; ++u) ldv<VEC>(zp + static_cast<size_t>(c + u * CG) * HW, x[u]);
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    float mn = m[k];
#pragma unroll
                    for (int u = 0; u < U; ++u) mn = fmaxf(mn, x[u][k]);
                    float acc = s[k] * ex2_approx((m[k] - mn) * LOG2E);
#pragma unroll
                    for (int u = 0; u < U; ++u) acc += ex2_approx((x[u][k] - mn) * LOG2E);
                    s[k] = acc;
                    m[k] = mn;
                }
            }
            for (; c < P; c += CG) {
                float x[VEC];
                ldv<VEC>(zp + static_cast<size_t>(c) * HW, x);
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const float mn = fmaxf(m[k], x[k]);
                    s[k] = s[k] * ex2_approx((m[k] - mn) * LOG2E) + ex2_approx((x[k] - mn) * LOG2E);
                    m[k] = mn;
                }
            }
        }
        float mt[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) mt[k] = m[k];
        tile_reduce<VEC, true>(m, s_red, s_col);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            m[k] = s_col[hwl * VEC + k];
            // a thread without channels (or beyond the map's edge) has mt = -inf and s = 0: keep it 0, not 0 * 2^(-inf - -inf)
            s[k] = s[k] == 0.f ? 0.f : s[k] * ex2_approx((mt[k] - m[k]) * LOG2E);
        }
    }
    tile_reduce<VEC, false>(s, s_red, s_col);
    float inv[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) inv[k] = 1.0f / s_col[hwl * VEC + k];

    float* op = pf + base;
    float* pv = part_val + (static_cast<size_t>(n) * gridDim.x + blockIdx.x) * P;
    int* pi = part_idx == nullptr ? nullptr : part_idx + (static_cast<size_t>(n) * gridDim.x + blockIdx.x) * P;
    // pass 3, in chunks of `rows` channel rows (all P rows when the tile is cached): probabilities out to global memory and
    // into the thread's own slots of the staging rows (zeros beyond the map's edge); then thread = channel row: max (first
    // position on ties) or sum over the tile's 32 positions, in position order.
    const int rows = CACHE ? P : (P < STAGE_ROWS ? P : STAGE_ROWS);
    for (int c0 = 0; c0 < P; c0 += rows) {
        const int c1 = c0 + rows < P ? c0 + rows : P;
        int c = c0 + cg;
        if constexpr (!CACHE) {  // global loads U at a time
            for (; c + (U - 1) * CG < c1; c += U * CG) {
                float x[U][VEC];
#pragma unroll
                for (int u = 0; u < U; ++u) {
#pragma unroll
                    for (int k = 0; k < VEC; ++k) x[u][k] = 0.f;
                    if (ok) ldv<VEC>(zp + static_cast<size_t>(c + u * CG) * HW, x[u]);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (ok) {
#pragma unroll
                        for (int k = 0; k < VEC; ++k) x[u][k] = ex2_approx((x[u][k] - m[k]) * LOG2E) * inv[k];
                        stv_cs<VEC>(op + static_cast<size_t>(c + u * CG) * HW, x[u]);
                    }
                    stv<VEC>(slot(c + u * CG - c0), x[u]);
                }
            }
        }
        for (; c < c1; c += CG) {
            float y[VEC];
#pragma unroll
            for (int k = 0; k < VEC; ++k) y[k] = 0.f;
            if (ok) {
                float x[VEC];
                reload(c, x);
#pragma unroll
                for (int k = 0; k < VEC; ++k) y[k] = ex2_approx((x[k] - m[k]) * LOG2E) * inv[k];
                stv_cs<VEC>(op + static_cast<size_t>(c) * HW, y);
            }
            stv<VEC>(slot(c - c0), y);
        }
        __syncthreads();
        for (int r = tid; r < c1 - c0; r += blockDim.x) {
            const float* row = s_tile + r * TRAIN_TILE;
            float v[TRAIN_TILE];
#pragma unroll
            for (int j = 0; j < TRAIN_TILE / 4; ++j) {
                const float4 t = *reinterpret_cast<const float4*>(row + ((j ^ (r & 7)) << 2));
                v[4 * j] = t.x, v[4 * j + 1] = t.y, v[4 * j + 2] = t.z, v[4 * j + 3] = t.w;
            }
            if (mode_sum) {
                float t = v[0];
#pragma unroll
                for (int j = 1; j < TRAIN_TILE; ++j) t += v[j];
                pv[c0 + r] = t;
            } else {
                float best = v[0];
#pragma unroll
                for (int j = 1; j < TRAIN_TILE; ++j) best = fmaxf(best, v[j]);
                int idx = 0;
#pragma unroll
                for (int j = TRAIN_TILE - 1; j >= 0; --j) idx = (v[j] == best) ? j : idx;
                pv[c0 + r] = best;
                pi[c0 + r] = blockIdx.x * TRAIN_TILE + idx;
            }
        }
        if (c1 < P) __syncthreads();  // the next chunk overwrites the staging rows
    }
}

// pooled [N, P] (and argmax [N, P] in max mode) from the per-tile partials, in tile order.
// grid = (ceil(P / 128), N), block = 128.
__global__ void __launch_bounds__(128)
    pool_fold_nchw_kernel(const float* __restrict__ part_val, const int* __restrict__ part_idx, int mode_sum,
                          float* __restrict__ pooled, int* __restrict__ argmax, int P, int tiles) {
    const int n = blockIdx.y, p = blockIdx.x * 128 + threadIdx.x;
    if (p >= P) return;
    const size_t base = static_cast<size_t>(n) * tiles * P + p;
    if (mode_sum) {
        float a = 0.f;
        for (int t = 0; t < tiles; ++t) a += part_val[base + static_cast<size_t>(t) * P];
        pooled[static_cast<size_t>(n) * P + p] = a;
    } else {
        ValIdx best{part_val[base], part_idx[base]};
        for (int t = 1; t < tiles; ++t)
            best = better(best, ValIdx{part_val[base + static_cast<size_t>(t) * P], part_idx[base + static_cast<size_t>(t) * P]});
        pooled[static_cast<size_t>(n) * P + p] = best.v;
        argmax[static_cast<size_t>(n) * P + p] = best.i;
    }
}

// ---- backward ---------------------------------------------------------------------------------------------------------
// g[c] = grad_pf[n, c, hw] (0 if grad_pf == nullptr) + grad_pooled[n, c] * (sum mode ? 1 : hw == argmax[n, c]);
// grad_z = pf * (g - sum_c g * pf)   (softmax backward of the sum of both incoming gradients).
// grid = (ceil(HW / 32), N), block = (32 / VEC) * channel groups.
template <int VEC>
__global__ void __launch_bounds__(TRAIN_THREADS, 2)
    softmax_pool_nchw_bwd_kernel(const float* __restrict__ pf, const float* __restrict__ grad_pf,
                                 const float* __restrict__ grad_pooled, const int* __restrict__ argmax, int mode_sum,
                                 float* __restrict__ grad_z, int P, int HW) {
    constexpr int L = 32 / VEC;
    __shared__ float s_red[TRAIN_THREADS / 32][TRAIN_TILE + 1];
    __shared__ float s_col[TRAIN_TILE];
    const int tid = threadIdx.x, hwl = tid % L, cg = tid / L, CG = blockDim.x / L;
    const int n = blockIdx.y, hw0 = blockIdx.x * TRAIN_TILE + hwl * VEC;
    const bool ok = hw0 < HW;
    const size_t base = static_cast<size_t>(n) * P * HW + hw0;
    const float* yp = pf + base;
    const float* gp = grad_pf == nullptr ? nullptr : grad_pf + base;
    const float* gpool = grad_pooled == nullptr ? nullptr : grad_pooled + static_cast<size_t>(n) * P;
    const int* am = (argmax == nullptr || mode_sum) ? nullptr : argmax + static_cast<size_t>(n) * P;
    auto incoming = [&](int c, float (&g)[VEC]) {
        if (gp != nullptr) {
            ldv<VEC>(gp + static_cast<size_t>(c) * HW, g);
        } else {
#pragma unroll
            for (int k = 0; k < VEC; ++k) g[k] = 0.f;
        }
        if (gpool != nullptr) {
            const float t = __ldg(gpool + c);
            if (mode_sum) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) g[k] += t;
            } else {
                const int a = __ldg(am + c) - hw0;
#pragma unroll
                for (int k = 0; k < VEC; ++k) g[k] += (a == k) ? t : 0.f;
            }
        }
    };
    float acc[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) acc[k] = 0.f;
    if (ok) {
        int c = cg;
        for (; c + CG < P; c += 2 * CG) {
            float y0[VEC], y1[VEC], g0[VEC], g1[VEC];
            ldv<VEC>(yp + static_cast<size_t>(c) * HW, y0);
            ldv<VEC>(yp + static_cast<size_t>(c + CG) * HW, y1);
            incoming(c, g0);
            incoming(c + CG, g1);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = fmaf(g0[k], y0[k], acc[k]);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = fmaf(g1[k], y1[k], acc[k]);
        }
        if (c < P) {
            float y0[VEC], g0[VEC];
            ldv<VEC>(yp + static_cast<size_t>(c) * HW, y0);
            incoming(c, g0);
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = fmaf(g0[k], y0[k], acc[k]);
        }
    }
    tile_reduce<VEC, false>(acc, s_red, s_col);
    if (!ok) return;
    float dot[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) dot[k] = s_col[hwl * VEC + k];
    float* out = grad_z + base;
    int c = cg;
    for (; c + CG < P; c += 2 * CG) {
        float y0[VEC], y1[VEC], g0[VEC], g1[VEC];
        ldv<VEC>(yp + static_cast<size_t>(c) * HW, y0);
        ldv<VEC>(yp + static_cast<size_t>(c + CG) * HW, y1);
        incoming(c, g0);
        incoming(c + CG, g1);
#pragma unroll
        for (int k = 0; k < VEC; ++k) g0[k] = y0[k] * (g0[k] - dot[k]);
#pragma unroll
        for (int k = 0; k < VEC; ++k) g1[k] = y1[k] * (g1[k] - dot[k]);
        stv_cs<VEC>(out + static_cast<size_t>(c) * HW, g0);
        stv_cs<VEC>(out + static_cast<size_t>(c + CG) * HW, g1);
    }
    if (c < P) {
        float y0[VEC], g0[VEC];
        ldv<VEC>(yp + static_cast<size_t>(c) * HW, y0);
        incoming(c, g0);
#pragma unroll
        for (int k = 0; k < VEC; ++k) g0[k] = y0[k] * (g0[k] - dot[k]);
        stv_cs<VEC>(out + static_cast<size_t>(c) * HW, g0);
    }
}

}  // namespace cpn
