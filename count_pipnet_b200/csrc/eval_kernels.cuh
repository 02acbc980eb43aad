// Per-batch evaluation statistics of the reference's test loop (pipnet/test.py:67-142, compute_local_explanation_sizes
// :266-318, util/func.py:22-37), computed from the forward's outputs without building the [K, B, P] `scores` tensor or
// walking the batch on the host for the confusion matrix.
//
//   pooled [B, P] fp32, out [B, K] fp32 (the module's 2nd and 3rd return values), w [K, P] fp32 = the weights the
//   reference multiplies pooled with (`_classification.weight` as stored -- signed, not clamped -- for PIP-Net;
//   get_prototype_importance_per_class stacked to [K, P] for Count-PIP-Net), ys [B] int64 labels.
//
// One CTA per image. All counts are integer results of fp32 compares on individually rounded products
// (__fmul_rn / __fsub_rn: the compiler must not contract `pooled * w - 1e-3` into an FMA, the reference rounds the
// product first), so they equal the reference's bit for bit; only pred_prob goes through exp/log (1e-5 relative).
// HBM/L2-bound integer work: 4 B per (image, class, prototype) read out of the L2-resident weight table.
#pragma once
#include "convnext_kernels.cuh"

namespace cpn {

struct EvalParams {
    const float* pooled;
    const float* out;
    const float* w;
    const long long* ys;
    float mult;  // _classification.normalization_multiplier (test.py:83)
    float thr;   // 1e-3 (test.py:104, 110, 122)
    int B, P, K;
    int* pred;             // [B] arg-max class, first maximum (torch.max, test.py:82)
    float* pred_prob;      // [B] amax softmax(log1p(out ** mult)) (test.py:83)
    float* max_score;      // [B] out[b, pred]; 0 = abstained (test.py:85)
    int* any_size;         // [B] prototypes with |pooled * w[k]| > thr for ANY class (test.py:295-297)
    int* pred_size;        // [B] ... for the predicted class (test.py:307-316)
    int* anz;              // [B] prototypes with |pooled| > thr (test.py:121-124)
    int* top1;             // [B] 1 if the label is the top class, [B] top5 likewise (util/func.py:22-37)
    int* top5;
    unsigned long long* cm;   // [K, K] confusion matrix, cm[y_true][y_pred] += 1 (test.py:135-137); accumulates
    unsigned char* relevant;  // [K, P] 1 where relu(pooled * w - thr) > 0 for some image of the batch (test.py:108-113)
};

constexpr int EVAL_THREADS = 1024;
constexpr int EVAL_QLANES = 256;                          // prototype lanes ...
constexpr int EVAL_KSLICES = EVAL_THREADS / EVAL_QLANES;  // ... times class slices

__device__ __forceinline__ int block_sum_int(int v, int* s_slot) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd(s_slot, v);
    __syncthreads();
    return *s_slot;
}

__global__ void __launch_bounds__(EVAL_THREADS)
eval_batch_kernel(const EvalParams p) {
    __shared__ float s_val[EVAL_THREADS / 32];
    __shared__ int s_idx[EVAL_THREADS / 32];
    __shared__ float s_f[2];
    __shared__ int s_pred;
    __shared__ int s_cnt[5];
    extern __shared__ unsigned char s_any[];  // [P]: prototype relevant for any class of this image
    const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float* ob = p.out + static_cast<size_t>(b) * p.K;
    if (tid < 5) s_cnt[tid] = 0;

    // ---- arg-max class (first maximum) ----
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int k = tid; k < p.K; k += EVAL_THREADS) {
        const float v = ob[k];
        if (v > bv || bi == 0x7fffffff) { bv = v; bi = k; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
    }
    if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
        float v = s_val[0];
        int i = s_idx[0];
        for (int w = 1; w < EVAL_THREADS / 32; ++w)
            if (s_idx[w] != 0x7fffffff && (i == 0x7fffffff || s_val[w] > v || (s_val[w] == v && s_idx[w] < i))) {
                v = s_val[w];
                i = s_idx[w];
            }
        s_pred = i;
        p.pred[b] = i;
        p.max_score[b] = v;
        // log1p and x ** mult are increasing for x >= 0 (logits of the non-negative classifier), but a negative bias can
        // produce negative logits whose even power is large: take the maximum of the transformed values explicitly.
    }
    __syncthreads();
    const int pred = s_pred;

    // ---- confidence of the prediction: amax softmax(log1p(out ** mult)) = 1 / sum exp(l - l_max) ----
    float lm = -INFINITY;
    for (int k = tid; k < p.K; k += EVAL_THREADS) lm = fmaxf(lm, log1pf(powf(ob[k], p.mult)));
    lm = warp_max(lm);
    if (lane == 0) s_val[warp] = lm;
    __syncthreads();
    if (tid == 0) {
        float v = s_val[0];
        for (int w = 1; w < EVAL_THREADS / 32; ++w) v = fmaxf(v, s_val[w]);
        s_f[0] = v;
    }
    __syncthreads();
    lm = s_f[0];
    float se = 0.f;
    for (int k = tid; k < p.K; k += EVAL_THREADS) se += expf(log1pf(powf(ob[k], p.mult)) - lm);
    se = warp_sum(se);
    __syncthreads();
    if (lane == 0) s_val[warp] = se;
    __syncthreads();
    if (tid == 0) {
        float v = 0.f;
        for (int w = 0; w < EVAL_THREADS / 32; ++w) v += s_val[w];
        p.pred_prob[b] = 1.0f / v;
    }

    // ---- rank of the label among the logits (ties broken towards the lower class index) ----
    long long y = -1;
    if (p.ys != nullptr) y = p.ys[b];
    const bool y_ok = y >= 0 && y < p.K;
    int rank = 0;
    if (y_ok) {
        const float vt = ob[y];
        for (int k = tid; k < p.K; k += EVAL_THREADS) {
            const float v = ob[k];
            rank += (v > vt || (v == vt && k < static_cast<int>(y))) ? 1 : 0;
        }
    }
    rank = block_sum_int(rank, &s_cnt[0]);

    // ---- explanation sizes: thread = (prototype lane, class slice); a warp covers 32 consecutive prototypes of one
    // class slice, so the weight reads are coalesced 128-byte rows and a warp whose 32 pooled values are all zero
    // (most of a PIP-Net's prototypes after the inference threshold) skips its classes altogether ----
    const float* pb = p.pooled + static_cast<size_t>(b) * p.P;
    for (int q = tid; q < p.P; q += EVAL_THREADS) s_any[q] = 0;
    __syncthreads();
    const int kslice = tid / EVAL_QLANES, ql = tid % EVAL_QLANES;
    const int kper = (p.K + EVAL_KSLICES - 1) / EVAL_KSLICES;
    const int k0 = kslice * kper, k1 = min(p.K, k0 + kper);
    int n_anz = 0, n_any = 0, n_pred = 0;
    for (int q0 = ql - lane; q0 < p.P; q0 += EVAL_QLANES) {  // q0: warp-uniform
        const int q = q0 + lane;
        const bool valid = q < p.P;
        const float pl = valid ? pb[q] : 0.f;
        if (kslice == 0) n_anz += fabsf(pl) > p.thr ? 1 : 0;
        if (!__any_sync(0xffffffffu, pl != 0.f)) continue;  // 0 * w is 0 (or NaN): never above the threshold
        const float* wq = p.w + (valid ? q : 0);
        bool any = false;
        int k = k0;
        for (; k + 4 <= k1; k += 4) {
            float wv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) wv[u] = __ldg(wq + static_cast<size_t>(k + u) * p.P);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float s = __fmul_rn(pl, wv[u]);
                if (fabsf(s) > p.thr) {
                    any = true;
                    n_pred += (k + u == pred) ? 1 : 0;
                }
                if (valid && __fsub_rn(s, p.thr) > 0.f) p.relevant[static_cast<size_t>(k + u) * p.P + q] = 1;
            }
        }
        for (; k < k1; ++k) {
            const float s = __fmul_rn(pl, __ldg(wq + static_cast<size_t>(k) * p.P));
            if (fabsf(s) > p.thr) {
                any = true;
                n_pred += (k == pred) ? 1 : 0;
            }
            if (valid && __fsub_rn(s, p.thr) > 0.f) p.relevant[static_cast<size_t>(k) * p.P + q] = 1;  // same value from every image
        }
        if (any && valid) s_any[q] = 1;  // several class slices may store the same 1
    }
    __syncthreads();
    for (int q = tid; q < p.P; q += EVAL_THREADS) n_any += s_any[q];
    n_anz = block_sum_int(n_anz, &s_cnt[1]);
    n_any = block_sum_int(n_any, &s_cnt[2]);
    n_pred = block_sum_int(n_pred, &s_cnt[3]);
    if (tid == 0) {
        p.anz[b] = n_anz;
        p.any_size[b] = n_any;
        p.pred_size[b] = n_pred;
        p.top1[b] = (y_ok && rank < 1) ? 1 : 0;
        p.top5[b] = (y_ok && rank < min(5, p.K)) ? 1 : 0;
        if (y_ok && p.cm != nullptr) atomicAdd(p.cm + static_cast<size_t>(y) * p.K + pred, 1ULL);
    }
}

// prototypes_per_class[k] = number of prototypes whose thresholded contribution to class k is non-zero for at least one
// image of the batch (test.py:108-113: the batch mean of relu(scores - thr) is > 0 exactly then).
__global__ void __launch_bounds__(128)
eval_class_count_kernel(const unsigned char* __restrict__ relevant, int* __restrict__ per_class, int P) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const unsigned char* r = relevant + static_cast<size_t>(blockIdx.x) * P;
    int n = 0;
    for (int q = threadIdx.x; q < P; q += blockDim.x) n += r[q] != 0 ? 1 : 0;
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0 && n != 0) atomicAdd(&s_cnt, n);
    __syncthreads();
    if (threadIdx.x == 0) per_class[blockIdx.x] = s_cnt;
}

}  // namespace cpn
