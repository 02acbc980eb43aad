// C-ABI entry points of the training side of the prototype head (include/cpn_b200.h, "training" section): the PIP-Net
// losses with their gradients (reference pipnet/train.py:165-243, 253-265) and softmax + pooling forward / backward on
// the NCHW map of the autograd path (pipnet/pipnet.py:31-35, pipnet/count_pipnet.py:356-360, 88). Nothing is recorded
// in plans here: a training step changes its weights every iteration and is not replayed.
#include "api_common.h"
#include "train_kernels.cuh"

#include <cstdint>
#include <initializer_list>

using namespace cpn_api;

namespace {

inline int tiles_of(int HW) { return (HW + cpn::TRAIN_TILE - 1) / cpn::TRAIN_TILE; }

// The map kernels read 4 neighbouring positions per lane when rows and bases allow 16-byte accesses.
inline bool aligned16(const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int vec_of(int HW, std::initializer_list<const void*> maps) {
    if (HW % 4 != 0) return 1;
    for (const void* p : maps)
        if (!aligned16(p)) return 1;
    return 4;
}
// Threads per CTA: (32 / vec) lanes per channel row x as many channel groups as there are channels, at most 512 threads,
// a whole number of warps.
inline int threads_of(int vec, int P) {
    const int L = 32 / vec, cg_max = cpn::TRAIN_THREADS / L;
    int cg = P < cg_max ? P : cg_max;
    int threads = L * cg;
    threads = (threads + 31) / 32 * 32;
    return threads;
}
inline int tanh_blocks(int P) { return (P + 127) / 128; }

struct LossWs {  // layout of the caller's workspace, in floats
    long long align, tanh, row_loss, row_gmult, row_correct, ticket, total;
};
LossWs loss_ws(int B, int P, int HW) {
    LossWs w;
    w.align = 0;
    w.tanh = w.align + static_cast<long long>(B) * tiles_of(HW);
    w.row_loss = w.tanh + 2LL * tanh_blocks(P);
    w.row_gmult = w.row_loss + 2LL * B;
    w.row_correct = w.row_gmult + 2LL * B;
    w.ticket = w.row_correct + 2LL * B;
    w.total = w.ticket + 1;
    return w;
}

}  // namespace

extern "C" long long cpn_pipnet_loss_ws_floats(int B, int P, int HW) {
    if (B <= 0 || P <= 0 || HW <= 0) return -1;
    return loss_ws(B, P, HW).total;
}

extern "C" int cpn_pipnet_loss(const float* proto_features, const float* pooled, const float* out, const long long* ys,
                               const float* norm_mult, int B, int P, int HW, int K, float align_w, float tanh_w,
                               float class_w, float tanh_coeff, float eps_tanh, int pretrain, int finetune,
                               int use_log1p, float* grad_pf, float* grad_pooled, float* grad_out, float* result,
                               float* workspace, void* stream) {
    if (proto_features == nullptr || pooled == nullptr || result == nullptr || workspace == nullptr)
        return fail("cpn_pipnet_loss: missing operand");
    if (B <= 0 || B > 32767 || P <= 0 || HW <= 0) return fail("cpn_pipnet_loss: bad B=%d P=%d HW=%d", B, P, HW);
    if (!pretrain) {
        if (out == nullptr || ys == nullptr || K <= 0) return fail("cpn_pipnet_loss: out / ys / K needed unless pretrain");
        if (use_log1p && norm_mult == nullptr) return fail("cpn_pipnet_loss: norm_mult needed for log1p(out^m)");
    }
    if (pretrain && finetune) return fail("cpn_pipnet_loss: pretrain and finetune exclude each other");
    if (finetune && (grad_pf != nullptr || grad_pooled != nullptr))
        return fail("cpn_pipnet_loss: the fine-tuning loss has no align / tanh term (train.py:224-225); pass NULL for "
                    "grad_pf and grad_pooled");
    const LossWs w = loss_ws(B, P, HW);
    const int tiles = tiles_of(HW);
    const double n_patches = static_cast<double>(B) * HW;
    const float align_eps = 1e-12f;  // align_loss's own default (train.py:254), not calculate_loss's EPS
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    set_info("pipnet_loss", 4.0 * B * P * HW, (grad_pf ? 16.0 : 8.0) * B * P * HW);

    // One kernel: the first CTAs do the tanh / classification terms on 128 threads while the others stream the map; the
    // CTA that takes the last ticket does the weighting. The ticket is zeroed by a memset node ahead of it.
    cpn::LossTermsArgs a;
    unsigned* ticket = reinterpret_cast<unsigned*>(workspace + w.ticket);
    a.pooled = pooled;
    a.out = out;
    a.ys = ys;
    a.mult = norm_mult;
    a.grad_pooled = grad_pooled;
    a.grad_out = grad_out;
    a.align_partial = workspace + w.align;
    a.tanh_partial = workspace + w.tanh;
    a.row_loss = workspace + w.row_loss;
    a.row_gmult = workspace + w.row_gmult;
    a.row_correct = workspace + w.row_correct;
    a.ticket = ticket;
    a.result = result;
    a.B = B, a.P = P, a.K = K, a.n_align = B * tiles, a.tanh_blocks = tanh_blocks(P);
    a.use_log1p = use_log1p, a.pretrain = pretrain, a.finetune = finetune;
    a.coeff = tanh_coeff, a.eps = eps_tanh;
    a.tanh_scale = static_cast<float>(-0.5 * tanh_w / P);
    a.class_scale = static_cast<float>(class_w / (2.0 * B));
    a.n_patches = static_cast<float>(n_patches);
    a.align_w = align_w, a.tanh_w = tanh_w, a.class_w = class_w;
    const int n_terms = 2 * tanh_blocks(P) + (pretrain ? 0 : 2 * B);
    const long long ctas = static_cast<long long>(n_terms) + static_cast<long long>(tiles) * B;
    if (ctas > 0x7fffffffLL) return fail("cpn_pipnet_loss: %lld CTAs", ctas);
    const int vec = vec_of(HW, {proto_features, grad_pf});
    const float align_scale = static_cast<float>(-0.5 * align_w / n_patches);
    int threads = threads_of(vec, P);
    if (threads < 128) threads = 128;  // the terms roles and the weighting run on 128 threads
    CPN_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned), st));
    if (vec == 4)
        cpn::pipnet_loss_kernel<4><<<static_cast<unsigned>(ctas), threads, 0, st>>>(
            proto_features, grad_pf, workspace + w.align, n_terms, tiles, B, P, HW, align_eps, align_scale, a);
    else
        cpn::pipnet_loss_kernel<1><<<static_cast<unsigned>(ctas), threads, 0, st>>>(
            proto_features, grad_pf, workspace + w.align, n_terms, tiles, B, P, HW, align_eps, align_scale, a);
    CPN_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int cpn_scale_unless_one(float* buf, long long n, const float* scale, void* stream) {
    if (buf == nullptr || scale == nullptr || n < 0) return fail("cpn_scale_unless_one: bad operand");
    if (n == 0) return 0;
    long long blocks = (n + 1023) / 1024;
    const long long cap = 8LL * num_sms();
    if (blocks > cap) blocks = cap;
    cpn::scale_unless_one_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(buf, n, scale);
    CPN_CUDA(cudaGetLastError());
    return 0;
}

extern "C" long long cpn_softmax_pool_nchw_ws_floats(int N, int P, int HW) {
    if (N <= 0 || P <= 0 || HW <= 0) return -1;
    return 2LL * N * tiles_of(HW) * P;  // values, then indices (int32 in the same words)
}

extern "C" int cpn_softmax_pool_nchw(const float* z, int mode, float* proto_features, float* pooled, int* argmax, int N,
                                     int P, int HW, float* workspace, void* stream) {
    if (z == nullptr || proto_features == nullptr || pooled == nullptr || workspace == nullptr)
        return fail("cpn_softmax_pool_nchw: missing operand");
    if (mode != CPN_POOL_MAX && mode != CPN_POOL_SUM) return fail("cpn_softmax_pool_nchw: mode %d", mode);
    if (mode == CPN_POOL_MAX && argmax == nullptr) return fail("cpn_softmax_pool_nchw: max pooling needs argmax [N, P]");
    if (N <= 0 || N > 65535 || P <= 0 || HW <= 0) return fail("cpn_softmax_pool_nchw: bad N=%d P=%d HW=%d", N, P, HW);
    const int tiles = tiles_of(HW);
    float* part_val = workspace;
    int* part_idx = reinterpret_cast<int*>(workspace + static_cast<size_t>(N) * tiles * P);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    set_info("softmax_pool_nchw", 0.0, 8.0 * N * P * HW);
    int* idx = mode == CPN_POOL_MAX ? part_idx : nullptr;
    const int vec = vec_of(HW, {z, proto_features});
    const dim3 grid(tiles, N);
    const int threads = threads_of(vec, P), sum = mode == CPN_POOL_SUM;
    // the [P, 32] tile of z in shared memory when two CTAs of it fit one SM (P <= 768), else 512 staging rows
    const size_t tile_bytes = static_cast<size_t>(P) * cpn::TRAIN_TILE * sizeof(float);
    const bool cache = tile_bytes <= 96 * 1024;
    const size_t smem = cache ? tile_bytes
                              : static_cast<size_t>(P < cpn::STAGE_ROWS ? P : cpn::STAGE_ROWS) * cpn::TRAIN_TILE * sizeof(float);
    if (cache && vec == 4) {
        CPN_CUDA(ensure_smem(cpn::softmax_pool_nchw_kernel<4, true>, smem));
        cpn::softmax_pool_nchw_kernel<4, true><<<grid, threads, smem, st>>>(z, proto_features, part_val, idx, sum, P, HW);
    } else if (cache) {
        CPN_CUDA(ensure_smem(cpn::softmax_pool_nchw_kernel<1, true>, smem));
        cpn::softmax_pool_nchw_kernel<1, true><<<grid, threads, smem, st>>>(z, proto_features, part_val, idx, sum, P, HW);
    } else if (vec == 4) {
        CPN_CUDA(ensure_smem(cpn::softmax_pool_nchw_kernel<4, false>, smem));
        cpn::softmax_pool_nchw_kernel<4, false><<<grid, threads, smem, st>>>(z, proto_features, part_val, idx, sum, P, HW);
    } else {
        CPN_CUDA(ensure_smem(cpn::softmax_pool_nchw_kernel<1, false>, smem));
        cpn::softmax_pool_nchw_kernel<1, false><<<grid, threads, smem, st>>>(z, proto_features, part_val, idx, sum, P, HW);
    }
    CPN_CUDA(cudaGetLastError());
    cpn::pool_fold_nchw_kernel<<<dim3((P + 127) / 128, N), 128, 0, st>>>(part_val, part_idx, mode == CPN_POOL_SUM, pooled,
                                                                        argmax, P, tiles);
    CPN_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int cpn_softmax_pool_nchw_bwd(const float* proto_features, const float* grad_pf, const float* grad_pooled,
                                         const int* argmax, int mode, float* grad_z, int N, int P, int HW,
                                         void* stream) {
    if (proto_features == nullptr || grad_z == nullptr) return fail("cpn_softmax_pool_nchw_bwd: missing operand");
    if (mode != CPN_POOL_MAX && mode != CPN_POOL_SUM) return fail("cpn_softmax_pool_nchw_bwd: mode %d", mode);
    if (mode == CPN_POOL_MAX && grad_pooled != nullptr && argmax == nullptr)
        return fail("cpn_softmax_pool_nchw_bwd: max pooling needs the forward's argmax");
    if (N <= 0 || N > 65535 || P <= 0 || HW <= 0) return fail("cpn_softmax_pool_nchw_bwd: bad N=%d P=%d HW=%d", N, P, HW);
    set_info("softmax_pool_nchw_bwd", 0.0, (grad_pf ? 12.0 : 8.0) * N * P * HW);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec_of(HW, {proto_features, grad_pf, grad_z}) == 4)
        cpn::softmax_pool_nchw_bwd_kernel<4><<<dim3(tiles_of(HW), N), threads_of(4, P), 0, st>>>(
            proto_features, grad_pf, grad_pooled, argmax, mode == CPN_POOL_SUM, grad_z, P, HW);
    else
        cpn::softmax_pool_nchw_bwd_kernel<1><<<dim3(tiles_of(HW), N), threads_of(1, P), 0, st>>>(
            proto_features, grad_pf, grad_pooled, argmax, mode == CPN_POOL_SUM, grad_z, P, HW);
    CPN_CUDA(cudaGetLastError());
    return 0;
}
