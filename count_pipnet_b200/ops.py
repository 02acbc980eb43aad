"""Tensor-level wrappers over the C ABI (one function per entry point of include/cpn_b200.h).

Every function takes CUDA tensors that already have the layout the kernel wants, enqueues the kernel on torch's
current stream (or records it into `plan`) and returns nothing: outputs are caller-allocated. These wrappers do the
shape bookkeeping only; there is no torch arithmetic here and no fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import ptr


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.CpnError("count_pipnet_b200 kernels need CUDA tensors; there is no CPU fallback")
        if t is not None and not t.is_contiguous():
            raise _lib.CpnError("count_pipnet_b200 kernels need contiguous tensors")


class Plan:
    """A recorded launch sequence (cpn_plan). Tensors referenced by recorded ops must outlive the plan."""

    def __init__(self):
        h = C.c_void_p()
        _lib.check(_lib.load().cpn_plan_create(C.byref(h)))
        self.handle = h
        self.keepalive = []

    def run(self) -> None:
        _lib.check(_lib.load().cpn_plan_run(self.handle, _stream()))

    def capture(self) -> None:
        """Turn the recorded launches into one CUDA graph (captured on torch's current stream); run() replays it."""
        _lib.check(_lib.load().cpn_plan_capture(self.handle, _stream()))

    @property
    def has_graph(self) -> bool:
        return bool(_lib.load().cpn_plan_has_graph(self.handle))

    def __len__(self) -> int:
        return _lib.load().cpn_plan_num_ops(self.handle)

    def run_timed(self):
        """Replay with an event after every op; returns [(name, ms, flops, bytes)] (synchronises)."""
        lib = _lib.load()
        n = len(self)
        ms = (C.c_float * n)()
        _lib.check(lib.cpn_plan_run_timed(self.handle, _stream(), ms, n))
        out = []
        for i in range(n):
            name, fl, by = C.c_char_p(), C.c_double(), C.c_double()
            _lib.check(lib.cpn_plan_op_info(self.handle, i, C.byref(name), C.byref(fl), C.byref(by)))
            out.append((name.value.decode(), float(ms[i]), fl.value, by.value))
        return out

    def details(self):
        """Shape string of every recorded op (same order as run_timed)."""
        lib = _lib.load()
        return [(lib.cpn_plan_op_detail(self.handle, i) or b"").decode() for i in range(len(self))]

    def __del__(self):
        try:
            if self.handle:
                _lib.load().cpn_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def _ph(plan):
    return None if plan is None else plan.handle


def check_device() -> None:
    _lib.check(_lib.load().cpn_check_device())


def stem_ln(x, w, b, ln_w, ln_b, out, eps=1e-6, plan=None):
    """x [B,3,H,W] fp32 -> out [B,H/4,W/4,96] fp32 (channels-last)."""
    _need_cuda(x, w, b, ln_w, ln_b, out)
    B, _, H, W = x.shape
    _lib.call("cpn_stem_ln", _ph(plan), ptr(x), ptr(w), ptr(b), ptr(ln_w), ptr(ln_b), ptr(out), B, H, W, eps, _stream())


def dwconv7_ln(x, w49, b, ln_w, ln_b, out_hi, out_lo=None, eps=1e-6, plan=None):
    """x [B,H,W,C] fp32, w49 [49,C] -> out [B*H*W, C] bf16 (hi[, lo])."""
    _need_cuda(x, w49, b, ln_w, ln_b, out_hi, out_lo)
    B, H, W, Cc = x.shape
    _lib.call("cpn_dwconv7_ln", _ph(plan), ptr(x), ptr(w49), ptr(b), ptr(ln_w), ptr(ln_b), ptr(out_hi), ptr(out_lo),
              B, H, W, Cc, eps, _stream())


def dwconv7_stats(x, w49, b, out, stats, plan=None):
    """x [B,H,W,C] fp32 -> out [B*H*W, C] bf16 (conv + bias, not normalised), stats [B*H*W, C/32, 2] fp32."""
    _need_cuda(x, w49, b, out, stats)
    B, H, W, Cc = x.shape
    _lib.call("cpn_dwconv7_stats", _ph(plan), ptr(x), ptr(w49), ptr(b), ptr(out), ptr(stats), B, H, W, Cc, _stream())


def dwconv7_stats_split(x, w49, b, out_hi, out_lo, stats, plan=None):
    """fp32 mode: out_hi + out_lo [B*H*W, C] bf16 planes of conv + bias, stats of hi + lo per 32-channel group."""
    _need_cuda(x, w49, b, out_hi, out_lo, stats)
    B, H, W, Cc = x.shape
    _lib.call("cpn_dwconv7_stats_split", _ph(plan), ptr(x), ptr(w49), ptr(b), ptr(out_hi), ptr(out_lo), ptr(stats), B, H,
              W, Cc, _stream())


def gemm_ln_gelu_split(a_hi, a_lo, bw_hi, bw_lo, M, N, K, stats, nparts, colsum, bias, out_hi, out_lo, eps=1e-6, plan=None):
    """fp32 mode of gemm_ln_gelu: hi/lo operand planes, exact GELU, hi/lo output planes."""
    _need_cuda(a_hi, a_lo, bw_hi, bw_lo, stats, colsum, bias, out_hi, out_lo)
    _lib.call("cpn_gemm_ln_gelu_split", _ph(plan), ptr(a_hi), ptr(a_lo), ptr(bw_hi), ptr(bw_lo), M, N, K, ptr(stats),
              nparts, ptr(colsum), ptr(bias), float(eps), ptr(out_hi), ptr(out_lo), _stream())


def gemm_ln_gelu(a, bw, M, N, K, stats, nparts, colsum, bias, out_bf16, eps=1e-6, plan=None):
    """out = gelu(LayerNorm-folded fc1): see cpn_gemm_ln_gelu in include/cpn_b200.h."""
    _need_cuda(a, bw, stats, colsum, bias, out_bf16)
    _lib.call("cpn_gemm_ln_gelu", _ph(plan), ptr(a), ptr(bw), M, N, K, ptr(stats), nparts, ptr(colsum), ptr(bias),
              float(eps), ptr(out_bf16), _stream())


def mlp_fused(y_bf16, w1g, w2, M, C, stats, colsum, bias1, bias2, scale, x, eps=1e-6, plan=None):
    """x += scale * fc2(gelu(fc1(LN(y)))) in one kernel: see cpn_mlp_fused in include/cpn_b200.h."""
    _need_cuda(y_bf16, w1g, w2, stats, colsum, bias1, bias2, scale, x)
    _lib.call("cpn_mlp_fused", _ph(plan), ptr(y_bf16), ptr(w1g), ptr(w2), M, C, ptr(stats), C // 32, ptr(colsum),
              ptr(bias1), ptr(bias2), ptr(scale), float(eps), ptr(x), _stream())


def ln_im2col2(x, ln_w, ln_b, out_hi, out_lo=None, stride=2, eps=1e-6, plan=None):
    """x [B,H,W,C] fp32 -> out [B*Ho*Wo, 4C] bf16."""
    _need_cuda(x, ln_w, ln_b, out_hi, out_lo)
    B, H, W, Cc = x.shape
    _lib.call("cpn_ln_im2col2", _ph(plan), ptr(x), ptr(ln_w), ptr(ln_b), ptr(out_hi), ptr(out_lo), B, H, W, Cc, stride,
              eps, _stream())


def gemm_bf16(a, bw, M, N, K, *, a_lo=None, bw_lo=None, bias=None, scale=None, res_f32=None, res_bf16=None,
              out_f32=None, out_bf16=None, out_bf16_lo=None, act=_lib.ACT_NONE, relu_after_res=False, plan=None):
    """out = epilogue(a[M,K] @ bw[N,K]^T); a, bw bf16 row-major."""
    _need_cuda(a, bw, a_lo, bw_lo, bias, scale, res_f32, res_bf16, out_f32, out_bf16, out_bf16_lo)
    _lib.call("cpn_gemm_bf16", _ph(plan), ptr(a), ptr(a_lo), ptr(bw), ptr(bw_lo), M, N, K, ptr(bias), ptr(scale),
              ptr(res_f32), ptr(res_bf16), ptr(out_f32), ptr(out_bf16), ptr(out_bf16_lo), int(act),
              int(bool(relu_after_res)), _stream())


def addon_conv1x1(feat, w, b, z, T, Cc, P, plan=None):
    _need_cuda(feat, w, b, z)
    _lib.call("cpn_addon_conv1x1", _ph(plan), ptr(feat), ptr(w), ptr(b), ptr(z), T, Cc, P, _stream())


def addon_softmax_pool(feat, w, b, partial, B, HW, Cc, P, chunks, *, gumbel=None, tau=1.0, hard=False,
                       mode=_lib.POOL_MAX, map_out=None, partial_arg=None, plan=None):
    """Add-on 1x1 conv + softmax / hard Gumbel + spatial pooling in one kernel (P = 16 or 32): see the header."""
    _need_cuda(feat, w, b, partial, gumbel, map_out, partial_arg)
    _lib.call("cpn_addon_softmax_pool", _ph(plan), ptr(feat), ptr(w), ptr(b), ptr(gumbel), float(tau), int(bool(hard)),
              int(mode), ptr(map_out), ptr(partial), ptr(partial_arg), B, HW, Cc, P, chunks, _stream())


def softmax_pool(z, partial, B, HW, P, chunks, *, gumbel=None, tau=1.0, hard=False, mode=_lib.POOL_MAX, map_out=None,
                 partial_arg=None, plan=None):
    """partial_arg [B,chunks,P] int32 (max pooling): token of every chunk maximum, see pool_argmax_reduce."""
    _need_cuda(z, partial, gumbel, map_out, partial_arg)
    _lib.call("cpn_softmax_pool_arg", _ph(plan), ptr(z), ptr(gumbel), float(tau), int(bool(hard)), int(mode),
              ptr(map_out), ptr(partial), ptr(partial_arg), B, HW, P, chunks, _stream())


def pool_argmax_reduce(partial, partial_arg, chunks, arg_out, B, P, plan=None):
    """arg_out [B,P] int32: token (h * W + w) of each prototype's spatial maximum per image."""
    _need_cuda(partial, partial_arg, arg_out)
    _lib.call("cpn_pool_argmax_reduce", _ph(plan), ptr(partial), ptr(partial_arg), chunks, ptr(arg_out), B, P, _stream())


def head_tail(partial, chunks, pooled_out, logits_out, cls_w, cls_b, B, P, D, K, *, mode=_lib.POOL_MAX,
              inference=True, use_ste=False, max_count=3, encoding=_lib.ENC_NONE, enc_w0=None, enc_w1=None,
              enc_w2=None, plan=None):
    _need_cuda(partial, pooled_out, logits_out, cls_w, cls_b, enc_w0, enc_w1, enc_w2)
    _lib.call("cpn_head_tail", _ph(plan), ptr(partial), chunks, int(mode), int(bool(inference)), int(bool(use_ste)),
              int(max_count), int(encoding), ptr(enc_w0), ptr(enc_w1), ptr(enc_w2), ptr(cls_w), ptr(cls_b),
              ptr(pooled_out), ptr(logits_out), B, P, D, K, _stream())


def nhwc_to_nchw(x, out, B, HW, Cc, plan=None):
    _need_cuda(x, out)
    _lib.call("cpn_nhwc_to_nchw", _ph(plan), ptr(x), ptr(out), B, HW, Cc, _stream())


def nchw_to_nhwc(x, out, B, HW, Cc, plan=None):
    _need_cuda(x, out)
    _lib.call("cpn_nchw_to_nhwc", _ph(plan), ptr(x), ptr(out), B, HW, Cc, _stream())


def conv3x3_tile_rows(H: int, W: int) -> int:
    """Image rows per M tile of the implicit 3x3 conv, 0 if this map size needs the im2col path."""
    return int(_lib.load().cpn_conv3x3_tile_rows(int(H), int(W)))


def conv_bn_act(x_bf16, w_bf16, bias, B, H, W, Cin, Cout, ksize, *, relu, res_f32=None, out_f32=None, out_bf16=None,
                plan=None):
    """ResNet conv + folded BN (+ReLU / + identity) as one GEMM without im2col: see cpn_conv_bn_act in the header."""
    _need_cuda(x_bf16, w_bf16, bias, res_f32, out_f32, out_bf16)
    _lib.call("cpn_conv_bn_act", _ph(plan), ptr(x_bf16), ptr(w_bf16), ptr(bias), ptr(res_f32), ptr(out_f32),
              ptr(out_bf16), B, H, W, Cin, Cout, int(ksize), int(bool(relu)), _stream())


def conv_bn_act_resb(x_bf16, w_bf16, bias, res_bf16, B, H, W, Cin, Cout, ksize, *, out_f32=None, out_bf16=None, plan=None):
    """Block-end conv + folded BN + bf16 identity + ReLU (bf16 residual stream): see cpn_conv_bn_act_resb in the header."""
    _need_cuda(x_bf16, w_bf16, bias, res_bf16, out_f32, out_bf16)
    _lib.call("cpn_conv_bn_act_resb", _ph(plan), ptr(x_bf16), ptr(w_bf16), ptr(bias), ptr(res_bf16), ptr(out_f32),
              ptr(out_bf16), B, H, W, Cin, Cout, int(ksize), _stream())


def im2col_bf16(x, out, B, H, W, Cc, k, stride, pad, Kpad, plan=None):
    _need_cuda(x, out)
    _lib.call("cpn_im2col_bf16", _ph(plan), ptr(x), ptr(out), B, H, W, Cc, k, stride, pad, Kpad, _stream())


def im2col_nchw_f32(x, out, B, Cc, H, W, k, stride, pad, Kpad, out_lo=None, plan=None):
    _need_cuda(x, out, out_lo)
    _lib.call("cpn_im2col_nchw_f32", _ph(plan), ptr(x), ptr(out), ptr(out_lo), B, Cc, H, W, k, stride, pad, Kpad,
              _stream())


def maxpool3s2_bf16(x, out, B, H, W, Cc, plan=None):
    _need_cuda(x, out)
    _lib.call("cpn_maxpool3s2_bf16", _ph(plan), ptr(x), ptr(out), B, H, W, Cc, _stream())


def maxpool3s2(x, B, H, W, Cc, out_f32=None, out_hi=None, out_lo=None, plan=None):
    _need_cuda(x, out_f32, out_hi, out_lo)
    _lib.call("cpn_maxpool3s2", _ph(plan), ptr(x), ptr(out_f32), ptr(out_hi), ptr(out_lo), B, H, W, Cc, _stream())


def subsample2_bf16(x, out, B, H, W, Cc, plan=None):
    _need_cuda(x, out)
    _lib.call("cpn_subsample2_bf16", _ph(plan), ptr(x), ptr(out), B, H, W, Cc, _stream())


def eval_batch(pooled, out, weights, ys, mult, thr, res, cm=None, plan=None):
    """Per-batch evaluation statistics: see cpn_eval_batch in include/cpn_b200.h. `res` is a dict of caller-allocated
    outputs (pred, pred_prob, max_score, any_size, pred_size, anz, top1, top5 [B]; relevant [K, P] uint8;
    protos_per_class [K]); cm [K, K] int64 accumulates."""
    _need_cuda(pooled, out, weights, ys, cm, *res.values())
    B, P = pooled.shape
    K = out.shape[1]
    if tuple(weights.shape) != (K, P):
        raise _lib.CpnError(f"eval_batch: weights {tuple(weights.shape)} != (K, P) = ({K}, {P})")
    _lib.call("cpn_eval_batch", _ph(plan), ptr(pooled), ptr(out), ptr(weights), ptr(ys), float(mult), float(thr), B, P,
              K, ptr(res["pred"]), ptr(res["pred_prob"]), ptr(res["max_score"]), ptr(res["any_size"]),
              ptr(res["pred_size"]), ptr(res["anz"]), ptr(res["top1"]), ptr(res["top5"]), ptr(cm),
              ptr(res["relevant"]), ptr(res["protos_per_class"]), _stream())


def topk_update(pooled, argmax, image_base, top_val, top_img, top_tok, top_cnt, plan=None):
    """Merges a batch into the running per-prototype top-k lists [k, P]: see cpn_topk_update in include/cpn_b200.h."""
    _need_cuda(pooled, argmax, top_val, top_img, top_tok, top_cnt)
    B, P = pooled.shape
    _lib.call("cpn_topk_update", _ph(plan), ptr(pooled), ptr(argmax), B, P, top_val.shape[0], int(image_base),
              ptr(top_val), ptr(top_img), ptr(top_tok), ptr(top_cnt), _stream())


# ---- training side of the head (no plans) --------------------------------------------------------------------------
def pipnet_loss_ws_floats(B, P, HW) -> int:
    return int(_lib.load().cpn_pipnet_loss_ws_floats(B, P, HW))


def pipnet_loss(pf, pooled, out, ys, norm_mult, B, P, HW, K, align_w, tanh_w, class_w, tanh_coeff, eps_tanh, pretrain,
                finetune, use_log1p, grad_pf, grad_pooled, grad_out, result, workspace):
    """Loss values and gradients of pipnet/train.py's calculate_loss: see cpn_pipnet_loss in include/cpn_b200.h."""
    _need_cuda(pf, pooled, out, ys, norm_mult, grad_pf, grad_pooled, grad_out, result, workspace)
    _lib.call("cpn_pipnet_loss", ptr(pf), ptr(pooled), ptr(out), ptr(ys), ptr(norm_mult), B, P, HW, K, float(align_w),
              float(tanh_w), float(class_w), float(tanh_coeff), float(eps_tanh), int(bool(pretrain)),
              int(bool(finetune)), int(bool(use_log1p)), ptr(grad_pf), ptr(grad_pooled), ptr(grad_out), ptr(result),
              ptr(workspace), _stream())


def scale_unless_one(buf, scale):
    _need_cuda(buf, scale)
    _lib.call("cpn_scale_unless_one", ptr(buf), buf.numel(), ptr(scale), _stream())


def softmax_pool_nchw_ws_floats(N, P, HW) -> int:
    return int(_lib.load().cpn_softmax_pool_nchw_ws_floats(N, P, HW))


def softmax_pool_nchw(z, mode, pf, pooled, argmax, workspace):
    """z [N, P, H, W] -> pf (softmax over P), pooled [N, P] (max with argmax, or sum): cpn_softmax_pool_nchw."""
    _need_cuda(z, pf, pooled, argmax, workspace)
    N, P = z.shape[0], z.shape[1]
    _lib.call("cpn_softmax_pool_nchw", ptr(z), mode, ptr(pf), ptr(pooled), ptr(argmax), N, P, z[0, 0].numel(),
              ptr(workspace), _stream())


def softmax_pool_nchw_bwd(pf, grad_pf, grad_pooled, argmax, mode, grad_z):
    _need_cuda(pf, grad_pf, grad_pooled, argmax, grad_z)
    N, P = pf.shape[0], pf.shape[1]
    _lib.call("cpn_softmax_pool_nchw_bwd", ptr(pf), ptr(grad_pf), ptr(grad_pooled), ptr(argmax), mode, ptr(grad_z), N, P,
              pf[0, 0].numel(), _stream())
