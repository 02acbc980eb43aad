"""ResNet program: features/resnet_features.py (reference) as im2col gathers + tcgen05 GEMMs with eval-mode
BatchNorm folded into the weights (scale) and the epilogue bias, ReLU and the residual add fused in the epilogue.

Layout: channels-last; the block-to-block residual stream is fp32 with a bf16 copy (the next conv's A operand).
"""
from __future__ import annotations

import os
from typing import List, Tuple

import torch
import torch.nn as nn

from . import _lib, ops
from .features import resnet_features as rsn
from .programs import Emit, _split


# CPN_CONV_GEMM=0: every conv through im2col + cpn_gemm_bf16 as in round 1 (A/B only)
USE_CONV_GEMM = os.environ.get("CPN_CONV_GEMM", "1") != "0"


def _fold(conv: nn.Conv2d, bn: nn.BatchNorm2d, split: bool, k_order: str, kpad: int = 0):
    """conv weight * bn scale -> [Cout, K] bf16 (hi, lo) and the fp32 bias of the folded affine
    (features/resnet_features.py:97-117: BN in eval mode is y = (x - mean) / sqrt(var + eps) * w + b)."""
    scale = bn.weight.detach() / torch.sqrt(bn.running_var.detach() + bn.eps)
    bias = (bn.bias.detach() - bn.running_mean.detach() * scale).contiguous()
    w = conv.weight.detach() * scale.view(-1, 1, 1, 1)
    if k_order == "khwc":      # matches cpn_im2col_bf16 columns (kh, kw, c)
        w = w.permute(0, 2, 3, 1).reshape(w.shape[0], -1)
    else:                       # "chw": matches cpn_im2col_nchw_f32 columns (c, kh, kw)
        w = w.reshape(w.shape[0], -1)
    if kpad > w.shape[1]:
        w = torch.cat([w, w.new_zeros(w.shape[0], kpad - w.shape[1])], dim=1)
    hi, lo = _split(w.contiguous(), split)
    return hi, lo, bias


class ResNetProgram:
    def __init__(self, net: rsn.ResNet_features):
        self.net = net
        self.blocks: List[nn.Module] = []
        for layer in (net.layer1, net.layer2, net.layer3, net.layer4):
            self.blocks.extend(list(layer.children()))
        last = self.blocks[-1]
        self.out_channels = (last.conv3 if isinstance(last, rsn.Bottleneck) else last.conv2).out_channels
        self.packed = None

    def pack(self, split: bool) -> None:
        net = self.net
        pk = {"stem": _fold(net.conv1, net.bn1, split, "chw", kpad=152), "blocks": []}
        for blk in self.blocks:
            d = {}
            convs = [(blk.conv1, blk.bn1), (blk.conv2, blk.bn2)]
            if isinstance(blk, rsn.Bottleneck):
                convs.append((blk.conv3, blk.bn3))
            d["convs"] = [_fold(c, b, split, "khwc") for c, b in convs]
            if blk.downsample is not None:
                d["down"] = _fold(blk.downsample[0], blk.downsample[1], split, "khwc")
            pk["blocks"].append(d)
        self.packed = pk

    # one conv + folded BN (+ReLU / + residual) on a channels-last bf16 map
    def _conv(self, em: Emit, tag: str, x_hi, x_lo, B, H, W, conv: nn.Conv2d, folded, *, relu: bool,
              res_f32=None, out_f32=None, want_bf16: bool = True):
        k, s, p = conv.kernel_size[0], conv.stride[0], conv.padding[0]
        Cin, Cout = conv.in_channels, conv.out_channels
        Ho, Wo = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        M = B * Ho * Wo
        w_hi, w_lo, bias = folded
        # bf16 mode: conv + BN (+ReLU / +identity) as one GEMM with a TMA epilogue; 3x3 taps read straight from the
        # activation (csrc/conv_gemm.cuh). Strided convs gather first (1x1: subsample) or keep the im2col path (3x3).
        direct = (not em.split and USE_CONV_GEMM and Cout % 64 == 0 and Cin % 64 == 0 and
                  ((k == 1) or (k == 3 and s == 1 and p == 1 and ops.conv3x3_tile_rows(H, W) > 0)) and
                  ((relu and res_f32 is None and out_f32 is None and want_bf16) or
                   (relu and res_f32 is not None and out_f32 is not None and want_bf16) or
                   (not relu and res_f32 is None and out_f32 is not None and not want_bf16)))
        if direct:
            xin = x_hi
            if k == 1 and s == 2:
                xin = em.buf("sub_hi", (M, Cin), torch.bfloat16)
                ops.subsample2_bf16(x_hi, xin, B, H, W, Cin, plan=em.plan)
            elif k == 1 and s != 1:
                direct = False
        if direct:
            o_hi = em.buf(tag + "_hi", (M, Cout), torch.bfloat16) if want_bf16 else None
            ops.conv_bn_act(xin, w_hi, bias, B, Ho, Wo, Cin, Cout, k, relu=relu, res_f32=res_f32, out_f32=out_f32,
                            out_bf16=o_hi, plan=em.plan)
            return o_hi, None, Ho, Wo
        if k == 1 and s == 1:
            a_hi, a_lo, K = x_hi, x_lo, Cin
        elif k == 1 and s == 2:
            a_hi, a_lo = em.pair("sub", (M, Cin))
            ops.subsample2_bf16(x_hi, a_hi, B, H, W, Cin, plan=em.plan)
            if a_lo is not None:
                ops.subsample2_bf16(x_lo, a_lo, B, H, W, Cin, plan=em.plan)
            K = Cin
        else:
            K = k * k * Cin
            a_hi, a_lo = em.pair("col", (M, K))
            ops.im2col_bf16(x_hi, a_hi, B, H, W, Cin, k, s, p, K, plan=em.plan)
            if a_lo is not None:
                ops.im2col_bf16(x_lo, a_lo, B, H, W, Cin, k, s, p, K, plan=em.plan)
        o_hi, o_lo = em.pair(tag, (M, Cout)) if want_bf16 else (None, None)
        ops.gemm_bf16(a_hi, w_hi, M, Cout, K, a_lo=a_lo, bw_lo=w_lo, bias=bias, res_f32=res_f32, out_f32=out_f32,
                      out_bf16=o_hi, out_bf16_lo=o_lo, act=_lib.ACT_RELU if relu else _lib.ACT_NONE,
                      relu_after_res=res_f32 is not None, plan=em.plan)
        return o_hi, o_lo, Ho, Wo

    def emit(self, em: Emit, x: torch.Tensor, B: int, H: int, W: int) -> Tuple[torch.Tensor, int, int, int]:
        net, pk, plan = self.net, self.packed, em.plan
        # conv1 7x7 s2 p3 + bn + relu: im2col from the NCHW fp32 input, K = 147 padded to 152
        H1, W1 = (H + 6 - 7) // 2 + 1, (W + 6 - 7) // 2 + 1
        M1 = B * H1 * W1
        c_hi, c_lo = em.pair("col0", (M1, 152))
        ops.im2col_nchw_f32(x, c_hi, B, 3, H, W, 7, 2, 3, 152, out_lo=c_lo, plan=plan)
        w_hi, w_lo, bias = pk["stem"]
        s_f32 = em.buf("stem_f32", (M1, 64), torch.float32)
        ops.gemm_bf16(c_hi, w_hi, M1, 64, 152, a_lo=c_lo, bw_lo=w_lo, bias=bias, out_f32=s_f32, act=_lib.ACT_RELU,
                      plan=plan)
        H2, W2 = (H1 + 2 - 3) // 2 + 1, (W1 + 2 - 3) // 2 + 1
        C = 64
        res = em.buf("res_a", (B * H2 * W2, C), torch.float32)
        r_hi, r_lo = em.pair("rcur_a", (B * H2 * W2, C))
        ops.maxpool3s2(s_f32, B, H1, W1, C, out_f32=res, out_hi=r_hi, out_lo=r_lo, plan=plan)
        H, W = H2, W2
        flip = 0
        for blk, d in zip(self.blocks, pk["blocks"]):
            flip ^= 1
            sfx = "_b" if flip else "_a"
            bott = isinstance(blk, rsn.Bottleneck)
            if blk.downsample is not None:
                dconv = blk.downsample[0]
                so = (H - 1) // dconv.stride[0] + 1
                ident = em.buf("ident", (B * so * ((W - 1) // dconv.stride[0] + 1), dconv.out_channels), torch.float32)
                self._conv(em, "unused", r_hi, r_lo, B, H, W, dconv, d["down"], relu=False, out_f32=ident,
                           want_bf16=False)
            else:
                ident = res
            o_hi, o_lo, Ho, Wo = self._conv(em, "o1", r_hi, r_lo, B, H, W, blk.conv1, d["convs"][0], relu=True)
            last = blk.conv3 if bott else blk.conv2
            if bott:
                o_hi, o_lo, Ho, Wo = self._conv(em, "o2", o_hi, o_lo, B, Ho, Wo, blk.conv2, d["convs"][1], relu=True)
            Cn = last.out_channels
            # the strided conv (conv2 of a Bottleneck, conv1 of a BasicBlock) has run: (Ho, Wo) is the block's output grid
            nres = em.buf("res" + sfx, (B * Ho * Wo, Cn), torch.float32)
            r_hi, r_lo, Ho, Wo = self._conv(em, "rcur" + sfx, o_hi, o_lo, B, Ho, Wo, last, d["convs"][-1], relu=True,
                                            res_f32=ident, out_f32=nres)
            res, H, W, C = nres, Ho, Wo, Cn
        return res, H, W, C
