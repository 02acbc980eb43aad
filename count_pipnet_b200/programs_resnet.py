"""ResNet program: features/resnet_features.py (reference) as im2col gathers + tcgen05 GEMMs with eval-mode
BatchNorm folded into the weights (scale) and the epilogue bias, ReLU and the residual add fused in the epilogue.

Layout: channels-last. fp32 mode: the block-to-block residual stream is fp32 with bf16 hi / lo planes (the next conv's A
operand). bf16 mode: the stream itself is bf16 when every block end can run as a conv_gemm kernel (`_bf16_stream_ok`): the
identity is then read and the block output written as 2-byte values (the block-end 1x1 convs are HBM-bound), only the last
block writes the fp32 features the head reads. One bf16 rounding per block on the stream, the same rounding conv1 of the
next block applied to its input anyway.
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
# CPN_RESNET_BF16_RES=0: keep the fp32 residual stream in bf16 mode as well (A/B only)
USE_BF16_STREAM = os.environ.get("CPN_RESNET_BF16_RES", "1") != "0"


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

    def _bf16_stream_ok(self, H: int, W: int) -> bool:
        """True when every block end (and projection shortcut) qualifies for cpn_conv_bn_act(_resb) at the maps a
        H x W input produces: 1x1 or 3x3 stride-1 convs with channel counts in multiples of 64 and, for 3x3, a tile shape."""
        H, W = (H + 6 - 7) // 2 + 1, (W + 6 - 7) // 2 + 1
        H, W = (H + 2 - 3) // 2 + 1, (W + 2 - 3) // 2 + 1
        for blk in self.blocks:
            bott = isinstance(blk, rsn.Bottleneck)
            if not bott:
                # BasicBlock ends are 3x3 convs (compute-bound: little to gain), and the 8-block ResNet-18 golden case
                # measured 1.8e-2 / 2.1e-2 on pooled / logits with the bf16 stream against 0.9e-2 with the fp32 stream
                # (tolerance 2e-2); the Bottleneck nets move from 3.9e-3 to 4.6e-3 (profiles/r02_resnet_bf16_stream.md)
                return False
            strided = blk.conv2 if bott else blk.conv1
            k, st, pd = strided.kernel_size[0], strided.stride[0], strided.padding[0]
            H, W = (H + 2 * pd - k) // st + 1, (W + 2 * pd - k) // st + 1
            last = blk.conv3 if bott else blk.conv2
            k = last.kernel_size[0]
            if last.stride[0] != 1 or k not in (1, 3) or last.in_channels % 64 or last.out_channels % 64:
                return False
            if k == 3 and (last.padding[0] != 1 or ops.conv3x3_tile_rows(H, W) <= 0):
                return False
            if blk.downsample is not None:
                d = blk.downsample[0]
                if d.kernel_size[0] != 1 or d.stride[0] not in (1, 2) or d.in_channels % 64 or d.out_channels % 64:
                    return False
        return True

    def emit(self, em: Emit, x: torch.Tensor, B: int, H: int, W: int) -> Tuple[torch.Tensor, int, int, int]:
        net, pk, plan = self.net, self.packed, em.plan
        bf16_stream = (not em.split) and USE_CONV_GEMM and USE_BF16_STREAM and self._bf16_stream_ok(H, W)
        # conv1 7x7 s2 p3 + bn + relu: im2col from the NCHW fp32 input, K = 147 padded to 152
        H1, W1 = (H + 6 - 7) // 2 + 1, (W + 6 - 7) // 2 + 1
        M1 = B * H1 * W1
        c_hi, c_lo = em.pair("col0", (M1, 152))
        ops.im2col_nchw_f32(x, c_hi, B, 3, H, W, 7, 2, 3, 152, out_lo=c_lo, plan=plan)
        w_hi, w_lo, bias = pk["stem"]
        H2, W2 = (H1 + 2 - 3) // 2 + 1, (W1 + 2 - 3) // 2 + 1
        C = 64
        r_hi, r_lo = em.pair("rcur_a", (B * H2 * W2, C))
        if bf16_stream:
            # stem output and pooling in bf16: max commutes with the (monotonic) rounding, so r_hi has the same bits
            res = None
            s_b16 = em.buf("stem_b16", (M1, 64), torch.bfloat16)
            ops.conv_bn_act(c_hi, w_hi, bias, B, H1, W1, 152, 64, 1, relu=True, out_bf16=s_b16, plan=plan)
            ops.maxpool3s2_bf16(s_b16, r_hi, B, H1, W1, C, plan=plan)
        else:
            s_f32 = em.buf("stem_f32", (M1, 64), torch.float32)
            ops.gemm_bf16(c_hi, w_hi, M1, 64, 152, a_lo=c_lo, bw_lo=w_lo, bias=bias, out_f32=s_f32, act=_lib.ACT_RELU,
                          plan=plan)
            res = em.buf("res_a", (B * H2 * W2, C), torch.float32)
            ops.maxpool3s2(s_f32, B, H1, W1, C, out_f32=res, out_hi=r_hi, out_lo=r_lo, plan=plan)
        H, W = H2, W2
        flip = 0
        for bi, (blk, d) in enumerate(zip(self.blocks, pk["blocks"])):
            flip ^= 1
            sfx = "_b" if flip else "_a"
            bott = isinstance(blk, rsn.Bottleneck)
            if bf16_stream:
                # identity, conv chain and block end on the bf16 stream (csrc/conv_gemm.cuh CEPI_RESB_*)
                ident_b = r_hi
                if blk.downsample is not None:
                    dconv = blk.downsample[0]
                    st = dconv.stride[0]
                    Hd, Wd = (H - 1) // st + 1, (W - 1) // st + 1
                    xin = r_hi
                    if st == 2:
                        xin = em.buf("sub_hi", (B * Hd * Wd, dconv.in_channels), torch.bfloat16)
                        ops.subsample2_bf16(r_hi, xin, B, H, W, dconv.in_channels, plan=plan)
                    ident_b = em.buf("ident_b", (B * Hd * Wd, dconv.out_channels), torch.bfloat16)
                    dw_hi, _, dbias = d["down"]
                    ops.conv_bn_act(xin, dw_hi, dbias, B, Hd, Wd, dconv.in_channels, dconv.out_channels, 1, relu=False,
                                    out_bf16=ident_b, plan=plan)
                o_hi, o_lo, Ho, Wo = self._conv(em, "o1", r_hi, None, B, H, W, blk.conv1, d["convs"][0], relu=True)
                last = blk.conv3 if bott else blk.conv2
                if bott:
                    o_hi, o_lo, Ho, Wo = self._conv(em, "o2", o_hi, None, B, Ho, Wo, blk.conv2, d["convs"][1], relu=True)
                Cn = last.out_channels
                lw_hi, _, lbias = d["convs"][-1]
                if bi == len(self.blocks) - 1:
                    res = em.buf("res" + sfx, (B * Ho * Wo, Cn), torch.float32)
                    ops.conv_bn_act_resb(o_hi, lw_hi, lbias, ident_b, B, Ho, Wo, last.in_channels, Cn, last.kernel_size[0],
                                         out_f32=res, plan=plan)
                else:
                    r_hi = em.buf("rcur" + sfx + "_hi", (B * Ho * Wo, Cn), torch.bfloat16)
                    ops.conv_bn_act_resb(o_hi, lw_hi, lbias, ident_b, B, Ho, Wo, last.in_channels, Cn, last.kernel_size[0],
                                         out_bf16=r_hi, plan=plan)
                H, W, C = Ho, Wo, Cn
                continue
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
