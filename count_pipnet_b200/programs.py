"""Backbone "programs": they read a reference-shaped backbone module (ConvNeXt-tiny variants, ResNets), repack its
weights for the kernels once per weight version, and emit the kernel sequence for a given input shape - either
launching immediately or recording into a plan. Feature maps are channels-last, residual stream fp32, GEMM operands
bf16 (plus the bf16 "lo" planes in fp32 mode).
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from . import _lib, ops
from .features import convnext_features as cnx
from .features import resnet_features as rsn


# Working-set target of one image chunk of a stage (bf16 mode); 0 = whole batch at once. Measured on B200 (B=64, cfg2):
# chunking for L2 residency LOSES (6.6 ms at 96 MB chunks vs 4.9 ms unchunked) - the smaller GEMMs and depthwise convs
# lose more to wave quantisation than the HBM round trips cost - so it is off unless CPN_L2_CHUNK_MB is set.
# CPN_FUSED_MLP=0 falls back to the two-GEMM MLP (fc1 + LN/GELU epilogue, fc2 + scale/residual epilogue)
FUSED_MLP = os.environ.get("CPN_FUSED_MLP", "1") != "0"
# CPN_FP32_LEGACY=1: fp32 mode as in round 1 (LayerNorm inside the depthwise kernel, direct-store GEMM epilogues); A/B only
LEGACY_FP32 = os.environ.get("CPN_FP32_LEGACY", "0") == "1"
L2_CHUNK_BYTES = float(os.environ.get("CPN_L2_CHUNK_MB", "0")) * 2 ** 20 or 1e18


class Emit:
    """Emission context: where buffers come from and whether launches are recorded."""

    def __init__(self, buffers, split: bool, plan: Optional[ops.Plan]):
        self.buffers = buffers
        self.split = split
        self.plan = plan

    def buf(self, name: str, shape, dtype) -> torch.Tensor:
        t = self.buffers.get(name, shape, dtype)
        if self.plan is not None:
            self.plan.keepalive.append(t)
        return t

    def pair(self, name: str, shape):
        """bf16 hi plane and, in fp32 mode, the lo plane."""
        hi = self.buf(name + "_hi", shape, torch.bfloat16)
        lo = self.buf(name + "_lo", shape, torch.bfloat16) if self.split else None
        return hi, lo


def _split(w: torch.Tensor, want_lo: bool):
    hi = w.to(torch.bfloat16).contiguous()
    lo = (w - hi.float()).to(torch.bfloat16).contiguous() if want_lo else None
    return hi, lo


class ConvNeXtProgram:
    """features = [stem, stage, down, stage, ...] as built by features/convnext_features.py (any truncation)."""

    def __init__(self, net: nn.Module):
        self.net = net
        self.steps: List[tuple] = []
        for layer in net.features:
            kids = list(layer.children())
            if len(kids) == 2 and isinstance(kids[0], nn.Conv2d) and kids[0].in_channels == 3:
                conv, ln = kids
                if conv.kernel_size != (4, 4) or conv.stride != (4, 4) or conv.out_channels != 96:
                    raise NotImplementedError("fused stem expects Conv2d(3, 96, kernel 4, stride 4)")
                self.steps.append(("stem", conv, ln))
            elif len(kids) > 0 and all(isinstance(k, cnx.CNBlock) for k in kids):
                for blk in kids:
                    self.steps.append(("block", blk))
            elif len(kids) == 2 and isinstance(kids[1], nn.Conv2d) and kids[1].kernel_size == (2, 2):
                self.steps.append(("down", kids[0], kids[1]))
            else:
                raise NotImplementedError(f"fused ConvNeXt path does not know feature layer {layer}")
        if not self.steps or self.steps[0][0] != "stem":
            raise NotImplementedError("fused ConvNeXt path needs the stem as first feature layer")
        self.packed: List[dict] = []
        self.out_channels = 96
        for s in self.steps:
            if s[0] == "down":
                self.out_channels = s[2].out_channels

    def pack(self, split: bool) -> None:
        """Kernel-side copies of the weights that need a different layout or dtype (everything else is read live)."""
        self.packed = []
        with torch.no_grad():
            for s in self.steps:
                d = {}
                if s[0] == "block":
                    blk = s[1]
                    dw, fc1, fc2 = blk.block[0], blk.block[3], blk.block[5]
                    c = dw.out_channels
                    d["w49"] = dw.weight.detach().reshape(c, 49).t().contiguous()
                    # LayerNorm folded into fc1 (csrc/gemm_tma_epi.cuh EPI_LN_GELU_BF16, csrc/gemm_split_epi.cuh):
                    #   fc1(LN(y)) = rstd * (y @ (W1*gamma)^T - mean * rowsum(W1*gamma)) + (b1 + W1 @ beta)
                    # with (W1*gamma) rounded to bf16 (bf16 mode) or kept as hi + lo planes (fp32 mode); the row sums are
                    # those of the values the GEMM multiplies with.
                    ln = blk.block[2]
                    w1 = fc1.weight.detach().float()
                    if c % 32 == 0 and not LEGACY_FP32:
                        d["w1g"], d["w1g_lo"] = _split(w1 * ln.weight.detach().float()[None, :], split)
                        wsum = d["w1g"].double() if not split else d["w1g"].double() + d["w1g_lo"].double()
                        d["w1g_sum"] = wsum.sum(1).float().contiguous()
                        d["b1f"] = (fc1.bias.detach().double() + w1.double() @ ln.bias.detach().double()).float().contiguous()
                    else:
                        d["w1"], d["w1_lo"] = _split(fc1.weight.detach(), split)
                    d["w2"], d["w2_lo"] = _split(fc2.weight.detach(), split)
                elif s[0] == "down":
                    conv = s[2]
                    w = conv.weight.detach().permute(0, 2, 3, 1).reshape(conv.out_channels, 4 * conv.in_channels)
                    d["w"], d["w_lo"] = _split(w, split)
                self.packed.append(d)

    def emit(self, em: Emit, x: torch.Tensor, B: int, H: int, W: int) -> Tuple[torch.Tensor, int, int, int]:
        """Returns (features [B*Ho*Wo, C] fp32 channels-last, Ho, Wo, C)."""
        plan = em.plan
        res = None
        C = 0
        n_res = 0
        steps = list(zip(self.steps, self.packed))
        i = 0
        while i < len(steps):
            s, pk = steps[i]
            if s[0] == "stem":
                _, conv, ln = s
                H, W = (H - 4) // 4 + 1, (W - 4) // 4 + 1
                C = conv.out_channels
                res = em.buf(f"res{n_res}", (B, H, W, C), torch.float32)
                n_res += 1
                ops.stem_ln(x, conv.weight.detach().view(C, 48), conv.bias.detach(), ln.weight.detach(),
                            ln.bias.detach(), res, eps=ln.eps, plan=plan)
                i += 1
            elif s[0] == "block":
                j = i
                while j < len(steps) and steps[j][0][0] == "block":
                    j += 1
                self._emit_stage(em, steps[i:j], res, B, H, W, C)
                i = j
            else:
                _, ln, conv = s
                stride = conv.stride[0]
                Ho, Wo = (H - 2) // stride + 1, (W - 2) // stride + 1
                M = B * Ho * Wo
                i_hi, i_lo = em.pair("im2col", (M, 4 * C))
                ops.ln_im2col2(res, ln.weight.detach(), ln.bias.detach(), i_hi, i_lo, stride=stride, eps=ln.eps,
                               plan=plan)
                Cn = conv.out_channels
                nxt = em.buf(f"res{n_res}", (B, Ho, Wo, Cn), torch.float32)
                n_res += 1
                ops.gemm_bf16(i_hi, pk["w"], M, Cn, 4 * C, a_lo=i_lo, bw_lo=pk["w_lo"], bias=conv.bias.detach(),
                              out_f32=nxt, plan=plan)
                res, H, W, C = nxt, Ho, Wo, Cn
                i += 1
        return res.view(B * H * W, C), H, W, C

    def _emit_stage(self, em: Emit, blocks, res: torch.Tensor, B: int, H: int, W: int, C: int) -> None:
        """The CNBlocks of one stage (same H, W, C), updating the fp32 residual stream `res` in place.

        bf16 mode runs the whole stage image chunk by image chunk (images are independent in the forward), with the
        chunk sized so that its working set - fp32 residual, bf16 conv output, bf16 hidden activations, 16 bytes per
        element - stays in the 126 MB L2 across the kernels and across the blocks of the stage, instead of
        round-tripping through HBM after every kernel."""
        plan = em.plan
        T = B * H * W
        a_hi, a_lo = em.pair("act", (T, C))
        folded = all("w1g" in pk for _, pk in blocks)
        fused = folded and not em.split and FUSED_MLP and C in (96, 192, 384)
        # the 4C-wide hidden plane exists in HBM only when the MLP runs as two GEMMs
        h_hi, h_lo = (None, None) if fused else em.pair("hid", (T, 4 * C))
        if not folded:
            for s, pk in blocks:
                blk = s[1]
                dw, ln, fc1, fc2 = blk.block[0], blk.block[2], blk.block[3], blk.block[5]
                ops.dwconv7_ln(res, pk["w49"], dw.bias.detach(), ln.weight.detach(), ln.bias.detach(), a_hi, a_lo,
                               eps=ln.eps, plan=plan)
                ops.gemm_bf16(a_hi, pk["w1"], T, 4 * C, C, a_lo=a_lo, bw_lo=pk["w1_lo"], bias=fc1.bias.detach(),
                              out_bf16=h_hi, out_bf16_lo=h_lo, act=_lib.ACT_GELU, plan=plan)
                ops.gemm_bf16(h_hi, pk["w2"], T, C, 4 * C, a_lo=h_lo, bw_lo=pk["w2_lo"], bias=fc2.bias.detach(),
                              scale=blk.layer_scale.detach().view(C), res_f32=res, out_f32=res, plan=plan)
            return
        stats = em.buf("ln_stats", (T, C // 32, 2), torch.float32)
        if em.split:
            # fp32 mode: hi/lo planes everywhere, three MMA passes per GEMM (csrc/gemm_split_epi.cuh)
            res4, res2 = res.view(B, H, W, C), res.view(T, C)
            for s, pk in blocks:
                blk = s[1]
                dw, ln, fc2 = blk.block[0], blk.block[2], blk.block[5]
                ops.dwconv7_stats_split(res4, pk["w49"], dw.bias.detach(), a_hi, a_lo, stats, plan=plan)
                ops.gemm_ln_gelu_split(a_hi, a_lo, pk["w1g"], pk["w1g_lo"], T, 4 * C, C, stats, C // 32, pk["w1g_sum"],
                                       pk["b1f"], h_hi, h_lo, eps=ln.eps, plan=plan)
                ops.gemm_bf16(h_hi, pk["w2"], T, C, 4 * C, a_lo=h_lo, bw_lo=pk["w2_lo"], bias=fc2.bias.detach(),
                              scale=blk.layer_scale.detach().view(C), res_f32=res2, out_f32=res2, plan=plan)
            return
        nb = max(1, min(B, int(L2_CHUNK_BYTES // (16 * H * W * C))))
        res4, res2 = res.view(B, H, W, C), res.view(T, C)
        for b0 in range(0, B, nb):
            b1 = min(B, b0 + nb)
            t0, t1 = b0 * H * W, b1 * H * W
            for s, pk in blocks:
                blk = s[1]
                dw, ln, fc2 = blk.block[0], blk.block[2], blk.block[5]
                ops.dwconv7_stats(res4[b0:b1], pk["w49"], dw.bias.detach(), a_hi[t0:t1], stats[t0:t1], plan=plan)
                if fused:
                    # fc1 -> GELU -> fc2 -> layer scale -> residual in one kernel, hidden activations stay on the SM
                    ops.mlp_fused(a_hi[t0:t1], pk["w1g"], pk["w2"], t1 - t0, C, stats[t0:t1], pk["w1g_sum"], pk["b1f"],
                                  fc2.bias.detach(), blk.layer_scale.detach().view(C), res2[t0:t1], eps=ln.eps,
                                  plan=plan)
                    continue
                ops.gemm_ln_gelu(a_hi[t0:t1], pk["w1g"], t1 - t0, 4 * C, C, stats[t0:t1], C // 32, pk["w1g_sum"],
                                 pk["b1f"], h_hi[t0:t1], eps=ln.eps, plan=plan)
                ops.gemm_bf16(h_hi[t0:t1], pk["w2"], t1 - t0, C, 4 * C, bias=fc2.bias.detach(),
                              scale=blk.layer_scale.detach().view(C), res_f32=res2[t0:t1], out_f32=res2[t0:t1],
                              plan=plan)
