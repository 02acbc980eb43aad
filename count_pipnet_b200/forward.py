"""The fused forward itself: `network_forward(model, xs, inference)` is what PIPNet.forward / CountPIPNet.forward
call in inference mode (reference pipnet/pipnet.py:31-41, pipnet/count_pipnet.py:70-110), and
`backbone_forward_nchw(net, x)` what `net._net(x)` calls when used on its own (reference callers:
util/interpret_idg.py:171-172, notebooks/main_interp.py:167-168)."""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.nn as nn

from . import _lib, ops
from . import engine as _e
from .features import resnet_features as rsn
from .pipnet import count_pipnet_utils as cpu_
from .programs import ConvNeXtProgram, Emit
from .programs_resnet import ResNetProgram

NUM_SMS = 148
# CTAs of the softmax/pool kernel per SM (each takes a chunk of one image's tokens); env for measurement only
POOL_CHUNK_FACTOR = float(os.environ.get("CPN_POOL_CHUNKS", "3"))
# CPN_FUSED_ADDON=0: add-on conv and softmax/pool as two kernels with the logits in HBM (A/B only)
FUSED_ADDON = os.environ.get("CPN_FUSED_ADDON", "1") != "0"


def _make_program(backbone: nn.Module):
    if isinstance(backbone, rsn.ResNet_features):
        return ResNetProgram(backbone)
    if hasattr(backbone, "features"):
        return ConvNeXtProgram(backbone)
    raise NotImplementedError(f"no fused program for backbone {type(backbone).__name__}")


def _pack_sig(backbone: nn.Module):
    return tuple((t.data_ptr(), t._version) for t in list(backbone.parameters()) + list(backbone.buffers()))


def _ptr_sig(model: nn.Module):
    return tuple(t.data_ptr() for t in list(model.parameters()) + list(model.buffers()))


def _ensure_program(st, backbone: nn.Module, split: bool):
    if st.program is None:
        st.program = _make_program(backbone)
    sig = (_pack_sig(backbone), split)
    if st.packed_sig != sig:
        st.program.pack(split)
        st.packed_sig = sig
        st.plans.clear()  # recorded launches point at the old packed weights
    return st.program


def _as_input(xs: torch.Tensor) -> torch.Tensor:
    if xs.dim() != 4 or xs.shape[1] != 3:
        raise _lib.CpnError(f"expected an image batch [B,3,H,W], got {tuple(xs.shape)}")
    if xs.dtype != torch.float32:
        xs = xs.float()
    return xs.contiguous()


# ------------------------------------------------------------------------------------------------ head
def _head_spec(model: nn.Module):
    """Reads the add-on / pooling / intermediate / classifier structure off the module tree."""
    mods = list(model._add_on.children())
    conv = mods[0] if mods and isinstance(mods[0], nn.Conv2d) else None
    act = mods[-1]
    if conv is not None and (conv.kernel_size != (1, 1) or conv.stride != (1, 1)):
        raise NotImplementedError("fused head expects a 1x1 add-on conv")
    if isinstance(act, cpu_.GumbelSoftmax):
        gumbel, tau = True, float(act.tau)
    elif isinstance(act, nn.Softmax):
        gumbel, tau = False, 1.0
    else:
        raise NotImplementedError(f"fused head does not know activation {type(act).__name__}")
    is_count = hasattr(model, "_max_count")
    enc, w0, w1, w2 = _lib.ENC_NONE, None, None, None
    if is_count:
        inter = model._intermediate
        if isinstance(inter, cpu_.OneHotEncoder):
            enc = _lib.ENC_ONEHOT
        elif isinstance(inter, cpu_.LinearIntermediate):
            enc, w0 = _lib.ENC_LINEAR, inter.linear.weight
        elif isinstance(inter, cpu_.LinearFull):
            enc, w0 = _lib.ENC_LINEAR_FULL, inter.linear.weight
        elif isinstance(inter, cpu_.BilinearIntermediate):
            enc, w0, w1, w2 = _lib.ENC_BILINEAR, inter.embed.weight, inter.W.weight, inter.V.weight
        elif isinstance(inter, cpu_.IdentityIntermediate):
            enc = _lib.ENC_IDENTITY
        else:
            raise NotImplementedError(f"fused head does not know intermediate {type(inter).__name__}")
    return conv, gumbel, tau, is_count, enc, w0, w1, w2


def _emit_head(em: Emit, model: nn.Module, feat: torch.Tensor, B: int, Hf: int, Wf: int, C: int, inference: bool,
               want_map: bool):
    conv, gumbel, tau, is_count, enc, w0, w1, w2 = _head_spec(model)
    plan = em.plan
    T, HW = B * Hf * Wf, Hf * Wf
    fuse_addon = conv is not None and conv.out_channels in (16, 32) and C % 64 == 0 and FUSED_ADDON
    if conv is not None:
        P = conv.out_channels
        if not fuse_addon:
            z = em.buf("z", (T, P), torch.float32)
            ops.addon_conv1x1(feat, conv.weight.detach().view(P, C), None if conv.bias is None else conv.bias.detach(), z,
                              T, C, P, plan=plan)
    else:
        P, z = C, feat
    noise = em.buf("gumbel", (B, P, Hf, Wf), torch.float32) if gumbel else None
    chunks = max(1, min(max(HW // 8, 1), math.ceil(POOL_CHUNK_FACTOR * NUM_SMS / B)))
    if fuse_addon:
        # the fused kernel's CTA = 4 warps x 32 tokens: one 128-token chunk per CTA unless the batch is small
        chunks = max(chunks, math.ceil(HW / 128))
    partial = em.buf("partial", (B, chunks, P), torch.float32)
    pmap = em.buf("map", (B, P, Hf, Wf), torch.float32) if want_map else None
    mode = _lib.POOL_SUM if is_count else _lib.POOL_MAX
    # per (image, prototype) location of the spatial maximum, on request (model.track_prototype_argmax): what the
    # reference's visualisation search reads off the full map (util/vis_pipnet.py:209-274) without writing the map
    want_arg = bool(getattr(model, "track_prototype_argmax", False)) and not is_count
    parg = em.buf("partial_arg", (B, chunks, P), torch.int32) if want_arg else None
    if fuse_addon:
        # add-on conv + softmax + pooling in one kernel: the patch logits never reach HBM
        ops.addon_softmax_pool(feat, conv.weight.detach().view(P, C), None if conv.bias is None else conv.bias.detach(),
                               partial, B, HW, C, P, chunks, gumbel=noise, tau=tau, hard=gumbel, mode=mode, map_out=pmap,
                               partial_arg=parg, plan=plan)
    else:
        ops.softmax_pool(z, partial, B, HW, P, chunks, gumbel=noise, tau=tau, hard=gumbel, mode=mode, map_out=pmap,
                         partial_arg=parg, plan=plan)
    argmax = None
    if want_arg:
        argmax = em.buf("argmax", (B, P), torch.int32)
        ops.pool_argmax_reduce(partial, parg, chunks, argmax, B, P, plan=plan)
    cls = model._classification
    K, D = cls.weight.shape
    pooled = em.buf("pooled", (B, P), torch.float32)
    logits = em.buf("logits", (B, K), torch.float32)
    det = lambda t: None if t is None else t.detach()
    ops.head_tail(partial, chunks, pooled, logits, cls.weight.detach(), det(cls.bias), B, P, D, K, mode=mode,
                  inference=inference, use_ste=bool(getattr(model, "_use_ste", False)),
                  max_count=int(getattr(model, "_max_count", 1)), encoding=enc, enc_w0=det(w0), enc_w1=det(w1),
                  enc_w2=det(w2), plan=plan)
    return {"noise": noise, "map": pmap, "pooled": pooled, "logits": logits, "argmax": argmax, "grid": (Hf, Wf)}


class _Recorded:
    def __init__(self, sig, plan, x_in, outs):
        self.sig, self.plan, self.x_in, self.outs = sig, plan, x_in, outs
        self.runs = 0


def _fill_gumbel(noise: Optional[torch.Tensor], model: nn.Module) -> None:
    """Gumbel noise for this forward. Default: the draw F.gumbel_softmax makes (-log(Exp(1)) on a contiguous
    [B,P,H,W] tensor from torch's generator for the device). A caller can pin it by setting
    `model.cpn_gumbel_noise` to a [B,P,H,W] tensor (parity tests inject the reference run's noise this way)."""
    if noise is None:
        return
    fixed = getattr(model, "cpn_gumbel_noise", None)
    if fixed is not None:
        if tuple(fixed.shape) != tuple(noise.shape):
            raise _lib.CpnError(f"cpn_gumbel_noise has shape {tuple(fixed.shape)}, forward needs {tuple(noise.shape)}")
        noise.copy_(fixed)
    else:
        noise.exponential_().log_().neg_()


def network_forward(model: nn.Module, xs: torch.Tensor, inference: bool = False):
    x = _as_input(xs)
    B, _, H, W = x.shape
    st = _e.state_for(model, x.device)
    split = _e.precision_of(model) == "fp32"
    want_map = bool(getattr(model, "materialize_proto_map", True))
    with _e._LOCK, torch.cuda.device(x.device):
        prog = _ensure_program(st, model._net, split)
        if _e.USE_PLANS:
            tau = _head_spec(model)[2]
            key = (B, H, W, bool(inference), want_map, split, tau, bool(getattr(model, "track_prototype_argmax", False)))
            sig = _ptr_sig(model)
            rec = st.plans.get(key)
            if rec is None or rec.sig != sig:
                plan = ops.Plan()
                em = Emit(st.buffers, split, plan)
                x_in = em.buf("x_in", x.shape, torch.float32)
                feat, Hf, Wf, C = prog.emit(em, x_in, B, H, W)
                outs = _emit_head(em, model, feat, B, Hf, Wf, C, bool(inference), want_map)
                rec = _Recorded(sig, plan, x_in, outs)
                st.remember(key, rec)
            else:
                st.plans.move_to_end(key)
            rec.x_in.copy_(x)
            _fill_gumbel(rec.outs["noise"], model)
            if _e.USE_GRAPHS and rec.runs == 1 and not rec.plan.has_graph:
                # second forward of this shape: the kernels are loaded and their attributes set, so the launch
                # sequence can be captured; every later forward is one cudaGraphLaunch. A plan that cannot be captured
                # keeps replaying its launch list (same kernels, same results); the reason is kept for inspection.
                try:
                    rec.plan.capture()
                except _lib.CpnError as exc:
                    model.cpn_graph_error = str(exc)
            rec.plan.run()
            rec.runs += 1
            outs = rec.outs
        else:
            em = Emit(st.buffers, split, None)
            feat, Hf, Wf, C = prog.emit(em, x, B, H, W)
            # the noise buffer must be filled before the head kernels read it
            conv, gumbel = _head_spec(model)[0:2]
            if gumbel:
                P = conv.out_channels if conv is not None else C
                _fill_gumbel(st.buffers.get("gumbel", (B, P, Hf, Wf), torch.float32), model)
            outs = _emit_head(em, model, feat, B, Hf, Wf, C, bool(inference), want_map)
        proto = outs["map"].clone() if outs["map"] is not None else None
        if outs.get("argmax") is not None:
            # [B, P] int32 token index h * Wf + w of each prototype's maximum, and the patch grid it refers to
            model.cpn_prototype_argmax = outs["argmax"].clone()
            model.cpn_patch_grid = outs["grid"]
        return proto, outs["pooled"].clone(), outs["logits"].clone()


def backbone_forward_nchw(net: nn.Module, xs: torch.Tensor) -> torch.Tensor:
    x = _as_input(xs)
    B, _, H, W = x.shape
    st = _e.state_for(net, x.device)
    split = _e.precision_of(net) == "fp32"
    with _e._LOCK, torch.cuda.device(x.device):
        prog = _ensure_program(st, net, split)
        em = Emit(st.buffers, split, None)
        feat, Hf, Wf, C = prog.emit(em, x, B, H, W)
        out = torch.empty(B, C, Hf, Wf, device=x.device, dtype=torch.float32)
        ops.nhwc_to_nchw(feat, out, B, Hf * Wf, C)
        return out
