"""Host-side engine of the fused forward: decides when the sm_100a path runs, keeps per-(module, device) packed
weights, workspaces and recorded launch plans, and chains the C-ABI kernels for

    backbone (ConvNeXt-tiny variants / ResNet)  ->  add-on 1x1 conv  ->  softmax | hard Gumbel-softmax
    ->  spatial max | count pooling  ->  threshold | round+clamp+encoding  ->  non-negative classifier

i.e. PIPNet.forward (reference pipnet/pipnet.py:31-41) and CountPIPNet.forward (pipnet/count_pipnet.py:70-110).

All arithmetic happens in libcpnb200.so; torch is used for device memory, streams, the Gumbel noise draw (so the RNG
stream is the one `F.gumbel_softmax` would consume) and one-time weight repacking. There is no CPU path here: a
non-CUDA tensor in inference mode raises.
"""
from __future__ import annotations

import os
import threading
import weakref
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from . import _lib, ops

# "fp32" (the drop-in default): bf16 hi/lo split operands, three MMA passes per GEMM -- the mode that meets the north
#         star's exact rules (predicted class, arg-max prototype and counts equal to the reference's; pooled / logits
#         within 1e-4) on every parity case.
# "bf16" (opt-in: `model.cpn_precision = "bf16"` or CPN_PRECISION=bf16): bf16 tensor-core operands, fp32 accumulate and
#         residual stream; about twice as fast, pooled / logits within 2e-2, and a count or a thresholded score that
#         sits within that distance of its rounding boundary / the 0.1 threshold may land on the other side.
DEFAULT_PRECISION = os.environ.get("CPN_PRECISION", "fp32")
USE_PLANS = os.environ.get("CPN_NO_PLAN", "0") != "1"
# Replay a recorded plan as one CUDA graph from its second run on (small batches are launch-bound otherwise).
USE_GRAPHS = os.environ.get("CPN_GRAPHS", "1") != "0"
# Under autograd on a CUDA tensor, run softmax + pooling of the head (forward and backward) in the training kernels.
TRAIN_KERNELS = os.environ.get("CPN_TRAIN_KERNELS", "1") != "0"
MAX_PLANS = 8  # recorded plans kept per (module, device), least recently used dropped first

_LOCK = threading.RLock()
_STATES: "weakref.WeakKeyDictionary[nn.Module, Dict[int, object]]" = weakref.WeakKeyDictionary()


def precision_of(module: nn.Module) -> str:
    p = getattr(module, "cpn_precision", None) or DEFAULT_PRECISION
    if p not in ("bf16", "fp32"):
        raise ValueError(f"cpn_precision must be 'bf16' or 'fp32', got {p!r}")
    return p


def use_fused(module: nn.Module, x: torch.Tensor) -> bool:
    """True when the fused kernels must run: eval mode and no autograd graph is being built. With autograd enabled
    (training, attribution) the composed modules run so gradients exist. A CPU tensor in inference mode is an error:
    this package has no CPU implementation of the hot path."""
    if module.training:
        return False
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in module.parameters())):
        return False
    if not x.is_cuda:
        raise _lib.CpnError(
            "count_pipnet_b200 runs its inference forward on a B200 only (input is on %s); there is no CPU fallback. "
            "Move the module and input to CUDA, or enable autograd to use the composed PyTorch modules." % x.device)
    return True


def use_train_head(module: nn.Module, features: torch.Tensor, want_max: bool) -> bool:
    """True when the autograd (composed) forward should hand its head to the training kernels of pipnet/train.py:
    a CUDA fp32 map, the add-on ending in nn.Softmax(dim=1) (not the Gumbel variant, whose noise draw and
    straight-through estimator stay torch ops), and for PIP-Net the stock AdaptiveMaxPool2d((1, 1)) + Flatten pool.
    CPN_TRAIN_KERNELS=0 keeps the torch modules (A/B)."""
    if not TRAIN_KERNELS or not features.is_cuda or features.dtype != torch.float32 or features.dim() != 4:
        return False
    add_on = getattr(module, "_add_on", None)
    if not isinstance(add_on, nn.Sequential) or len(add_on) == 0:
        return False
    last = add_on[len(add_on) - 1]
    if type(last) is not nn.Softmax or last.dim != 1:
        return False
    if want_max:
        pool = getattr(module, "_pool", None)
        if not isinstance(pool, nn.Sequential) or len(pool) != 2:
            return False
        if type(pool[0]) is not nn.AdaptiveMaxPool2d or pool[0].output_size not in (1, (1, 1)) \
                or pool[0].return_indices or type(pool[1]) is not nn.Flatten:
            return False
    return True


class Buffers:
    """Named device buffers that persist across forwards (recorded plans hold their addresses)."""

    def __init__(self, device):
        self.device = device
        self.t: Dict[str, torch.Tensor] = {}
        self.regrown = False  # a buffer that recorded plans may point at has been replaced by a larger one

    def get(self, name: str, shape, dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        cur = self.t.get(name)
        if cur is None or cur.dtype != dtype or cur.numel() < n:
            if cur is not None:
                self.regrown = True
            cur = torch.empty(max(n, 1), dtype=dtype, device=self.device)
            self.t[name] = cur
        return cur[:n].view(*shape)


def split_bf16(w: torch.Tensor, want_lo: bool) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """fp32 -> (bf16 hi, bf16 lo) with w ~= hi + lo."""
    hi = w.to(torch.bfloat16).contiguous()
    lo = (w - hi.float()).to(torch.bfloat16).contiguous() if want_lo else None
    return hi, lo


def param_signature(module: nn.Module):
    """(data_ptr, version) of every parameter and buffer: changes when weights are replaced or mutated in place."""
    sig = []
    for t in list(module.parameters()) + list(module.buffers()):
        sig.append((t.data_ptr(), t._version))
    return tuple(sig)


class State:
    """Everything cached for one (module, device)."""

    def __init__(self, device):
        self.device = device
        self.buffers = Buffers(device)
        self.program = None
        self.packed_sig = None
        self.packed_precision = None
        self.plans: "OrderedDict[tuple, object]" = OrderedDict()

    def remember(self, key, rec) -> None:
        """Keep `rec` as the most recent plan. Plans recorded against buffers that have since been replaced by larger
        ones pin the old allocations (Plan.keepalive): when that happened during this recording, drop them; otherwise
        keep at most MAX_PLANS."""
        if self.buffers.regrown:
            self.plans.clear()
            self.buffers.regrown = False
        self.plans[key] = rec
        self.plans.move_to_end(key)
        while len(self.plans) > MAX_PLANS:
            self.plans.popitem(last=False)


def state_for(module: nn.Module, device: torch.device) -> State:
    if getattr(module, "_is_replica", False):
        # nn.DataParallel over several devices (reference main.py:118) builds throw-away replicas with fresh weight
        # copies on every forward: packed weights, TMA descriptors and plans could never be reused. One device under
        # DataParallel calls the module itself and is fine.
        raise _lib.CpnError(
            "the fused B200 forward keeps per-(module, device) packed weights and launch plans and cannot run inside "
            "multi-device nn.DataParallel replicas; run one process per GPU (torchrun) and shard the batch with "
            "count_pipnet_b200.sharding, or pass a single device_id")
    with _LOCK:
        per_dev = _STATES.get(module)
        if per_dev is None:
            per_dev = {}
            _STATES[module] = per_dev
        idx = device.index if device.index is not None else torch.cuda.current_device()
        st = per_dev.get(idx)
        if st is None:
            ops.check_device()
            st = State(device)
            per_dev[idx] = st
        return st


from .forward import backbone_forward_nchw, network_forward  # noqa: E402,F401  (after the definitions it uses)
