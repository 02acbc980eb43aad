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
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from . import _lib, ops

# "bf16": bf16 tensor-core operands, fp32 accumulate/residual (2e-2 parity bar).
# "fp32": bf16 hi/lo split operands, three MMA passes (1e-4 parity bar).
DEFAULT_PRECISION = os.environ.get("CPN_PRECISION", "bf16")
USE_PLANS = os.environ.get("CPN_NO_PLAN", "0") != "1"

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


class Buffers:
    """Named device buffers that persist across forwards (recorded plans hold their addresses)."""

    def __init__(self, device):
        self.device = device
        self.t: Dict[str, torch.Tensor] = {}

    def get(self, name: str, shape, dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        cur = self.t.get(name)
        if cur is None or cur.dtype != dtype or cur.numel() < n:
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
        self.plans: Dict[tuple, ops.Plan] = {}


def state_for(module: nn.Module, device: torch.device) -> State:
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
