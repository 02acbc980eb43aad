"""Count-PIP-Net building blocks with the reference's names and semantics
(pipnet/count_pipnet_utils.py of TarasKutsyk/Count_PIPNet; file:line cited per symbol).

These are the composed (autograd-capable) forms. The inference forward of a whole network does not call them: the
fused sm_100a head kernels (csrc/head_kernels.cuh) read their parameters directly. They stay real, individually
callable modules because reference callers use them on their own (`net._intermediate(counts)`,
`prototype_to_classifier_input_weights`).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


class GumbelSoftmax(nn.Module):
    """Soft Gumbel-softmax sample in training, hard (one-hot, straight-through form) in eval
    (count_pipnet_utils.py:7-38). `tau` is a plain attribute, not saved in checkpoints."""

    def __init__(self, dim: int = 1, tau: float = 1.0):
        super().__init__()
        self.dim = dim
        self.tau = tau

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return F.gumbel_softmax(x, tau=self.tau, hard=not self.training, dim=self.dim)


class STE_Round(torch.autograd.Function):
    """round() forward (half to even), identity backward (count_pipnet_utils.py:41-55)."""

    @staticmethod
    def forward(ctx, x):
        return x.round()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output


class ClampSTE(torch.autograd.Function):
    """clamp forward; backward is identity or gated by the in-range mask (count_pipnet_utils.py:58-84)."""

    @staticmethod
    def forward(ctx, input, min_val, max_val, is_backward_identity: bool):
        ctx.save_for_backward(input)
        ctx.bounds = (min_val, max_val)
        ctx.is_backward_identity = is_backward_identity
        return input.clamp(min_val, max_val)

    @staticmethod
    def backward(ctx, grad_output):
        if not ctx.is_backward_identity:
            (inp,) = ctx.saved_tensors
            lo, hi = ctx.bounds
            grad_output = grad_output * ((inp >= lo) & (inp <= hi)).float()
        return grad_output, None, None, None


def create_modified_encoding(x: torch.Tensor, max_count: int) -> torch.Tensor:
    """Counts -> [B, P, max_count] fp32 codes: 0 -> zeros, c>=1 -> unit vector at min(c, max_count) - 1
    (count_pipnet_utils.py:141-185: mask x > 0.1, index clamp(long(x) - 1, 0, max_count - 1))."""
    b, p = x.shape
    idx = torch.clamp(x.long() - 1, 0, max_count - 1)
    codes = F.one_hot(idx, max_count).to(torch.float32)
    return codes * (x > 0.1).unsqueeze(-1).to(torch.float32)


class ModifiedSTEFunction(torch.autograd.Function):
    """One-hot count encoding with a straight-through backward (count_pipnet_utils.py:188-321).

    Forward: round, then `create_modified_encoding` (:195-217). Backward reproduces the reference's *effective*
    rule: for a non-zero count the gradient follows the most negative slot - positive (decrease the count) if that
    slot is below the active one, negative (increase) if above, zero if it is the active slot; `current_grad` /
    `max_grad` change the magnitude when every slot gradient is positive; `respect_active_grad` zeroes entries whose
    active-slot gradient is already negative. Zero counts receive no gradient: the reference's assignments through a
    chained boolean index (:313-319, and :262-263 in the `max_grad` branch) write to temporaries, and that behaviour is
    kept so training dynamics match.
    """

    @staticmethod
    def forward(ctx, counts, max_count: int, respect_active_grad: bool, positive_grad_strategy: Optional[str] = None):
        rounded = counts.round()
        ctx.save_for_backward(rounded)
        ctx.cfg = (max_count, respect_active_grad, positive_grad_strategy)
        return create_modified_encoding(rounded, max_count)

    @staticmethod
    def backward(ctx, grad_output):
        (rounded,) = ctx.saved_tensors
        max_count, respect_active, strategy = ctx.cfg
        if grad_output.shape != (*rounded.shape, max_count):
            raise ValueError(f"Unexpected grad_output shape. Expected [{rounded.shape[0]}, {rounded.shape[1]}, "
                             f"{max_count}], got {grad_output.shape}")
        active = torch.clamp(rounded.long() - 1, 0, max_count - 1)
        nonzero = ~(rounded < 0.1)
        min_val, min_idx = grad_output.min(dim=2)
        grad_at_active = grad_output.gather(2, active.unsqueeze(-1)).squeeze(-1)
        all_pos = (min_val > 0) & nonzero
        zeros = torch.zeros_like(min_val)
        if strategy == "max_grad" and bool(all_pos.any()):
            grad = torch.where(all_pos, grad_output.max(dim=2)[0], zeros)
        else:
            mag = min_val.abs()
            if strategy == "current_grad" and bool(all_pos.any()):
                mag = torch.where(all_pos, grad_at_active, mag)
            grad = torch.where(min_idx < active, mag, torch.where(min_idx > active, -mag, zeros))
        if respect_active:
            grad = torch.where(grad_at_active < 0, zeros, grad)
        grad = torch.where(nonzero, grad, zeros)
        return grad, None, None, None


class OneHotEncoder(nn.Module):
    """Counts [B,P] -> flattened codes [B, P*num_bins] (count_pipnet_utils.py:86-139)."""

    def __init__(self, num_bins: int = 4, use_ste: bool = False, respect_active_grad: bool = False,
                 num_prototypes: int = None, device: Optional[torch.device] = None,
                 positive_grad_strategy: Optional[str] = None):
        super().__init__()
        self.num_bins = num_bins
        self.num_prototypes = num_prototypes
        self.device = device
        self.use_ste = use_ste
        self.respect_active_grad = respect_active_grad
        self.positive_grad_strategy = positive_grad_strategy

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.use_ste:
            codes = ModifiedSTEFunction.apply(x, self.num_bins, self.respect_active_grad, self.positive_grad_strategy)
        else:
            codes = create_modified_encoding(x, self.num_bins)
        return codes.view(codes.size(0), -1)

    def prototype_to_classifier_input_weights(self, prototype_idx):
        v = torch.zeros(self.num_prototypes * self.num_bins, device=self.device)
        v[prototype_idx * self.num_bins:(prototype_idx + 1) * self.num_bins] = 1.0
        return v


class BilinearIntermediate(nn.Module):
    """e = embed(x); W(e) * V(e), all bias-free (count_pipnet_utils.py:323-385). embed starts as the count ramp
    embed[p*M + c, p] = c + 1 (:349-358)."""

    def __init__(self, num_prototypes, max_count, expanded_dim=None, custom_init=False):
        super().__init__()
        self.num_prototypes = num_prototypes
        self.max_count = max_count
        self.expanded_dim = num_prototypes * max_count if expanded_dim is None else expanded_dim
        self.embed = nn.Linear(num_prototypes, self.expanded_dim, bias=False)
        self.W = nn.Linear(self.expanded_dim, self.expanded_dim, bias=False)
        self.V = nn.Linear(self.expanded_dim, self.expanded_dim, bias=False)
        with torch.no_grad():
            self.embed.weight.copy_(_count_ramp(num_prototypes, max_count, self.expanded_dim))
            if custom_init:
                eye = 0.1 * torch.eye(self.expanded_dim)
                self.W.weight.normal_(0.0, 0.1).add_(eye)
                self.V.weight.normal_(0.0, 0.1).add_(eye)

    def forward(self, x):
        e = self.embed(x)
        return self.W(e) * self.V(e)


def _count_ramp(num_prototypes: int, max_count: int, rows: int) -> torch.Tensor:
    """[rows, P] matrix with entry (p*M + c, p) = c + 1 and zeros elsewhere."""
    w = torch.zeros(rows, num_prototypes)
    p = torch.arange(num_prototypes).repeat_interleave(max_count)
    c = torch.arange(max_count).repeat(num_prototypes)
    w[p * max_count + c, p] = (c + 1).to(torch.float32)
    return w


class LinearFull(nn.Module):
    """Dense bias-free map P -> P*M (count_pipnet_utils.py:387-444); init: count ramp on the own prototype,
    0.1*(c+1)/P on the others (:410-427)."""

    def __init__(self, num_prototypes, max_count, expanded_dim=None):
        super().__init__()
        self.num_prototypes = num_prototypes
        self.max_count = max_count
        self.expanded_dim = num_prototypes * max_count if expanded_dim is None else expanded_dim
        self.linear = nn.Linear(num_prototypes, self.expanded_dim, bias=False)
        with torch.no_grad():
            ramp = _count_ramp(num_prototypes, max_count, self.expanded_dim)
            c1 = (torch.arange(max_count).repeat(num_prototypes) + 1).to(torch.float32)
            side = torch.zeros_like(ramp)
            side[:num_prototypes * max_count] = (0.1 * c1 / num_prototypes).unsqueeze(1)
            self.linear.weight.copy_(torch.where(ramp != 0, ramp, side))

    def forward(self, x):
        return self.linear(x)

    def prototype_to_classifier_input_weights(self, prototype_idx):
        return self.linear.weight[:, prototype_idx]


class IdentityIntermediate(nn.Module):
    """Identity with the intermediate-layer interface (count_pipnet_utils.py:446-469)."""

    def __init__(self, num_prototypes, device):
        super().__init__()
        self.identity = nn.Identity()
        self.num_prototypes = num_prototypes
        self.device = device

    def forward(self, x):
        return self.identity(x)

    def prototype_to_classifier_input_weights(self, prototype_idx):
        return torch.eye(self.num_prototypes, device=self.device)[prototype_idx]


class LinearIntermediate(nn.Module):
    """out[b, p*E + i] = w[i] * x[b, p] with a shared [E,1] weight, init (i+1)/M (count_pipnet_utils.py:471-538)."""

    def __init__(self, num_prototypes, max_count, expansion_factor=None):
        super().__init__()
        self.num_prototypes = num_prototypes
        self.max_count = max_count
        self.expansion_factor = max_count if expansion_factor is None else expansion_factor
        self.linear = nn.Linear(1, self.expansion_factor, bias=False)
        with torch.no_grad():
            ramp = (torch.arange(self.expansion_factor, dtype=torch.float32) + 1) / self.max_count
            self.linear.weight.copy_(ramp.unsqueeze(1))

    def forward(self, x):
        b = x.shape[0]
        return self.linear(x.view(b * self.num_prototypes, 1)).view(b, self.num_prototypes * self.expansion_factor)

    def prototype_to_classifier_input_weights(self, prototype_idx):
        e = self.expansion_factor
        v = torch.zeros(self.num_prototypes * e, device=self.linear.weight.device, dtype=self.linear.weight.dtype)
        v[prototype_idx * e:(prototype_idx + 1) * e] = self.linear.weight[:, 0]
        return v
