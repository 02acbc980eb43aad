"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's training losses and of the
head's autograd forward, as plain functions over tensors.

  pipnet_losses(...)      pipnet/train.py:165-229 calculate_loss (terms and weighting) + :253-265 align_loss,
                          :233-235 accuracy; criterion = nn.NLLLoss(reduction='mean') (main.py:208)
  loss_gradients(...)     d loss / d (proto_features, pooled, out, normalization_multiplier) by autograd over the
                          restatement (what `loss.backward()`, train.py:103, hands to the head)
  head_forward(...)       pipnet/pipnet.py:32-33 (nn.Softmax(dim=1), AdaptiveMaxPool2d((1,1)) + Flatten: :96, :105-108)
                          or pipnet/count_pipnet.py:88 (sum over the spatial axes)
  head_gradients(...)     autograd of head_forward for given output gradients

Parity pinning: tests/golden/losses_golden.npz holds loss values, accuracies, components and autograd gradients of the
reference's OWN calculate_loss (tests/golden/make_golden_losses.py imports pipnet.train from /root/reference and runs it
on the seeded cases of tests/golden/golden_loss_cases.py); tests/test_losses_oracle.py checks this restatement against it.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch


def align_term(e1: torch.Tensor, e2: torch.Tensor, eps: float = 1e-12) -> torch.Tensor:
    """train.py:253-265 with targets detached: -mean_n log(<e1_n, e2_n> + eps)."""
    return -torch.log((e1 * e2.detach()).sum(dim=1) + eps).mean()


def pipnet_losses(proto_features: torch.Tensor, pooled: torch.Tensor, out: Optional[torch.Tensor],
                  ys1: Optional[torch.Tensor], align_w: float, tanh_w: float, class_w: float,
                  norm_mult: Optional[torch.Tensor], pretrain: bool, finetune: bool, is_count: bool = False,
                  eps: float = 1e-10, enforce_weight_sparsity: bool = True, tanh_coeff: float = 1.0
                  ) -> Tuple[torch.Tensor, float, Dict[str, float]]:
    """-> (loss, acc, components) as train.py:165-243 returns them."""
    n2, P = proto_features.shape[0], proto_features.shape[1]
    B = n2 // 2
    pf1, pf2 = proto_features[:B], proto_features[B:]          # :168 chunk(2)
    e1 = pf1.reshape(B, P, -1).transpose(1, 2).reshape(-1, P)   # :170-171: one row per patch
    e2 = pf2.reshape(B, P, -1).transpose(1, 2).reshape(-1, P)
    align = (align_term(e1, e2) + align_term(e2, e1)) / 2.      # :173
    c = tanh_coeff if is_count else 1.0                          # :175-189
    p1, p2 = pooled[:B] * c, pooled[B:] * c
    tanh = -(torch.log(torch.tanh(p1.sum(dim=0)) + eps).mean() + torch.log(torch.tanh(p2.sum(dim=0)) + eps).mean()) / 2.
    comps = {'align': align.item(), 'align_weighted': align.item() * align_w, 'tanh': tanh.item(),
             'tanh_weighted': tanh.item() * tanh_w, 'class': 0.0, 'class_weighted': 0.0}
    loss = None
    if not finetune:
        loss = align_w * align + tanh_w * tanh                   # :212-214
    acc = 0.
    if not pretrain:
        ys = torch.cat([ys1, ys1])                               # :166
        v = torch.log1p(out ** norm_mult) if enforce_weight_sparsity else out   # :217-220
        logp = torch.log_softmax(v, dim=1)
        cl = -logp[torch.arange(n2, device=out.device), ys].mean()                  # NLLLoss(reduction='mean')
        comps['class'] = cl.item()
        comps['class_weighted'] = cl.item() * class_w
        loss = class_w * cl if finetune else loss + class_w * cl   # :224-229
        acc = (out.argmax(dim=1) == ys).sum().item() / float(n2)  # :233-235
    return loss, acc, comps


def loss_gradients(proto_features, pooled, out, ys1, align_w, tanh_w, class_w, norm_mult, pretrain, finetune, **kw):
    """-> (loss value, acc, components, {name: gradient or None}) with fresh leaves."""
    leaves = {'proto_features': proto_features.detach().clone().requires_grad_(True),
              'pooled': pooled.detach().clone().requires_grad_(True)}
    if not pretrain:
        leaves['out'] = out.detach().clone().requires_grad_(True)
        leaves['norm_mult'] = norm_mult.detach().clone().requires_grad_(True)
    loss, acc, comps = pipnet_losses(leaves['proto_features'], leaves['pooled'], leaves.get('out'), ys1, align_w, tanh_w,
                                     class_w, leaves.get('norm_mult'), pretrain, finetune, **kw)
    loss.backward()
    return loss.item(), acc, comps, {k: (None if v.grad is None else v.grad.clone()) for k, v in leaves.items()}


def head_forward(z: torch.Tensor, mode: str) -> Tuple[torch.Tensor, torch.Tensor]:
    """z [N, P, H, W] -> (softmax over dim 1, max or sum over the spatial axes)."""
    pf = torch.softmax(z, dim=1)
    if mode == "max":
        pooled = torch.nn.functional.adaptive_max_pool2d(pf, (1, 1)).flatten(1)
    else:
        pooled = pf.sum(dim=(2, 3))
    return pf, pooled


def head_gradients(z: torch.Tensor, mode: str, g_pf: Optional[torch.Tensor], g_pooled: Optional[torch.Tensor]):
    """Gradient w.r.t. z of <g_pf, pf> + <g_pooled, pooled>."""
    zl = z.detach().clone().requires_grad_(True)
    pf, pooled = head_forward(zl, mode)
    total = 0.
    if g_pf is not None:
        total = total + (pf * g_pf).sum()
    if g_pooled is not None:
        total = total + (pooled * g_pooled).sum()
    total.backward()
    return zl.grad.clone()
