"""Seeded inputs of the loss parity cases, shared by make_golden_losses.py (reference side) and the tests."""
import torch

# name -> dict(B pairs, P, H, W, K, flags and weights as train.py:53-61 sets them)
CASES = {
    "pipnet_pretrain": dict(B=3, P=12, H=5, W=5, K=7, pretrain=True, finetune=False, align_w=0.3, tanh_w=5., class_w=0.,
                            is_count=False, enforce=True, mult=1.0, coeff=1.0, seed=1),
    "pipnet_train": dict(B=4, P=16, H=6, W=6, K=5, pretrain=False, finetune=False, align_w=5., tanh_w=2., class_w=2.,
                         is_count=False, enforce=True, mult=1.0, coeff=1.0, seed=2),
    "pipnet_train_mult": dict(B=2, P=40, H=7, W=5, K=9, pretrain=False, finetune=False, align_w=5., tanh_w=2.,
                              class_w=2., is_count=False, enforce=True, mult=1.7, coeff=1.0, seed=3),
    "pipnet_finetune": dict(B=3, P=10, H=4, W=4, K=6, pretrain=False, finetune=True, align_w=5., tanh_w=2., class_w=2.,
                            is_count=False, enforce=True, mult=1.3, coeff=1.0, seed=4),
    "count_train_noenforce": dict(B=3, P=8, H=6, W=6, K=4, pretrain=False, finetune=False, align_w=5., tanh_w=2.,
                                  class_w=2., is_count=True, enforce=False, mult=1.0, coeff=0.05, seed=5),
    "count_pretrain": dict(B=2, P=33, H=3, W=11, K=3, pretrain=True, finetune=False, align_w=1.0, tanh_w=5., class_w=0.,
                           is_count=True, enforce=True, mult=1.0, coeff=0.3, seed=6),
}


def make_inputs(B, P, H, W, K, is_count, mult, seed, scale=3.0, **_):
    """-> proto_features [2B, P, H, W], pooled [2B, P], out [2B, K] (>= 0, with exact zeros), ys [B], norm_mult [1]."""
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(2 * B, P, H, W, generator=g) * scale
    pf = torch.softmax(z, dim=1)
    pooled = pf.sum(dim=(2, 3)) if is_count else pf.amax(dim=(2, 3))
    Wc = torch.relu(torch.randn(K, P, generator=g))
    Wc[0] = 0.  # a class without prototypes: out == 0 exactly (pow / log1p at zero)
    out = pooled @ Wc.t()
    ys = torch.randint(0, K, (B,), generator=g)
    return pf.contiguous(), pooled.contiguous(), out.contiguous(), ys, torch.tensor([float(mult)])
