"""Seeded inputs of the straight-through parity cases (make_golden_ste.py on the reference side, the tests on ours)."""
import torch

MAX_COUNT = 3
STRATEGIES = {"none": None, "current": "current_grad", "max": "max_grad"}


def make_case(B=6, P=40, seed=9):
    """counts [B, P] in [0, 4.5) with exact zeros, halves and values beyond max_count; gradients with ties and zeros."""
    g = torch.Generator().manual_seed(seed)
    counts = torch.rand(B, P, generator=g) * 4.5
    counts[0, :8] = torch.tensor([0.0, 0.04, 0.5, 1.5, 2.5, 3.0, 3.49, 4.4])
    grad3 = torch.randn(B, P, MAX_COUNT, generator=g)
    grad3[1, :5] = 0.0                      # all-equal slots
    grad3[2, :5, 1] = grad3[2, :5, 0]       # ties between two slots
    grad2 = torch.randn(B, P, generator=g)
    return counts, grad3, grad2
