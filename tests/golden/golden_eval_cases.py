"""Seeded inputs of the evaluation-statistics golden file (shared by make_golden_eval.py, which runs the reference on
them, and tests/test_eval_metrics.py)."""
import torch


def make_case(seed, n_batches, B, P, K, neg_weights=False, zero_rows=0):
    """Sparse PIP-Net-like tensors: pooled in [0, 1] with most entries exactly 0 (the inference threshold), a sparse
    classifier with entries near the 1e-3 threshold, out = pooled @ relu(W).T (+ ties / abstentions)."""
    g = torch.Generator().manual_seed(seed)
    W = torch.zeros(K, P)
    for k in range(K):
        cols = torch.randperm(P, generator=g)[:6]
        W[k, cols] = torch.rand(6, generator=g) * 10 + 1
    near = torch.rand(K, P, generator=g) < 0.05
    W = torch.where(near & (W == 0), torch.rand(K, P, generator=g) * 4e-3, W)  # values around the threshold
    if neg_weights:
        W = torch.where(torch.rand(K, P, generator=g) < 0.02, -torch.rand(K, P, generator=g) * 5, W)
    batches = []
    for i in range(n_batches):
        pooled = torch.rand(B, P, generator=g)
        pooled = torch.where(pooled < 0.7, torch.zeros(()), pooled)
        pooled = torch.where(torch.rand(B, P, generator=g) < 0.05, torch.rand(B, P, generator=g) * 3e-3, pooled)
        if zero_rows:
            pooled[:zero_rows] = 0  # abstained images: all logits 0 (ties over every class)
        out = pooled @ torch.relu(W).t()
        ys = torch.randint(0, K, (B,), generator=g)
        ys[::2] = out[::2].argmax(dim=1)  # half of the labels are right
        batches.append((pooled, out, ys))
    return W, batches


CASES = {
    "small_k5": dict(seed=1, n_batches=3, B=8, P=64, K=5),
    "cub_like": dict(seed=2, n_batches=2, B=16, P=768, K=200),
    "two_class_neg": dict(seed=3, n_batches=4, B=7, P=48, K=2, neg_weights=True),
    "abstain": dict(seed=4, n_batches=2, B=12, P=96, K=9, zero_rows=3),
}
MULT = 2.0
