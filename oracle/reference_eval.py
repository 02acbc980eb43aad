"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the per-batch statistics of the reference's
test loop, pipnet/test.py:82-137, compute_local_explanation_sizes (:266-318) and topk_accuracy (util/func.py:22-37).

Parity pinning: tests/golden/eval_metrics.npz holds the `info` dict of the reference's OWN eval_pipnet run in the dev
container on seeded tensors (tests/golden/make_golden_eval.py imports pipnet.test from /root/reference and feeds it a
stand-in network that returns pre-computed pooled / out); tests/test_eval_metrics.py checks this restatement against it.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch


def local_explanation_sizes(scores: torch.Tensor, ys_pred: torch.Tensor, threshold: float = 1e-3):
    """scores [K, B, P] -> (any-class sizes [B], predicted-class sizes [B]) as pipnet/test.py:266-318."""
    relevant = torch.abs(scores) > threshold
    any_sizes = relevant.any(dim=0).sum(dim=1)
    per_class = relevant.sum(dim=2)  # [K, B]
    pred_sizes = per_class[ys_pred, torch.arange(scores.shape[1])]  # the diagonal of index_select(..., ys_pred)
    return any_sizes, pred_sizes


def batch_statistics(pooled: torch.Tensor, out: torch.Tensor, weights: torch.Tensor, ys: torch.Tensor,
                     mult: float) -> Dict[str, torch.Tensor]:
    """One batch of pipnet/test.py:82-137; weights [K, P] as the reference multiplies them (unclamped)."""
    pooled, out, weights = pooled.float(), out.float(), weights.float()
    K = out.shape[1]
    max_out, ys_pred = torch.max(out, dim=1)  # :82
    pred_prob = torch.amax(torch.softmax(torch.log1p(out ** torch.tensor(float(mult))), dim=1), dim=1)  # :83
    scores = pooled.unsqueeze(0) * weights.unsqueeze(1)  # [K, B, P], :99-104
    any_sizes, pred_sizes = local_explanation_sizes(scores, ys_pred, 1e-3)
    protos_per_class = torch.count_nonzero(torch.gt(torch.relu(scores - 1e-3).mean(dim=1), 0.).float(), dim=1)  # :108-113
    anz = torch.count_nonzero(torch.gt(torch.abs(pooled), 1e-3).float(), dim=1)  # :121-124
    cm = np.zeros((K, K), dtype=np.int64)
    for y_pred, y_true in zip(ys_pred.tolist(), ys.tolist()):  # :135-137
        cm[y_true][y_pred] += 1
    # label rank with ties towards the lower class index (the reference's topk leaves tie order open)
    vt = out.gather(1, ys.view(-1, 1))
    idx = torch.arange(K).view(1, -1)
    rank = ((out > vt) | ((out == vt) & (idx < ys.view(-1, 1)))).sum(dim=1)
    return {"pred": ys_pred, "pred_prob": pred_prob, "max_score": max_out, "any_size": any_sizes,
            "pred_size": pred_sizes, "protos_per_class": protos_per_class, "anz": anz, "cm": cm,
            "top1": (rank < 1).long(), "top5": (rank < min(5, K)).long()}


def eval_info(batches, weights: torch.Tensor, mult: float) -> dict:
    """The accumulation of pipnet/test.py:126-159 over `batches` = [(pooled, out, ys), ...]."""
    K = weights.shape[0]
    cm = np.zeros((K, K), dtype=np.int64)
    acc = {"pred_size": 0., "any_size": 0., "protos_per_class": 0., "anz": 0., "top1": 0.}
    abstained = 0
    for pooled, out, ys in batches:
        st = batch_statistics(pooled, out, weights, ys, mult)
        cm += st["cm"]
        for k in acc:
            acc[k] += st[k].float().mean().item()
        abstained += int((st["max_score"] == 0).sum())
    n = len(batches)
    total = cm.sum()
    return {"confusion_matrix": cm, "test_accuracy": 1 if total == 0 else np.trace(cm) / total,
            "top1_accuracy": acc["top1"] / n, "local_size_for_true_class": acc["pred_size"] / n,
            "local_size_for_all_classes": acc["any_size"] / n, "prototypes_per_class": acc["protos_per_class"] / n,
            "almost_nonzeros": acc["anz"] / n, "abstained": abstained}
