"""Evaluation statistics on the GPU: the part of the reference's test loop that runs right
after the forward on the same tensors (pipnet/test.py:67-159).

The reference builds `scores = pooled * weight.repeat(...)` as a [K, B, P] tensor every batch, reduces it three ways
(compute_local_explanation_sizes, prototypes per class, almost-non-zeros) with an `.item()` after each, and fills the
confusion matrix in a Python loop over the images. Here one kernel per batch (`cpn_eval_batch`) reads pooled [B, P],
out [B, K] and the [K, P] weight table and leaves every per-image statistic plus the confusion matrix on the device;
the host reads the accumulated sums once at the end. `eval_pipnet` keeps the reference's signature and the keys of its
`info` dict. The AUC / balanced-accuracy printing of the two-class case (sklearn, pipnet/test.py:159-183) stays with the
caller: `info` additionally carries `y_trues`, `y_preds`, `y_preds_classes` for it.

Ties: the arg-max class is the first maximum (as torch.max); the label's rank for top-1 / top-5 breaks ties towards the
lower class index (the reference's `topk` leaves tie order unspecified).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from . import ops

THRESHOLD = 1e-3  # pipnet/test.py:104, 110, 122


def acc_from_cm(cm) -> float:
    """Accuracy from a confusion matrix (pipnet/test.py:253-264): trace / total, 1 for an empty matrix."""
    cm = np.asarray(cm)
    assert cm.ndim == 2 and cm.shape[0] == cm.shape[1]
    total = cm.sum()
    return 1 if total == 0 else float(np.trace(cm)) / float(total)


def classification_weights(module) -> torch.Tensor:
    """[K, P] weights the reference multiplies `pooled` with (pipnet/test.py:51-65): the stored classifier weight for
    PIP-Net, get_prototype_importance_per_class stacked and transposed for Count-PIP-Net."""
    if hasattr(module, "_max_count"):
        per_proto = [module.get_prototype_importance_per_class(i) for i in range(module._num_prototypes)]
        w = torch.stack(per_proto, dim=0)  # [P, K]
        assert w.shape[0] == module._num_prototypes and w.shape[1] == module._num_classes
        return w.t().contiguous()
    return module._classification.weight


class BatchEvaluator:
    """Device-side accumulators for one evaluation pass over K classes and P prototypes."""

    def __init__(self, num_classes: int, num_prototypes: int, device, threshold: float = THRESHOLD):
        self.K, self.P, self.thr = int(num_classes), int(num_prototypes), float(threshold)
        self.device = torch.device(device)
        self.cm = torch.zeros(self.K, self.K, dtype=torch.int64, device=self.device)
        self.relevant = torch.empty(self.K, self.P, dtype=torch.uint8, device=self.device)
        self.protos_per_class = torch.empty(self.K, dtype=torch.int32, device=self.device)
        # running sums of the per-batch means the reference accumulates with .item() (pipnet/test.py:126-130, 146)
        self.sums = torch.zeros(6, dtype=torch.float64, device=self.device)
        self.abstained = torch.zeros((), dtype=torch.int64, device=self.device)
        self.batches = 0
        self._res: Dict[int, Dict[str, torch.Tensor]] = {}

    def _buffers(self, B: int) -> Dict[str, torch.Tensor]:
        if B not in self._res:
            i32 = lambda: torch.empty(B, dtype=torch.int32, device=self.device)
            f32 = lambda: torch.empty(B, dtype=torch.float32, device=self.device)
            self._res[B] = {"pred": i32(), "pred_prob": f32(), "max_score": f32(), "any_size": i32(),
                            "pred_size": i32(), "anz": i32(), "top1": i32(), "top5": i32(),
                            "relevant": self.relevant, "protos_per_class": self.protos_per_class}
        return self._res[B]

    def update(self, pooled: torch.Tensor, out: torch.Tensor, weights: torch.Tensor, ys: Optional[torch.Tensor],
               mult: float) -> Dict[str, torch.Tensor]:
        """One batch. Returns the per-image results (views of reused buffers: copy what must outlive the next call)."""
        B = pooled.shape[0]
        res = self._buffers(B)
        ys64 = None if ys is None else ys.to(device=self.device, dtype=torch.int64).contiguous()
        ops.eval_batch(pooled.float().contiguous(), out.float().contiguous(), weights.detach().float().contiguous(),
                       ys64, mult, self.thr, res, cm=self.cm if ys is not None else None)
        means = torch.stack([res["pred_size"].double().mean(), res["any_size"].double().mean(),
                             res["protos_per_class"].double().mean(), res["anz"].double().mean(),
                             res["top1"].double().mean(), res["top5"].double().mean()])
        self.sums += means
        self.abstained += (res["max_score"] == 0).sum()
        self.batches += 1
        return res

    def info(self) -> dict:
        n = max(self.batches, 1)
        s = self.sums.cpu().numpy()
        cm = self.cm.cpu().numpy()
        return {"confusion_matrix": cm, "test_accuracy": acc_from_cm(cm), "top1_accuracy": float(s[4]) / n,
                "top5_accuracy": float(s[5]) / n, "local_size_for_true_class": float(s[0]) / n,
                "local_size_for_all_classes": float(s[1]) / n, "prototypes_per_class": float(s[2]) / n,
                "almost_nonzeros": float(s[3]) / n, "abstained": int(self.abstained.item())}


@torch.no_grad()
def eval_pipnet(net, test_loader, epoch, device, log=None, progress_prefix: str = "Eval Epoch",
                enforce_weight_sparsity: bool = True, args=None) -> dict:
    """Same arguments and `info` keys as the reference's eval_pipnet (pipnet/test.py:12-186); `net` may be wrapped
    (`net.module`, as the reference expects) or bare."""
    net = net.to(device)
    net.eval()
    module = net.module if hasattr(net, "module") else net
    K = module._num_classes
    is_count = hasattr(module, "_max_count")
    # Count-PIP-Net: the table is built once before the loop (pipnet/test.py:51-63) and does not follow the in-loop
    # sparsity clamp; PIP-Net reads the live parameter every batch (:65, :99)
    weights = classification_weights(module)
    ev = BatchEvaluator(K, weights.shape[1], device)
    y_trues, y_preds, y_preds_classes = [], [], []
    for xs, ys in test_loader:
        xs, ys = xs.to(device), ys.to(device)
        if enforce_weight_sparsity:  # pipnet/test.py:72-74
            module._classification.weight.copy_(torch.clamp(module._classification.weight.data - 1e-3, min=0.))
        if not is_count:
            weights = module._classification.weight
        _, pooled, out = net(xs, inference=True)
        res = ev.update(pooled, out, weights, ys, float(module._classification.normalization_multiplier))
        y_preds.append(res["pred_prob"].clone())
        y_preds_classes.append(res["pred"].clone())
        y_trues.append(ys)
    info = ev.info()
    print("PIP-Net abstained from a decision for", info["abstained"], "images", flush=True)
    info["num non-zero prototypes"] = torch.gt(weights, 1e-3).any(dim=0).sum().item()
    cw = module._classification.weight
    print("sparsity ratio: ", (torch.numel(cw) - torch.count_nonzero(torch.relu(cw - 1e-3)).item()) / torch.numel(cw),
          flush=True)
    info["y_trues"] = torch.cat(y_trues).tolist() if y_trues else []
    info["y_preds"] = torch.cat(y_preds).tolist() if y_preds else []
    info["y_preds_classes"] = torch.cat(y_preds_classes).tolist() if y_preds_classes else []
    return info
