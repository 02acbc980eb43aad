"""Generates tests/golden/losses_golden.npz with the reference's OWN calculate_loss (pipnet/train.py:165-243), imported
from /root/reference, on the seeded cases of golden_loss_cases.py: loss, accuracy, components and the autograd gradients
w.r.t. proto_features, pooled, out and the normalization multiplier. Run in the dev container only:
    python tests/golden/make_golden_losses.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from pipnet.train import calculate_loss  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden_loss_cases import CASES, make_inputs  # noqa: E402


def main():
    store = {}
    crit = torch.nn.NLLLoss(reduction='mean')
    for name, kw in CASES.items():
        pf, pooled, out, ys, mult = make_inputs(**kw)
        leaves = [t.clone().requires_grad_(True) for t in (pf, pooled, out, mult)]
        loss, acc, comps = calculate_loss(leaves[0], leaves[1], leaves[2], ys, kw["align_w"], kw["tanh_w"], 0.,
                                          kw["class_w"], leaves[3], kw["pretrain"], kw["finetune"], crit, None,
                                          is_count_pipnet=kw["is_count"], verbose=False, EPS=1e-8,
                                          enforce_weight_sparsity=kw["enforce"], net=None, tanh_loss_coeff=kw["coeff"])
        loss.backward()
        for key, t in zip(("pf", "pooled", "out", "ys", "mult"), (pf, pooled, out, ys, mult)):
            store[f"{name}/in_{key}"] = t.numpy()
        store[f"{name}/loss"] = np.float64(loss.item())
        store[f"{name}/acc"] = np.float64(acc)
        for k, v in comps.items():
            store[f"{name}/comp_{k}"] = np.float64(v)
        for key, leaf in zip(("pf", "pooled", "out", "mult"), leaves):
            if leaf.grad is not None:
                store[f"{name}/grad_{key}"] = leaf.grad.numpy()
        print(name, "loss", loss.item(), "acc", acc, {k: round(v, 5) for k, v in comps.items()},
              "grads", [k for k, l in zip(("pf", "pooled", "out", "mult"), leaves) if l.grad is not None])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "losses_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
