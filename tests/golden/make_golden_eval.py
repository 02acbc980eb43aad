"""Generates tests/golden/eval_metrics.npz with the reference's OWN eval_pipnet (pipnet/test.py:12-186), imported from
/root/reference and run on seeded tensors through a stand-in network that returns pre-computed (pooled, out) pairs --
the statistics under test start from those tensors, the backbone is not involved. Run in the dev container only:
    python tests/golden/make_golden_eval.py
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from pipnet.test import eval_pipnet  # noqa: E402


from golden_eval_cases import CASES, MULT, make_case  # noqa: E402


class _Cls(torch.nn.Module):
    def __init__(self, W, mult):
        super().__init__()
        self.weight = torch.nn.Parameter(W.clone(), requires_grad=False)
        self.normalization_multiplier = torch.nn.Parameter(torch.tensor(float(mult)), requires_grad=False)


class _Inner(torch.nn.Module):
    def __init__(self, W, mult):
        super().__init__()
        self._num_classes, self._num_prototypes = W.shape
        self._classification = _Cls(W, mult)


class StandInNet(torch.nn.Module):
    """What eval_pipnet touches: .module._num_classes / ._classification, __call__(xs, inference=True)."""

    def __init__(self, W, mult, batches):
        super().__init__()
        self.module = _Inner(W, mult)
        self.batches, self.i = batches, 0

    def forward(self, xs, inference=False):
        pooled, out, _ = self.batches[self.i]
        self.i += 1
        return None, pooled, out


class _Dataset:
    class_to_idx = {"positive": 0, "negative": 1}  # read by the two-class branch (pipnet/test.py:170)


class Loader(list):
    dataset = _Dataset()


def main():
    store = {}
    for name, kw in CASES.items():
        W, batches = make_case(**kw)
        net = StandInNet(W, MULT, batches)
        loader = Loader((torch.zeros(b[0].shape[0], 1), b[2]) for b in batches)
        with contextlib.redirect_stdout(io.StringIO()):
            info = eval_pipnet(net, loader, 0, "cpu", enforce_weight_sparsity=False)
        store[name + "/confusion_matrix"] = np.asarray(info["confusion_matrix"], dtype=np.int64)
        store[name + "/scalars"] = np.asarray([info[k] for k in ("test_accuracy", "top1_accuracy", "local_size_for_true_class", "local_size_for_all_classes", "prototypes_per_class", "almost_nonzeros", "num non-zero prototypes")], dtype=np.float64)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "eval_metrics.npz")
    np.savez_compressed(out, **store)
    print(out, {k: v.shape for k, v in store.items()})


if __name__ == "__main__":
    main()
