"""Generates tests/golden/ste_golden.npz with the reference's OWN straight-through functions (pipnet/count_pipnet_utils.py:
STE_Round :41-55, ClampSTE :58-84, ModifiedSTEFunction :188-321, OneHotEncoder :86-139), imported from /root/reference:
forward values and the gradients their backward returns for seeded counts and incoming gradients, for every
positive_grad_strategy x respect_active_grad combination and both clamp strategies. Run in the dev container only:
    python tests/golden/make_golden_ste.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from pipnet.count_pipnet_utils import ClampSTE, ModifiedSTEFunction, OneHotEncoder, STE_Round  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from golden_ste_cases import MAX_COUNT, STRATEGIES, make_case  # noqa: E402


def main():
    store = {}
    counts, grad3, grad2 = make_case()
    store["counts"], store["grad3"], store["grad2"] = counts.numpy(), grad3.numpy(), grad2.numpy()
    # STE_Round / ClampSTE
    c = counts.clone().requires_grad_(True)
    r = STE_Round.apply(c)
    r.backward(grad2)
    store["round/out"], store["round/grad"] = r.detach().numpy(), c.grad.numpy()
    for ident in (True, False):
        c = (counts * 1.7 - 0.5).clone().requires_grad_(True)
        y = ClampSTE.apply(c, 0, MAX_COUNT, ident)
        y.backward(grad2)
        store[f"clamp_{int(ident)}/out"], store[f"clamp_{int(ident)}/grad"] = y.detach().numpy(), c.grad.numpy()
    # ModifiedSTEFunction, every strategy x respect_active_grad; the all-positive variant makes the strategies differ
    for sname, strat in STRATEGIES.items():
        for respect in (False, True):
            for gname, g in (("mixed", grad3), ("pos", grad3.abs() + 0.01)):
                c = counts.clone().requires_grad_(True)
                y = ModifiedSTEFunction.apply(c, MAX_COUNT, respect, strat)
                y.backward(g)
                key = f"mste_{sname}_{int(respect)}_{gname}"
                store[key + "/out"], store[key + "/grad"] = y.detach().numpy(), c.grad.numpy()
                print(key, "nonzero grads", int((c.grad != 0).sum()), "of", c.grad.numel())
    # OneHotEncoder module (flattened output), with and without STE
    for use_ste in (True, False):
        enc = OneHotEncoder(MAX_COUNT, use_ste=use_ste, respect_active_grad=False, num_prototypes=counts.shape[1],
                            positive_grad_strategy=None)
        c = counts.clone().requires_grad_(True)
        y = enc(c)
        store[f"onehot_{int(use_ste)}/out"] = y.detach().numpy()
        if use_ste:
            y.backward(grad3.reshape(y.shape))
            store[f"onehot_{int(use_ste)}/grad"] = c.grad.numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ste_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
