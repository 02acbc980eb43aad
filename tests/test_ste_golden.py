"""CPU: the package's straight-through functions (count_pipnet_b200/pipnet/count_pipnet_utils.py) against forward values
and backward gradients of the reference's OWN STE_Round, ClampSTE, ModifiedSTEFunction and OneHotEncoder
(tests/golden/ste_golden.npz, made by tests/golden/make_golden_ste.py). This pins the *effective* backward of the
reference, including the assignments through chained boolean indices that never reach the gradient
(pipnet/count_pipnet_utils.py:262-263, 313-319): zero counts get no gradient, and with `max_grad` only all-positive
entries do."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from golden_ste_cases import MAX_COUNT, STRATEGIES, make_case  # noqa: E402

from count_pipnet_b200.pipnet.count_pipnet_utils import (ClampSTE, ModifiedSTEFunction, OneHotEncoder,  # noqa: E402
                                                          STE_Round)

GOLD = np.load(os.path.join(HERE, "golden", "ste_golden.npz"))


def _inputs():
    counts, grad3, grad2 = make_case()
    np.testing.assert_array_equal(counts.numpy(), GOLD["counts"])
    np.testing.assert_array_equal(grad3.numpy(), GOLD["grad3"])
    np.testing.assert_array_equal(grad2.numpy(), GOLD["grad2"])
    return counts, grad3, grad2


def test_ste_round_and_clamp():
    counts, _, grad2 = _inputs()
    c = counts.clone().requires_grad_(True)
    r = STE_Round.apply(c)
    r.backward(grad2)
    np.testing.assert_array_equal(r.detach().numpy(), GOLD["round/out"])
    np.testing.assert_array_equal(c.grad.numpy(), GOLD["round/grad"])
    for ident in (True, False):
        c = (counts * 1.7 - 0.5).clone().requires_grad_(True)
        y = ClampSTE.apply(c, 0, MAX_COUNT, ident)
        y.backward(grad2)
        np.testing.assert_array_equal(y.detach().numpy(), GOLD[f"clamp_{int(ident)}/out"])
        np.testing.assert_array_equal(c.grad.numpy(), GOLD[f"clamp_{int(ident)}/grad"])


@pytest.mark.parametrize("gname", ["mixed", "pos"])
@pytest.mark.parametrize("respect", [False, True])
@pytest.mark.parametrize("sname", list(STRATEGIES))
def test_modified_ste_backward_is_the_references(sname, respect, gname):
    counts, grad3, _ = _inputs()
    g = grad3 if gname == "mixed" else grad3.abs() + 0.01
    c = counts.clone().requires_grad_(True)
    y = ModifiedSTEFunction.apply(c, MAX_COUNT, respect, STRATEGIES[sname])
    y.backward(g)
    key = f"mste_{sname}_{int(respect)}_{gname}"
    np.testing.assert_array_equal(y.detach().numpy(), GOLD[key + "/out"])
    np.testing.assert_array_equal(c.grad.numpy(), GOLD[key + "/grad"])


def test_onehot_encoder_module():
    counts, grad3, _ = _inputs()
    for use_ste in (True, False):
        enc = OneHotEncoder(MAX_COUNT, use_ste=use_ste, respect_active_grad=False, num_prototypes=counts.shape[1],
                            positive_grad_strategy=None)
        c = counts.clone().requires_grad_(True)
        y = enc(c)
        np.testing.assert_array_equal(y.detach().numpy(), GOLD[f"onehot_{int(use_ste)}/out"])
        if use_ste:
            y.backward(grad3.reshape(y.shape))
            np.testing.assert_array_equal(c.grad.numpy(), GOLD[f"onehot_{int(use_ste)}/grad"])
