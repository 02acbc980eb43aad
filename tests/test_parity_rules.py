"""CPU checks of the parity rules themselves (tests/parity_rules.py): the reference's own outputs pass in both modes,
a permitted boundary flip is accounted for exactly, and a flip away from the boundary or a wrong class is rejected."""
import os

import numpy as np
import pytest

import cases
from count_pipnet_b200.pipnet import count_pipnet as my_count
from count_pipnet_b200.pipnet import pipnet as my_pipnet
from parity_rules import check_outputs

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SMALL = ["count_linear_softmax", "count_onehot_softmax_noste_p192", "pipnet_convnext26_mid5_noinference",
         "count_linear_full_softmax_train_outputs", "pipnet_resnet18_64"]


def load(name):
    case = cases.CASES[name]
    model = cases.build(case, my_pipnet, my_count)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    raw = None
    if case["inference"]:
        raw = g["map_sum"] if case["model"] == "count_pipnet" else g["map_max"]
    return case, sd, g, raw


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", SMALL)
def test_reference_outputs_pass(name, precision):
    case, sd, g, raw = load(name)
    st = check_outputs(case, sd, precision, g["pooled"], g["out"], g["pooled"], g["out"], raw)
    assert st["flips"] == 0 and st["class_checked"] > 0


def test_count_flip_is_accounted_or_rejected():
    case, sd, g, raw = load("count_onehot_softmax_noste_p192")
    raw = raw.copy()
    pooled = g["pooled"].copy()
    # pick an entry, move the reference's raw sum onto a rounding boundary, flip our count by one
    b, p = np.argwhere((g["pooled"] >= 1) & (g["pooled"] <= 2))[0]
    raw[b, p] = g["pooled"][b, p] + 0.499
    pooled[b, p] = g["pooled"][b, p] + 1.0
    import torch
    from parity_rules import _encode
    w = torch.relu(sd["_classification.weight"]).numpy()
    out = g["out"] + (_encode(sd, case, pooled) - _encode(sd, case, g["pooled"])) @ w.T
    st = check_outputs(case, sd, "bf16", pooled, out, g["pooled"], g["out"], raw)
    assert st["flips"] == 1 and st["flip_images"] == 1
    with pytest.raises(AssertionError):  # same flip, but the logits do not reflect it
        check_outputs(case, sd, "bf16", pooled, g["out"] + 5.0 * (np.arange(g["out"].shape[1]) == 0), g["pooled"], g["out"], raw)
    with pytest.raises(AssertionError):  # a flip where the reference is nowhere near the boundary
        check_outputs(case, sd, "bf16", pooled, out, g["pooled"], g["out"], g["map_sum"])
    with pytest.raises(AssertionError):  # fp32 mode accepts no flip at all
        check_outputs(case, sd, "fp32", pooled, out, g["pooled"], g["out"], raw)


def test_wrong_class_is_rejected():
    case, sd, g, raw = load("pipnet_resnet18_64")
    out = g["out"].copy()
    k = out.shape[1]
    wrong = (g["pred"][0] + 1) % k
    out[0, wrong] = out[0].max() + 1.0
    for precision in ("fp32", "bf16"):
        with pytest.raises(AssertionError):
            check_outputs(case, sd, precision, g["pooled"], out, g["pooled"], g["out"], raw)
