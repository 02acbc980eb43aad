"""CPU: the oracle (oracle/) against the golden vectors produced by the real reference (tests/golden/make_golden.py).
This is what pins the oracle; the GPU parity tests then compare the CUDA path with both."""
import os

import numpy as np
import pytest
import torch

import cases
import oracle
from count_pipnet_b200.pipnet import count_pipnet as my_count
from count_pipnet_b200.pipnet import pipnet as my_pipnet

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# big-backbone cases take a few seconds each on 8 cores; keep the whole CPU suite in the minutes range
CPU_CASES = [n for n in cases.CASES if n != "cfg5_pipnet_convnext26_448_map"]


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def run_oracle(name):
    case = cases.CASES[name]
    model = cases.build(case, my_pipnet, my_count)  # only its state dict is used
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    x = cases.inputs(case)
    shape = None
    noise = None
    if case["args"].get("activation") == "gumbel_softmax":
        g = load_golden(name)
        shape = tuple(int(v) for v in g["map_shape"])
        noise = cases.gumbel_noise(shape)
    return oracle.forward(sd, cases.oracle_cfg(case), x, inference=case["inference"], gumbel_noise=noise)


@pytest.mark.parametrize("name", CPU_CASES)
def test_oracle_matches_reference_golden(name):
    g = load_golden(name)
    proto, pooled, out = run_oracle(name)
    case = cases.CASES[name]
    assert tuple(proto.shape) == tuple(int(v) for v in g["map_shape"])
    # classes, top prototype and (for count models at inference) the counts must be identical
    assert np.array_equal(out.argmax(1).numpy(), g["pred"])
    assert np.array_equal(pooled.argmax(1).numpy(), g["top_proto"])
    if case["model"] == "count_pipnet" and case["inference"]:
        assert np.array_equal(pooled.numpy(), g["pooled"])
    # floats: same arithmetic, possibly another CPU -> a few ulps
    def rel(a, b):
        return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))
    assert rel(pooled.numpy(), g["pooled"]) <= 1e-5
    assert rel(out.numpy(), g["out"]) <= 1e-5
    flat = proto.reshape(proto.shape[0], proto.shape[1], -1)
    assert rel(flat.amax(2).numpy(), g["map_max"]) <= 1e-5
    assert rel(flat.sum(2).numpy(), g["map_sum"]) <= 1e-5


def test_onehot_known_answers():
    """The value table of the reference's own tests/test-onehot-pass.py:21-26,49-60 (reshaped to [4,3,3])."""
    from oracle import reference_heads as rh
    x = np.array([[0, 1, 3], [0.05, 2, 2.9], [1, 0, 0.2], [3, 2, 1]], dtype=np.float32)
    enc = rh.onehot_encoding(np.rint(x), 3)
    expect = {0: (0, 0, 0), 1: (1, 0, 0), 2: (0, 1, 0), 3: (0, 0, 1)}
    for b in range(4):
        for p in range(3):
            assert tuple(enc[b, p]) == expect[int(np.rint(x[b, p]))], (b, p)
    # same table through this package's module (the STE path rounds first: 2.9 -> 3, 0.2 -> 0)
    from count_pipnet_b200.pipnet.count_pipnet_utils import OneHotEncoder
    mod = OneHotEncoder(3, use_ste=True, num_prototypes=3)
    assert np.array_equal(mod(torch.from_numpy(x)).detach().numpy().reshape(4, 3, 3), enc)


def test_counts_from_controlled_map():
    """tests/test-count-pipnet.py:253-293 (commented in the reference): a 1x2x4x4 map sums to [[3, 2]]."""
    from oracle import reference_heads as rh
    m = torch.zeros(1, 2, 4, 4)
    m[0, 0, 0, 0] = m[0, 0, 1, 2] = m[0, 0, 3, 3] = 1.0
    m[0, 1, 2, 2] = m[0, 1, 0, 3] = 1.0
    sd = {"_classification.weight": torch.ones(2, 6)}
    counts, _ = rh.count_head(sd, m, True, "onehot", 3, True)
    assert counts.tolist() == [[3.0, 2.0]]
