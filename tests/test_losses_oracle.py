"""CPU: the loss oracle (oracle/reference_losses.py) and the package's composed calculate_loss against the outputs of the
reference's OWN calculate_loss (tests/golden/losses_golden.npz, made by tests/golden/make_golden_losses.py), plus the
host logic of train_pipnet on a stand-in network."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from golden_loss_cases import CASES, make_inputs  # noqa: E402

from oracle import reference_losses as RL  # noqa: E402
from count_pipnet_b200.pipnet import train as T  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "losses_golden.npz"))
GRAD_KEYS = {"pf": "proto_features", "pooled": "pooled", "out": "out", "mult": "norm_mult"}


def _close(a, b, rtol=2e-6, atol=1e-7):
    np.testing.assert_allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rtol=rtol, atol=atol)


@pytest.mark.parametrize("name", list(CASES))
def test_seeded_inputs_are_the_golden_inputs(name):
    pf, pooled, out, ys, mult = make_inputs(**CASES[name])
    for key, t in zip(("pf", "pooled", "out", "ys", "mult"), (pf, pooled, out, ys, mult)):
        np.testing.assert_array_equal(t.numpy(), GOLD[f"{name}/in_{key}"])


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_matches_reference_calculate_loss(name):
    kw = CASES[name]
    pf, pooled, out, ys, mult = (torch.from_numpy(GOLD[f"{name}/in_{k}"]) for k in ("pf", "pooled", "out", "ys", "mult"))
    loss, acc, comps, grads = RL.loss_gradients(
        pf, pooled, out, ys, kw["align_w"], kw["tanh_w"], kw["class_w"], mult, kw["pretrain"], kw["finetune"],
        is_count=kw["is_count"], eps=1e-8, enforce_weight_sparsity=kw["enforce"], tanh_coeff=kw["coeff"])
    _close(loss, GOLD[f"{name}/loss"])
    assert acc == float(GOLD[f"{name}/acc"])
    for k, v in comps.items():
        _close(v, GOLD[f"{name}/comp_{k}"])
    for short, long in GRAD_KEYS.items():
        key = f"{name}/grad_{short}"
        if key in GOLD.files:
            _close(grads[long].numpy(), GOLD[key], rtol=1e-5, atol=1e-9)
        else:
            g = grads.get(long)
            assert g is None or float(g.abs().max()) == 0.0


@pytest.mark.parametrize("name", list(CASES))
def test_composed_calculate_loss_matches_reference(name):
    """The package's calculate_loss on CPU tensors (same signature as the reference's)."""
    kw = CASES[name]
    pf, pooled, out, ys, mult = (torch.from_numpy(GOLD[f"{name}/in_{k}"]) for k in ("pf", "pooled", "out", "ys", "mult"))
    leaves = [t.clone().requires_grad_(True) for t in (pf, pooled, out, mult)]
    loss, acc, comps = T.calculate_loss(leaves[0], leaves[1], leaves[2], ys, kw["align_w"], kw["tanh_w"], 0., kw["class_w"],
                                        leaves[3], kw["pretrain"], kw["finetune"], torch.nn.NLLLoss(reduction='mean'),
                                        None, is_count_pipnet=kw["is_count"], verbose=False, EPS=1e-8,
                                        enforce_weight_sparsity=kw["enforce"], net=None, tanh_loss_coeff=kw["coeff"])
    loss.backward()
    _close(loss.item(), GOLD[f"{name}/loss"])
    assert acc == float(GOLD[f"{name}/acc"])
    assert set(comps) == {"align", "align_weighted", "tanh", "tanh_weighted", "class", "class_weighted"}
    for k, v in comps.items():
        _close(v, GOLD[f"{name}/comp_{k}"])
    for short, leaf in zip(("pf", "pooled", "out", "mult"), leaves):
        key = f"{name}/grad_{short}"
        if key in GOLD.files:
            _close(leaf.grad.numpy(), GOLD[key], rtol=1e-5, atol=1e-9)


def test_head_oracle_gradients_are_autograd_of_the_modules():
    g = torch.Generator().manual_seed(0)
    z = torch.randn(3, 9, 4, 5, generator=g) * 2
    gpf = torch.randn(3, 9, 4, 5, generator=g)
    gpl = torch.randn(3, 9, generator=g)
    zl = z.clone().requires_grad_(True)
    pf = torch.nn.Softmax(dim=1)(zl)
    pooled = torch.nn.Sequential(torch.nn.AdaptiveMaxPool2d((1, 1)), torch.nn.Flatten())(pf)
    ((pf * gpf).sum() + (pooled * gpl).sum()).backward()
    np.testing.assert_allclose(RL.head_gradients(z, "max", gpf, gpl).numpy(), zl.grad.numpy(), rtol=1e-6, atol=1e-8)
    o_pf, o_pl = RL.head_forward(z, "sum")
    np.testing.assert_allclose(o_pl.numpy(), pf.detach().sum(dim=(2, 3)).numpy(), rtol=1e-6)


class _TinyNet(torch.nn.Module):
    """What train_pipnet touches: __call__(xs) -> (proto_features, pooled, out), ._classification."""

    def __init__(self, P=6, K=3):
        super().__init__()
        from count_pipnet_b200.pipnet.pipnet import NonNegLinear
        self._net = torch.nn.Conv2d(3, P, 3, padding=1)
        self._classification = NonNegLinear(P, K, bias=True)
        torch.nn.init.normal_(self._classification.weight, 1.0, 0.1)
        torch.nn.init.constant_(self._classification.bias, 0.)

    def forward(self, xs, inference=False):
        pf = torch.softmax(self._net(xs), dim=1)
        pooled = pf.amax(dim=(2, 3))
        return pf, pooled, self._classification(pooled)


@pytest.mark.parametrize("pretrain,finetune", [(True, False), (False, False), (False, True)])
def test_train_pipnet_epoch_on_cpu(pretrain, finetune, capsys):
    torch.manual_seed(0)
    net = _TinyNet()
    g = torch.Generator().manual_seed(1)
    loader = [(torch.randn(4, 3, 8, 8, generator=g), torch.randn(4, 3, 8, 8, generator=g),
               torch.randint(0, 3, (4,), generator=g)) for _ in range(3)]
    opt_net = torch.optim.SGD(net._net.parameters(), lr=0.05)
    opt_cls = torch.optim.SGD(net._classification.parameters(), lr=0.05)
    sch_net = torch.optim.lr_scheduler.LambdaLR(opt_net, lambda s: 1.0)
    sch_cls = torch.optim.lr_scheduler.CosineAnnealingWarmRestarts(opt_cls, T_0=5)
    before_net = net._net.weight.detach().clone()
    before_cls = net._classification.weight.detach().clone()
    info = T.train_pipnet(net, loader, opt_net, opt_cls, sch_net, sch_cls, torch.nn.NLLLoss(reduction='mean'), 1, 10,
                          torch.device("cpu"), pretrain=pretrain, finetune=finetune)
    capsys.readouterr()
    assert set(info) >= {"align_loss_raw", "tanh_loss_raw", "class_loss_raw", "align_loss_weighted", "tanh_loss_weighted",
                         "class_loss_weighted", "train_accuracy", "loss", "lrs_net", "lrs_class"}
    assert len(info["lrs_net"]) == 3 and np.isfinite(info["loss"])
    net_moved = not torch.equal(before_net, net._net.weight.detach())
    cls_moved = not torch.equal(before_cls, net._classification.weight.detach())
    assert net_moved == (not finetune)   # train.py:110-115: the backbone optimizer is skipped when fine-tuning
    assert cls_moved == (not pretrain)   # train.py:105-108: the classifier optimizer is skipped when pre-training
    if pretrain:
        assert info["class_loss_raw"] == 0.0 and info["train_accuracy"] == 0.0 and len(info["lrs_class"]) == 0
    else:
        assert float(net._classification.normalization_multiplier) >= 1.0   # train.py:136 clamp
        assert float(net._classification.weight.min()) >= 0.0


class _FakeCudaMap:
    """Stands in for a CUDA feature map in engine.use_train_head, which only looks at these three properties."""
    is_cuda = True
    dtype = torch.float32

    def dim(self):
        return 4


def test_use_train_head_decisions(capsys):
    """Which autograd forwards hand their head to the training kernels: plain nn.Softmax(dim=1) add-ons with the stock
    max pool (PIP-Net) or the spatial sum (Count-PIP-Net); never the Gumbel activation, never CPU tensors."""
    import torch.nn as nn
    from count_pipnet_b200 import engine, synth
    from count_pipnet_b200.pipnet import count_pipnet as my_count
    from count_pipnet_b200.pipnet import pipnet as my_pipnet
    pip, _ = my_pipnet.get_pipnet(5, synth.make_args(use_mid_layers=True, num_stages=1))
    pip_addon, _ = my_pipnet.get_pipnet(5, synth.make_args(num_features=8, use_mid_layers=True, num_stages=1))
    cnt_soft, _ = my_count.get_count_network(5, synth.make_args(num_features=8, use_mid_layers=True, num_stages=1,
                                                                activation="softmax"))
    cnt_gumbel, _ = my_count.get_count_network(5, synth.make_args(num_features=8, use_mid_layers=True, num_stages=1,
                                                                  activation="gumbel_softmax"))
    capsys.readouterr()
    fake = _FakeCudaMap()
    assert engine.use_train_head(pip, fake, want_max=True)
    assert engine.use_train_head(pip_addon, fake, want_max=True)
    assert engine.use_train_head(cnt_soft, fake, want_max=False)
    assert not engine.use_train_head(cnt_gumbel, fake, want_max=False)
    assert not engine.use_train_head(pip, torch.zeros(1, 4, 2, 2), want_max=True)          # CPU tensor
    pip._pool = nn.Sequential(nn.AdaptiveAvgPool2d((1, 1)), nn.Flatten())                   # not the stock pool
    assert not engine.use_train_head(pip, fake, want_max=True)
    pip_addon._add_on[-1].dim = 2                                                            # softmax over another axis
    assert not engine.use_train_head(pip_addon, fake, want_max=True)
    old = engine.TRAIN_KERNELS
    try:
        engine.TRAIN_KERNELS = False
        assert not engine.use_train_head(cnt_soft, fake, want_max=False)
    finally:
        engine.TRAIN_KERNELS = old
