"""GPU parity of the training side of the head (through the C ABI): cpn_pipnet_loss against the outputs of the
reference's OWN calculate_loss (tests/golden/losses_golden.npz) and against the CPU oracle at training shapes;
cpn_softmax_pool_nchw / _bwd against the oracle's autograd; the drop-in modules under autograd with the kernels on
against the same modules composed from torch ops.

Tolerances (north star: floating point within 1e-4 relative): loss terms 1e-5 relative; gradients max|delta| <= 1e-4 *
max|reference| per tensor (and element-wise rtol 1e-3 / atol 1e-4 * max for the large ones); accuracy counts and
arg-max positions exact."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from golden_loss_cases import CASES, make_inputs  # noqa: E402

from oracle import reference_losses as RL  # noqa: E402
from count_pipnet_b200 import _lib, engine, ops, synth  # noqa: E402
from count_pipnet_b200.pipnet import count_pipnet as my_count  # noqa: E402
from count_pipnet_b200.pipnet import pipnet as my_pipnet  # noqa: E402
from count_pipnet_b200.pipnet import train as T  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(HERE, "golden", "losses_golden.npz"))


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("name", list(CASES))
def test_calculate_loss_matches_reference_golden(cuda_device, name):
    kw = CASES[name]
    pf, pooled, out, ys, mult = (torch.from_numpy(GOLD[f"{name}/in_{k}"]).to(cuda_device)
                                 for k in ("pf", "pooled", "out", "ys", "mult"))
    leaves = [t.clone().requires_grad_(True) for t in (pf, pooled, out, mult)]
    loss, acc, comps = T.calculate_loss(leaves[0], leaves[1], leaves[2], ys, kw["align_w"], kw["tanh_w"], 0., kw["class_w"],
                                        leaves[3], kw["pretrain"], kw["finetune"], torch.nn.NLLLoss(reduction='mean'),
                                        None, is_count_pipnet=kw["is_count"], verbose=False, EPS=1e-8,
                                        enforce_weight_sparsity=kw["enforce"], net=None, tanh_loss_coeff=kw["coeff"])
    assert loss.is_cuda and loss.requires_grad
    loss.backward()
    assert abs(loss.item() - float(GOLD[f"{name}/loss"])) <= 1e-5 * abs(float(GOLD[f"{name}/loss"]))
    assert acc == float(GOLD[f"{name}/acc"])
    for k, v in comps.items():
        g = float(GOLD[f"{name}/comp_{k}"])
        assert abs(v - g) <= 1e-5 * abs(g) + 1e-9, (k, v, g)
    for short, leaf in zip(("pf", "pooled", "out", "mult"), leaves):
        key = f"{name}/grad_{short}"
        if key in GOLD.files:
            assert leaf.grad is not None, short
            assert rel(leaf.grad.cpu().numpy(), GOLD[key]) <= 1e-4, (short, rel(leaf.grad.cpu().numpy(), GOLD[key]))
        else:
            assert leaf.grad is None or float(leaf.grad.abs().max()) == 0.0, short


def _loss_call(pf, pooled, out, ys, mult, B, P, HW, K, flags, grads=True):
    dev = pf.device
    pretrain, finetune, use_log1p = flags
    g_pf = torch.full_like(pf, float("nan")) if grads and not finetune else None
    g_pl = torch.full_like(pooled, float("nan")) if grads and not finetune else None
    g_o = torch.full_like(out, float("nan")) if grads and not pretrain else None
    res = torch.full((8,), float("nan"), device=dev)
    ws = torch.full((ops.pipnet_loss_ws_floats(B, P, HW),), float("nan"), device=dev)
    ops.pipnet_loss(pf, pooled, out, ys, mult, B, P, HW, K, 5.0, 2.0, 2.0, 1.0, 1e-8, pretrain, finetune, use_log1p, g_pf,
                    g_pl, g_o, res, ws)
    torch.cuda.synchronize()
    return g_pf, g_pl, g_o, res


# (B pairs, P, H, W, K): the cfg2 training map at a small batch, the ResNet-50 map, ragged tiles, one position
@pytest.mark.parametrize("shape", [(4, 768, 26, 26, 200), (3, 2048, 7, 7, 196), (2, 33, 5, 9, 11), (5, 16, 24, 24, 9),
                                   (2, 7, 1, 1, 3), (1, 130, 6, 11, 1000), (2, 33, 5, 8, 4), (3, 5, 2, 2, 2),
                                   (1, 2048, 28, 28, 196)])
@pytest.mark.parametrize("flags", [(False, False, True), (True, False, True), (False, True, True), (False, False, False)])
def test_pipnet_loss_c_abi_matches_oracle(cuda_device, shape, flags):
    B, P, H, W, K = shape
    pretrain, finetune, use_log1p = flags
    pf, pooled, out, ys, mult = make_inputs(B=B, P=P, H=H, W=W, K=K, is_count=False, mult=1.4, seed=B * 1000 + P)
    o_loss, o_acc, o_comps, o_grads = RL.loss_gradients(pf, pooled, out, ys, 5.0, 2.0, 2.0, mult, pretrain, finetune,
                                                        eps=1e-8, enforce_weight_sparsity=use_log1p)
    d = [t.to(cuda_device) for t in (pf, pooled, out, ys, mult)]
    g_pf, g_pl, g_o, res = _loss_call(*d, B, P, H * W, K, flags)
    r = res.cpu().numpy().astype(np.float64)
    assert np.isfinite(r).all()
    assert abs(r[0] - o_comps["align"]) <= 1e-5 * abs(o_comps["align"])
    assert abs(r[1] - o_comps["tanh"]) <= 1e-5 * abs(o_comps["tanh"]) + 1e-9
    assert abs(r[2] - o_comps["class"]) <= 1e-5 * abs(o_comps["class"])
    assert abs(r[3] - o_loss) <= 1e-5 * abs(o_loss)
    assert r[4] == round(o_acc * 2 * B)
    if not finetune:
        assert rel(g_pf.cpu().numpy(), o_grads["proto_features"].numpy()) <= 1e-4
        assert rel(g_pl.cpu().numpy(), o_grads["pooled"].numpy()) <= 1e-4
    if not pretrain:
        assert rel(g_o.cpu().numpy(), o_grads["out"].numpy()) <= 1e-4
        if use_log1p:
            gm = float(o_grads["norm_mult"])
            assert abs(r[5] - gm) <= 1e-4 * abs(gm) + 1e-5
    # same call again: same bits (fixed-order reductions, no atomics)
    g_pf2, g_pl2, g_o2, res2 = _loss_call(*d, B, P, H * W, K, flags)
    assert torch.equal(res, res2)
    for a, b in ((g_pf, g_pf2), (g_pl, g_pl2), (g_o, g_o2)):
        assert (a is None and b is None) or torch.equal(a, b)
    # values without gradients
    _, _, _, res3 = _loss_call(*d, B, P, H * W, K, flags, grads=False)
    assert torch.equal(res[:5], res3[:5])


def test_pipnet_loss_rejects_bad_arguments(cuda_device):
    pf, pooled, out, ys, mult = (t.to(cuda_device) for t in make_inputs(B=2, P=8, H=3, W=3, K=4, is_count=False, mult=1.0,
                                                                       seed=0))
    res = torch.empty(8, device=cuda_device)
    ws = torch.empty(ops.pipnet_loss_ws_floats(2, 8, 9), device=cuda_device)
    with pytest.raises(_lib.CpnError, match="fine-tuning"):
        ops.pipnet_loss(pf, pooled, out, ys, mult, 2, 8, 9, 4, 5., 2., 2., 1., 1e-8, False, True, True, torch.empty_like(pf),
                        None, None, res, ws)
    with pytest.raises(_lib.CpnError, match="exclude"):
        ops.pipnet_loss(pf, pooled, out, ys, mult, 2, 8, 9, 4, 5., 2., 2., 1., 1e-8, True, True, True, None, None, None,
                        res, ws)
    with pytest.raises(_lib.CpnError, match="unless pretrain"):
        ops.pipnet_loss(pf, pooled, None, None, None, 2, 8, 9, 4, 5., 2., 2., 1., 1e-8, False, False, True, None, None,
                        None, res, ws)
    with pytest.raises(ValueError, match="NLLLoss"):
        T.calculate_loss(pf, pooled, out, ys, 5., 2., 0., 2., mult, False, False, torch.nn.CrossEntropyLoss(), None,
                         verbose=False)


def test_backward_scales_by_the_incoming_gradient(cuda_device):
    pf, pooled, out, ys, mult = (t.to(cuda_device) for t in make_inputs(B=3, P=20, H=5, W=5, K=6, is_count=False, mult=1.2,
                                                                       seed=11))
    grads = []
    for factor in (1.0, 0.25):
        leaves = [t.clone().requires_grad_(True) for t in (pf, pooled, out, mult)]
        loss, _, _ = T.calculate_loss(leaves[0], leaves[1], leaves[2], ys, 5., 2., 0., 2., leaves[3], False, False,
                                      torch.nn.NLLLoss(reduction='mean'), None, verbose=False, EPS=1e-8)
        (loss * factor).backward()
        grads.append([l.grad.clone() for l in leaves])
    for a, b in zip(*grads):
        torch.testing.assert_close(b, a * 0.25, rtol=1e-6, atol=0)
    buf = torch.arange(10000, device=cuda_device, dtype=torch.float32)
    ops.scale_unless_one(buf, torch.ones(1, device=cuda_device))
    assert torch.equal(buf, torch.arange(10000, device=cuda_device, dtype=torch.float32))
    ops.scale_unless_one(buf, torch.full((1,), 0.5, device=cuda_device))
    assert torch.equal(buf, torch.arange(10000, device=cuda_device, dtype=torch.float32) * 0.5)


@pytest.mark.parametrize("mode", ["max", "sum"])
@pytest.mark.parametrize("shape", [(4, 768, 26, 26), (3, 2048, 7, 7), (2, 33, 5, 9), (6, 16, 24, 24), (2, 7, 1, 1),
                                   (2, 33, 5, 8), (3, 5, 2, 2), (2, 130, 6, 22), (2, 2048, 28, 28), (1, 1100, 4, 8),
                                   (2, 800, 3, 3)])
def test_softmax_pool_head_forward_backward(cuda_device, mode, shape):
    g = torch.Generator().manual_seed(sum(shape))
    z = torch.randn(*shape, generator=g) * 3
    gpf = torch.randn(*shape, generator=g)
    gpl = torch.randn(shape[0], shape[1], generator=g)
    o_pf, o_pl = RL.head_forward(z, mode)
    zd = z.to(cuda_device).requires_grad_(True)
    pf, pooled = T.softmax_pool_head(zd, mode)
    assert rel(pf.detach().cpu().numpy(), o_pf.numpy()) <= 1e-5
    assert rel(pooled.detach().cpu().numpy(), o_pl.numpy()) <= 1e-5
    if mode == "max":  # pooled is an element of the map the kernel wrote, bit for bit
        assert torch.equal(pooled.detach(), pf.detach().flatten(2).amax(dim=2))
    ((pf * gpf.to(cuda_device)).sum() + (pooled * gpl.to(cuda_device)).sum()).backward()
    o_gz = RL.head_gradients(z, mode, gpf, gpl)
    assert rel(zd.grad.cpu().numpy(), o_gz.numpy()) <= 1e-4
    # only one of the two outputs used downstream
    for use_pf in (True, False):
        zd2 = z.to(cuda_device).requires_grad_(True)
        pf2, pl2 = T.softmax_pool_head(zd2, mode)
        ((pf2 * gpf.to(cuda_device)).sum() if use_pf else (pl2 * gpl.to(cuda_device)).sum()).backward()
        o = RL.head_gradients(z, mode, gpf if use_pf else None, None if use_pf else gpl)
        assert rel(zd2.grad.cpu().numpy(), o.numpy()) <= 1e-4


@pytest.mark.parametrize("hw", [(3, 4), (3, 3), (5, 8), (7, 7)])
def test_max_pool_ties_route_to_the_first_position(cuda_device, hw):
    """Equal columns give equal softmax values at several positions; AdaptiveMaxPool2d's backward picks the first.
    Both lane layouts (H*W % 4 == 0 and not) and maps wider than one 32-position tile."""
    N, P, (H, W) = 2, 5, hw
    g = torch.Generator().manual_seed(3)
    col = torch.randn(N, P, 1, 1, generator=g)
    z = col.expand(N, P, H, W).contiguous()
    gpl = torch.randn(N, P, generator=g)
    pf = torch.empty(N, P, H, W, device=cuda_device)
    pooled = torch.empty(N, P, device=cuda_device)
    arg = torch.full((N, P), -1, device=cuda_device, dtype=torch.int32)
    ws = torch.empty(ops.softmax_pool_nchw_ws_floats(N, P, H * W), device=cuda_device)
    ops.softmax_pool_nchw(z.to(cuda_device), _lib.POOL_MAX, pf, pooled, arg, ws)
    assert int(arg.abs().max()) == 0
    o = RL.head_gradients(z, "max", None, gpl)
    assert float(o[:, :, 0, 0].abs().max()) > 1e-3 and float(o.flatten(2)[:, :, 1:].abs().max()) == 0.0
    gz = torch.empty_like(pf)
    ops.softmax_pool_nchw_bwd(pf, None, gpl.to(cuda_device), arg, _lib.POOL_MAX, gz)
    assert rel(gz.cpu().numpy(), o.numpy()) <= 1e-4


def test_unaligned_maps_take_the_scalar_lane_layout(cuda_device):
    """A map whose base is not 16-byte aligned (a view into a larger buffer) cannot use 128-bit accesses."""
    B, P, H, W, K = 2, 24, 6, 6, 5
    pf, pooled, out, ys, mult = (t.to(cuda_device) for t in make_inputs(B=B, P=P, H=H, W=W, K=K, is_count=False, mult=1.1,
                                                                       seed=21))
    n = pf.numel()
    store_in, store_g = torch.empty(n + 1, device=cuda_device), torch.empty(n + 1, device=cuda_device)
    pf_u = store_in[1:].view_as(pf)
    pf_u.copy_(pf)
    g_u = store_g[1:].view_as(pf)
    assert pf_u.data_ptr() % 16 == 4 and pf_u.is_contiguous()
    res_a, res_u = torch.empty(8, device=cuda_device), torch.empty(8, device=cuda_device)
    ws = torch.empty(ops.pipnet_loss_ws_floats(B, P, H * W), device=cuda_device)
    g_a, gpl, go = torch.empty_like(pf), torch.empty_like(pooled), torch.empty_like(out)
    ops.pipnet_loss(pf, pooled, out, ys, mult, B, P, H * W, K, 5., 2., 2., 1., 1e-8, False, False, True, g_a, gpl, go, res_a, ws)
    ops.pipnet_loss(pf_u, pooled, out, ys, mult, B, P, H * W, K, 5., 2., 2., 1., 1e-8, False, False, True, g_u, gpl, go,
                    res_u, ws)
    torch.testing.assert_close(res_u, res_a, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(g_u, g_a, rtol=1e-5, atol=1e-9)
    z = torch.randn(2 * B, P, H, W, device=cuda_device)
    z_u = store_in[1:].view_as(pf)
    z_u.copy_(z)
    outs = []
    for src in (z, z_u):
        y, pl = torch.empty_like(z), torch.empty(2 * B, P, device=cuda_device)
        arg = torch.empty(2 * B, P, device=cuda_device, dtype=torch.int32)
        ws2 = torch.empty(ops.softmax_pool_nchw_ws_floats(2 * B, P, H * W), device=cuda_device)
        ops.softmax_pool_nchw(src, _lib.POOL_MAX, y, pl, arg, ws2)
        gz = store_g[1:].view_as(pf) if src is z_u else torch.empty_like(z)
        ops.softmax_pool_nchw_bwd(y, None, pl, arg, _lib.POOL_MAX, gz)
        outs.append((y.clone(), pl.clone(), arg.clone(), gz.clone()))
    torch.testing.assert_close(outs[1][0], outs[0][0], rtol=1e-6, atol=1e-9)
    assert torch.equal(outs[1][2], outs[0][2])
    torch.testing.assert_close(outs[1][3], outs[0][3], rtol=1e-5, atol=1e-9)


def _train_step(model, x, ys, is_count, kernels_on):
    engine.TRAIN_KERNELS = kernels_on
    try:
        model.zero_grad(set_to_none=True)
        pf, pooled, out = model(x)
        if kernels_on:
            loss, acc, comps = T.calculate_loss(pf, pooled, out, ys, 5., 2., 0., 2.,
                                                model._classification.normalization_multiplier, False, False,
                                                torch.nn.NLLLoss(reduction='mean'), None, is_count_pipnet=is_count,
                                                verbose=False, EPS=1e-8, tanh_loss_coeff=0.05)
        else:  # the same step composed from torch ops on the same GPU (the oracle's restatement of train.py)
            loss, acc, comps = RL.pipnet_losses(pf, pooled, out, ys, 5., 2., 2.,
                                                model._classification.normalization_multiplier, False, False,
                                                is_count=is_count, eps=1e-8, tanh_coeff=0.05)
        loss.backward()
        grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}
        return pf.detach(), pooled.detach(), out.detach(), loss.item(), acc, grads
    finally:
        engine.TRAIN_KERNELS = True


@pytest.mark.parametrize("kind", ["pipnet", "pipnet_addon", "count_softmax"])
def test_training_step_through_the_modules(cuda_device, kind):
    """Forward + loss + backward through the drop-in modules under autograd: head and loss kernels on, against the same
    modules with torch ops only. Same backbone code on both sides; cuDNN's backward is not run-to-run identical, hence the
    1e-3 bound on parameter gradients."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    if kind == "count_softmax":
        args = synth.make_args(num_features=16, use_mid_layers=True, num_stages=2, activation="softmax",
                               intermediate_layer="linear")
        model, _ = my_count.get_count_network(9, args, max_count=3, use_ste=True)
        K = 9
    else:
        args = synth.make_args(num_features=24 if kind == "pipnet_addon" else 0, use_mid_layers=True, num_stages=2)
        model, _ = my_pipnet.get_pipnet(7, args)
        K = 7
    synth.load_synth(model, seed=0)
    model = model.to(cuda_device).train()
    for m in model.modules():  # stochastic depth would draw different masks in the two runs
        if m.__class__.__name__ == "StochasticDepth":
            m.p = 0.0
    g = torch.Generator().manual_seed(5)
    x = torch.randn(4, 3, 96, 96, generator=g).to(cuda_device)
    ys = torch.randint(0, K, (2,), generator=g).to(cuda_device)
    a = _train_step(model, x, ys, kind == "count_softmax", kernels_on=True)
    b = _train_step(model, x, ys, kind == "count_softmax", kernels_on=False)
    assert rel(a[0].cpu().numpy(), b[0].cpu().numpy()) <= 1e-5
    assert rel(a[1].cpu().numpy(), b[1].cpu().numpy()) <= 1e-5
    assert rel(a[2].cpu().numpy(), b[2].cpu().numpy()) <= 1e-5
    assert abs(a[3] - b[3]) <= 1e-5 * abs(b[3]) and a[4] == b[4]
    assert set(a[5]) == set(b[5]) and len(a[5]) > 10
    worst = max(rel(a[5][n].cpu().numpy(), b[5][n].cpu().numpy()) for n in a[5] if float(b[5][n].abs().max()) > 1e-12)
    assert worst <= 1e-3, worst
