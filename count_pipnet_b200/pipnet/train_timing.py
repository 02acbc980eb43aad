"""CUDA-event timing of the training-side head kernels (pipnet/train.py mirror: cpn_pipnet_loss, cpn_softmax_pool_nchw,
cpn_softmax_pool_nchw_bwd) at a training shape, next to the same arithmetic composed from stock PyTorch ops on the same
GPU (the reference's calculate_loss, train.py:165-265, and its Softmax + pooling modules with their autograd). Used by
bench.py (`train_head` block) and tools/train_loss_timing.py.

GB/s = algorithmic bytes / time: loss = read the map once + write its gradient once (8 B per map element), head forward
8 B, head backward with both incoming gradients 12 B.
"""
from __future__ import annotations

import torch

from .. import _lib, ops


def _timed(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        flush.add_(1.0)  # > L2: the map of the previous iteration is gone from the cache
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        total += e0.elapsed_time(e1)
    return total / iters


def eager_loss(pf, pooled, out, ys1, mult):
    """pipnet/train.py:165-229 + 253-265 with stock torch ops (align 5, tanh 2, class 2; EPS 1e-8)."""
    ys = torch.cat([ys1, ys1])
    pooled1, pooled2 = pooled.chunk(2)
    pf1, pf2 = pf.chunk(2)
    e2 = pf2.flatten(start_dim=2).permute(0, 2, 1).flatten(end_dim=1)
    e1 = pf1.flatten(start_dim=2).permute(0, 2, 1).flatten(end_dim=1)

    def al(a, b):
        return -torch.log(torch.einsum("nc,nc->n", [a, b]) + 1e-12).mean()
    a_loss = (al(e1, e2.detach()) + al(e2, e1.detach())) / 2.
    t_loss = -(torch.log(torch.tanh(torch.sum(pooled1, dim=0)) + 1e-8).mean()
               + torch.log(torch.tanh(torch.sum(pooled2, dim=0)) + 1e-8).mean()) / 2.
    c_loss = torch.nn.functional.nll_loss(torch.log_softmax(torch.log1p(out ** mult), dim=1), ys)
    return 5. * a_loss + 2. * t_loss + 2. * c_loss


def time_train_head(B, P, H, W, K, mode="max", iters=20, device=None, eager=True, flush=None):
    """B image PAIRS (2B images in the forward). -> {op: {"ms", "gbs", "eager_ms"}} for ops loss / head_fwd / head_bwd."""
    dev = device or torch.device("cuda:0")
    if flush is None:
        flush = torch.zeros(64 * 1024 * 1024, device=dev)  # 256 MB
    g = torch.Generator().manual_seed(0)
    z = (torch.randn(2 * B, P, H, W, generator=g) * 3).to(dev)
    Wc = torch.relu(torch.randn(K, P, generator=g)).to(dev)
    ys = torch.randint(0, K, (B,), generator=g).to(dev)
    mult = torch.tensor([1.3], device=dev)
    pf = torch.softmax(z, dim=1)
    pooled = pf.amax(dim=(2, 3)) if mode == "max" else pf.sum(dim=(2, 3)) * 0.05
    out = pooled @ Wc.t()
    elems, HW = pf.numel(), H * W
    res = {}

    g_pf, g_pl, g_o = torch.empty_like(pf), torch.empty_like(pooled), torch.empty_like(out)
    r8 = torch.empty(8, device=dev)
    ws = torch.empty(ops.pipnet_loss_ws_floats(B, P, HW), device=dev)
    t = _timed(lambda: ops.pipnet_loss(pf, pooled, out, ys, mult, B, P, HW, K, 5., 2., 2., 1., 1e-8, False, False, True,
                                       g_pf, g_pl, g_o, r8, ws), iters, flush)
    res["loss"] = {"ms": t, "gbs": 8.0 * elems / t / 1e6}
    if eager:
        leaves = [x.clone().requires_grad_(True) for x in (pf, pooled, out, mult)]

        def run():
            for l in leaves:
                l.grad = None
            eager_loss(*leaves[:3], ys, leaves[3]).backward()
        res["loss"]["eager_ms"] = _timed(run, iters, flush)
        del leaves

    pf_o, pl_o = torch.empty_like(z), torch.empty(2 * B, P, device=dev)
    arg = torch.empty(2 * B, P, device=dev, dtype=torch.int32)
    ws2 = torch.empty(ops.softmax_pool_nchw_ws_floats(2 * B, P, HW), device=dev)
    m = _lib.POOL_MAX if mode == "max" else _lib.POOL_SUM
    t = _timed(lambda: ops.softmax_pool_nchw(z, m, pf_o, pl_o, arg, ws2), iters, flush)
    res["head_fwd"] = {"ms": t, "gbs": 8.0 * elems / t / 1e6}

    def pool(y):
        return torch.nn.functional.adaptive_max_pool2d(y, (1, 1)).flatten(1) if mode == "max" else y.sum(dim=(2, 3))
    if eager:
        def run_fwd():
            y = torch.softmax(z, dim=1)
            return y, pool(y)
        res["head_fwd"]["eager_ms"] = _timed(run_fwd, iters, flush)

    gz = torch.empty_like(z)
    gpl = torch.randn(2 * B, P, generator=g).to(dev)
    t = _timed(lambda: ops.softmax_pool_nchw_bwd(pf_o, g_pf, gpl, arg, m, gz), iters, flush)
    res["head_bwd"] = {"ms": t, "gbs": 12.0 * elems / t / 1e6}
    if eager:
        zl = z.clone().requires_grad_(True)
        y = torch.softmax(zl, dim=1)
        pl = pool(y)

        def run_bwd():
            zl.grad = None
            torch.autograd.backward([y, pl], [g_pf, gpl], retain_graph=True)
        res["head_bwd"]["eager_ms"] = _timed(run_bwd, iters, flush)
        del zl, y, pl
    del z, pf, g_pf, pf_o, gz
    torch.cuda.empty_cache()
    return res
