#!/usr/bin/env python
"""CUDA-event timing of the training-side head kernels at training shapes, next to the same arithmetic composed from
stock PyTorch ops on the same GPU (the reference's calculate_loss / Softmax + pooling modules and their autograd):

    python tools/train_loss_timing.py [--iters 20] > profiles/r02_train_loss_timing.md

Per shape: fused loss (values + all gradients, one call) vs eager loss forward + backward; head forward vs eager softmax +
pooling; head backward vs eager autograd. GB/s = algorithmic bytes / time (loss: read the map once, write its gradient
once = 8 B per map element; head forward: 8 B; head backward with both incoming gradients: 12 B).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from count_pipnet_b200 import _lib, ops  # noqa: E402
from count_pipnet_b200.pipnet import train as T  # noqa: E402

SHAPES = [  # name, image pairs B, P, H, W, K, pooling
    ("cfg2 PIP-Net ConvNeXt-26 224^2, 64 pairs", 64, 768, 26, 26, 200, "max"),
    ("cfg3 PIP-Net ResNet-50 224^2, 32 pairs", 32, 2048, 28, 28, 196, "max"),
    ("cfg1 Count-PIP-Net 192^2, 64 pairs", 64, 16, 24, 24, 9, "sum"),
    ("cfg5 PIP-Net 448^2, 16 pairs", 16, 768, 54, 54, 200, "max"),
]


def timed(fn, iters, flush):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    ops.check_device()
    peak = 6546.2
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    flush = torch.zeros(64 * 1024 * 1024, device=dev)  # 256 MB
    print("# Training-side head kernels vs stock PyTorch on the same B200 (CUDA events, L2 flushed between iterations)\n")
    print(f"HBM peak used for the fractions: {peak} GB/s (MEASURED_PEAKS.json, copy bandwidth)\n")
    print("| shape | op | kernels ms | GB/s (frac of peak) | PyTorch eager ms | speed-up |")
    print("|---|---|---|---|---|---|")
    for name, B, P, H, W, K, mode in SHAPES:
        g = torch.Generator().manual_seed(0)
        z = (torch.randn(2 * B, P, H, W, generator=g) * 3).to(dev)
        Wc = torch.relu(torch.randn(K, P, generator=g)).to(dev)
        ys = torch.randint(0, K, (B,), generator=g).to(dev)
        mult = torch.tensor([1.3], device=dev)
        pf = torch.softmax(z, dim=1)
        pooled = pf.amax(dim=(2, 3)) if mode == "max" else pf.sum(dim=(2, 3)) * 0.05
        out = pooled @ Wc.t()
        elems = pf.numel()
        HW = H * W

        # --- loss: values + gradients
        g_pf, g_pl, g_o = torch.empty_like(pf), torch.empty_like(pooled), torch.empty_like(out)
        res = torch.empty(8, device=dev)
        ws = torch.empty(ops.pipnet_loss_ws_floats(B, P, HW), device=dev)
        fused = lambda: ops.pipnet_loss(pf, pooled, out, ys, mult, B, P, HW, K, 5., 2., 2., 1., 1e-8, False, False, True,  # noqa: E731
                                        g_pf, g_pl, g_o, res, ws)
        leaves = [t.clone().requires_grad_(True) for t in (pf, pooled, out, mult)]

        def eager():
            for l in leaves:
                l.grad = None
            eager_loss(*leaves[:3], ys, leaves[3]).backward()
        t_f, t_e = timed(fused, a.iters, flush), timed(eager, a.iters, flush)
        gbs = 8.0 * elems / t_f / 1e6
        print(f"| {name} | loss + all gradients | {t_f:.4f} | {gbs:.0f} ({gbs / peak:.2f}) | {t_e:.4f} | {t_e / t_f:.1f}x |")

        # --- head forward
        pf_o, pl_o = torch.empty_like(z), torch.empty(2 * B, P, device=dev)
        arg = torch.empty(2 * B, P, device=dev, dtype=torch.int32)
        ws2 = torch.empty(ops.softmax_pool_nchw_ws_floats(2 * B, P, HW), device=dev)
        m = _lib.POOL_MAX if mode == "max" else _lib.POOL_SUM
        fwd = lambda: ops.softmax_pool_nchw(z, m, pf_o, pl_o, arg, ws2)  # noqa: E731

        def eager_fwd():
            y = torch.softmax(z, dim=1)
            return y, (torch.nn.functional.adaptive_max_pool2d(y, (1, 1)).flatten(1) if mode == "max" else y.sum(dim=(2, 3)))
        t_f, t_e = timed(fwd, a.iters, flush), timed(eager_fwd, a.iters, flush)
        gbs = 8.0 * elems / t_f / 1e6
        print(f"| | softmax + {mode} pool forward | {t_f:.4f} | {gbs:.0f} ({gbs / peak:.2f}) | {t_e:.4f} | {t_e / t_f:.1f}x |")

        # --- head backward (both incoming gradients)
        gz = torch.empty_like(z)
        gpl = torch.randn(2 * B, P, generator=g).to(dev)
        bwd = lambda: ops.softmax_pool_nchw_bwd(pf_o, g_pf, gpl, arg, m, gz)  # noqa: E731
        zl = z.clone().requires_grad_(True)
        y = torch.softmax(zl, dim=1)
        pl = torch.nn.functional.adaptive_max_pool2d(y, (1, 1)).flatten(1) if mode == "max" else y.sum(dim=(2, 3))

        def eager_bwd():
            zl.grad = None
            torch.autograd.backward([y, pl], [g_pf, gpl], retain_graph=True)
        t_f, t_e = timed(bwd, a.iters, flush), timed(eager_bwd, a.iters, flush)
        gbs = 12.0 * elems / t_f / 1e6
        print(f"| | softmax + {mode} pool backward | {t_f:.4f} | {gbs:.0f} ({gbs / peak:.2f}) | {t_e:.4f} | {t_e / t_f:.1f}x |")
        del z, pf, g_pf, pf_o, gz, zl, y, leaves
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
