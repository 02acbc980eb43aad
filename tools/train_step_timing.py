#!/usr/bin/env python
"""One training step of the reference's loop (pipnet/train.py:84-103: forward of both views, calculate_loss, backward)
through the drop-in modules on one B200, with the training-side head kernels on and off:

    python tools/train_step_timing.py [--config cfg2] [--pairs 32] [--iters 5]

"on"  = torch backbone under autograd + cpn_softmax_pool_nchw(_bwd) + cpn_pipnet_loss (count_pipnet_b200/pipnet/train.py)
"off" = the same modules with torch ops only (CPN_TRAIN_KERNELS=0 head, the reference's calculate_loss arithmetic)
The backbone's forward + backward is the same torch code on both sides; this shows what the head slice buys end to end."""
import argparse
import contextlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from count_pipnet_b200 import configs, engine  # noqa: E402
from count_pipnet_b200.pipnet import train as T  # noqa: E402
from count_pipnet_b200.pipnet.train_timing import eager_loss  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--pairs", type=int, default=32)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    cfg = configs.CONFIGS[a.config]
    with contextlib.redirect_stdout(sys.stderr):
        model = configs.build_model(a.config).to(dev).train()
    for m in model.modules():
        if m.__class__.__name__ == "StochasticDepth":
            m.p = 0.0
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2 * a.pairs, 3, cfg["size"], cfg["size"], generator=g).to(dev)
    ys = torch.randint(0, cfg["classes"], (a.pairs,), generator=g).to(dev)
    crit = torch.nn.NLLLoss(reduction='mean')
    mult = model._classification.normalization_multiplier

    def step(on):
        engine.TRAIN_KERNELS = on
        model.zero_grad(set_to_none=True)
        pf, pooled, out = model(x)
        if on:
            loss, _, _ = T.calculate_loss(pf, pooled, out, ys, 5., 2., 0., 2., mult, False, False, crit, None,
                                          verbose=False, EPS=1e-8)
        else:
            loss = eager_loss(pf, pooled, out, ys, mult)
        loss.backward()
        return loss

    res = {}
    for on in (True, False, True, False):
        for _ in range(2):
            step(on)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            loss = step(on)
        e1.record()
        torch.cuda.synchronize()
        res.setdefault(on, []).append((e0.elapsed_time(e1) / a.iters, loss.item()))
    engine.TRAIN_KERNELS = True
    print(f"{a.config}: one training step (forward of {2 * a.pairs} images, loss, backward), ms per step, two rounds each")
    for on in (True, False):
        print(f"  head + loss kernels {'on ' if on else 'off'}: " + ", ".join(f"{ms:.2f} ms (loss {l:.5f})" for ms, l in res[on]))
    print(f"  difference: {min(m for m, _ in res[False]) - min(m for m, _ in res[True]):.2f} ms per step")


if __name__ == "__main__":
    main()
