#!/usr/bin/env python
"""Per-launch device times of the recorded forward plan (CUDA events between ops, averaged over passes), with the
algorithmic TFLOP/s and GB/s of each launch. Run on the GPU box:

    python tools/plan_profile.py [--batch 64] [--precision bf16] [--size 224] > gpurun_out/plan_profile.txt
"""
import argparse
import contextlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--net", default="convnext_tiny_26")
    ap.add_argument("--passes", type=int, default=5)
    ap.add_argument("--map", action="store_true")
    ap.add_argument("--config", default="", help="cfg1..cfg5 of count_pipnet_b200/configs.py (overrides --net/--size/--map)")
    args = ap.parse_args()
    from count_pipnet_b200 import configs, engine, synth
    from count_pipnet_b200.pipnet import pipnet as my_pipnet
    dev = torch.device("cuda:0")
    if args.config:
        cfg = configs.CONFIGS[args.config]
        model = configs.build_model(args.config).to(dev)
        args.size, args.map, args.net = cfg["size"], cfg["materialize_map"], args.config
    else:
        with contextlib.redirect_stdout(sys.stderr):
            model, _ = my_pipnet.get_pipnet(200, synth.make_args(net=args.net, num_features=0))
        synth.load_synth(model, seed=0)
        model = model.eval().to(dev)
    model.cpn_precision = args.precision
    model.materialize_proto_map = args.map
    x = synth.synth_images(args.batch, args.size).to(dev)
    with torch.no_grad():
        for _ in range(3):
            model(x, inference=True)
    torch.cuda.synchronize()
    rec = next(reversed(engine.state_for(model, dev).plans.values()))
    det = rec.plan.details()
    acc = None
    for _ in range(args.passes):
        rows = rec.plan.run_timed()
        if acc is None:
            acc = [[r[0], 0.0, r[2], r[3]] for r in rows]
        for a, r in zip(acc, rows):
            a[1] += r[1] / args.passes
    total = sum(a[1] for a in acc)
    print(f"# {args.net} B={args.batch} {args.size}^2 {args.precision}: {len(acc)} launches, {total:.3f} ms "
          f"({args.batch / total * 1e3:.0f} img/s)")
    print(f"{'#':>3s} {'kernel':16s} {'ms':>8s} {'%':>6s} {'TFLOP/s':>8s} {'GB/s':>8s}  shape")
    for i, (a, d) in enumerate(zip(acc, det)):
        name, ms, fl, by = a
        print(f"{i:3d} {name:16s} {ms:8.4f} {100 * ms / total:6.2f} {fl / ms / 1e9:8.1f} {by / ms / 1e6:8.1f}  {d}")


if __name__ == "__main__":
    main()
