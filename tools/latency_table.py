#!/usr/bin/env python
"""Small-batch latency of `net(x, inference=True)` (the reference's visualisation / attribution callers run B = 1 forwards:
util/vis_pipnet.py:208-213, util/interpret_idg.py:168-172): this package with the recorded plan replayed as a CUDA graph,
the same plan launched kernel by kernel, and the reference arithmetic in stock PyTorch eager on the same GPU.

    python tools/latency_table.py > gpurun_out/latency.md        (on the GPU box)
"""
import argparse
import os
import sys
import time
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def wall_ms(fn, iters):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
        torch.cuda.synchronize()  # latency as a caller sees it: launch + execution + sync, one forward at a time
    return (time.perf_counter() - t0) / iters * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=100)
    args = ap.parse_args()
    import bench
    from count_pipnet_b200 import configs, engine, synth
    dev = torch.device("cuda:0")
    print("| config | batch | precision | CUDA graph replay | kernel-by-kernel plan | PyTorch eager (reference arithmetic) |")
    print("|---|---|---|---|---|---|")
    for cfg_name in ("cfg1", "cfg2"):
        cfg = configs.CONFIGS[cfg_name]
        for B in (1, 8):
            x = synth.synth_images(B, cfg["size"], seed=5).to(dev)
            ns = types.SimpleNamespace(config=cfg_name)
            eager = bench.time_gpu_eager(ns, dev, B, steps=30, warmup=5)
            for precision in ("fp32", "bf16"):
                row = []
                for graphs in (True, False):
                    engine.USE_GRAPHS = graphs
                    model = configs.build_model(cfg_name).to(dev)
                    model.cpn_precision = precision
                    model.materialize_proto_map = False

                    def fwd():
                        with torch.no_grad():
                            return model(x, inference=True)
                    row.append(wall_ms(fwd, args.iters))
                    del model
                print(f"| {cfg_name} | {B} | {precision} | {row[0]:.3f} ms | {row[1]:.3f} ms | {eager['ms_per_step']:.3f} ms (device time, "
                      f"tf32 conv {eager['allow_tf32']['cudnn']}) |", flush=True)
    engine.USE_GRAPHS = True


if __name__ == "__main__":
    main()
