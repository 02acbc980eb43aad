#!/usr/bin/env python
"""Training-side head kernels at the training shapes of the BASELINE configs, next to stock PyTorch on the same GPU:

    python tools/train_loss_timing.py [--iters 20] > profiles/r02_train_loss_timing.md

(the measuring code is count_pipnet_b200/pipnet/train_timing.py; bench.py reports the same numbers as `train_head`)."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from count_pipnet_b200 import ops  # noqa: E402
from count_pipnet_b200.pipnet.train_timing import time_train_head  # noqa: E402

SHAPES = [  # name, image pairs B, P, H, W, K, pooling
    ("cfg2 PIP-Net ConvNeXt-26 224^2, 64 pairs", 64, 768, 26, 26, 200, "max"),
    ("cfg3 PIP-Net ResNet-50 224^2, 32 pairs", 32, 2048, 28, 28, 196, "max"),
    ("cfg1 Count-PIP-Net 192^2, 64 pairs", 64, 16, 24, 24, 9, "sum"),
    ("cfg5 PIP-Net 448^2, 16 pairs", 16, 768, 54, 54, 200, "max"),
]
NAMES = {"loss": "loss + all gradients", "head_fwd": "softmax + {} pool forward", "head_bwd": "softmax + {} pool backward"}


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
        r = time_train_head(B, P, H, W, K, mode, a.iters, dev, flush=flush)
        for i, op in enumerate(("loss", "head_fwd", "head_bwd")):
            d = r[op]
            print(f"| {name if i == 0 else ''} | {NAMES[op].format(mode)} | {d['ms']:.4f} | {d['gbs']:.0f} "
                  f"({d['gbs'] / peak:.2f}) | {d['eager_ms']:.4f} | {d['eager_ms'] / d['ms']:.1f}x |")


if __name__ == "__main__":
    main()
