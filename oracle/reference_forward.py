"""ORACLE - test infrastructure only. Never imported by the product path (count_pipnet_b200/); only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.

A CPU fp32 restatement of the reference forward (TarasKutsyk/Count_PIPNet @ cf2985d5) written as plain functions over
a state dict - no nn.Module of the reference or of this package, no torchvision import - so it can travel to the GPU
box, where /root/reference does not exist. Each function cites the reference lines it restates. ConvNeXt arithmetic
lives in torchvision 0.26.0 (third party, not vendored by the reference; call sites features/convnext_features.py:
3,50-51,79-80); it is restated from torchvision/models/convnext.py:31-177.

Parity is PINNED: tests/golden/make_golden.py runs the real reference modules (imported from /root/reference in the
dev container) on the seeded weights/inputs of count_pipnet_b200/synth.py, and tests/test_oracle_golden.py checks
this restatement against those committed outputs (bit-level for counts/classes, <= 2e-6 relative for floats).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


# ------------------------------------------------------------------------------------------------ ConvNeXt
def layer_norm_2d(x, w, b, eps=1e-6):
    """LayerNorm2d: permute to NHWC, layer_norm over C, permute back (torchvision/models/convnext.py:31-36)."""
    return F.layer_norm(x.permute(0, 2, 3, 1), (x.shape[1],), w, b, eps).permute(0, 3, 1, 2)


def cn_block(sd: SD, p: str, x):
    """CNBlock.forward in eval mode (torchvision/models/convnext.py:51-67; stochastic depth is identity in eval,
    torchvision/ops/stochastic_depth.py:32-33)."""
    c = x.shape[1]
    y = F.conv2d(x, sd[p + "block.0.weight"], sd[p + "block.0.bias"], padding=3, groups=c)
    y = y.permute(0, 2, 3, 1)
    y = F.layer_norm(y, (c,), sd[p + "block.2.weight"], sd[p + "block.2.bias"], 1e-6)
    y = F.linear(y, sd[p + "block.3.weight"], sd[p + "block.3.bias"])
    y = F.gelu(y)  # exact erf form
    y = F.linear(y, sd[p + "block.5.weight"], sd[p + "block.5.bias"])
    y = y.permute(0, 3, 1, 2)
    return x + sd[p + "layer_scale"] * y


def convnext_features(sd: SD, prefix: str, x, stride_threshold: int):
    """features[0..n] present in the state dict (features/convnext_features.py:17-36 keeps 0..num_stages);
    stride-2 convs with in_channels > threshold run at stride 1 (features/convnext_features.py:5-15)."""
    idx = sorted({int(k[len(prefix):].split(".")[1]) for k in sd if k.startswith(prefix + "features.")})
    for i in idx:
        p = f"{prefix}features.{i}."
        if i == 0:  # stem: conv 4x4 s4 + LayerNorm2d (torchvision/models/convnext.py:120-131)
            x = F.conv2d(x, sd[p + "0.weight"], sd[p + "0.bias"], stride=4)
            x = layer_norm_2d(x, sd[p + "1.weight"], sd[p + "1.bias"])
        elif i % 2 == 1:  # stage of CNBlocks
            nblk = len({k[len(p):].split(".")[0] for k in sd if k.startswith(p)})
            for j in range(nblk):
                x = cn_block(sd, f"{p}{j}.", x)
        else:  # downsample: LayerNorm2d + conv 2x2 (torchvision/models/convnext.py:146-150)
            x = layer_norm_2d(x, sd[p + "0.weight"], sd[p + "0.bias"])
            stride = 1 if x.shape[1] > stride_threshold else 2
            x = F.conv2d(x, sd[p + "1.weight"], sd[p + "1.bias"], stride=stride)
    return x


# ------------------------------------------------------------------------------------------------ ResNet
def _bn(sd: SD, p: str, x):
    """Eval-mode BatchNorm2d (running statistics, eps 1e-5)."""
    return F.batch_norm(x, sd[p + "running_mean"], sd[p + "running_var"], sd[p + "weight"], sd[p + "bias"],
                        False, 0.0, 1e-5)


def resnet_features(sd: SD, prefix: str, x):
    """ResNet_features.forward (features/resnet_features.py:199-210) with Bottleneck (:97-117) or BasicBlock
    (:55-74) blocks; layer2 has stride 2, layer3/layer4 stride 1 (:152-154)."""
    x = F.conv2d(x, sd[prefix + "conv1.weight"], None, stride=2, padding=3)
    x = F.relu(_bn(sd, prefix + "bn1.", x))
    x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    for li, lstride in ((1, 1), (2, 2), (3, 1), (4, 1)):
        lp = f"{prefix}layer{li}."
        nblk = len({k[len(lp):].split(".")[0] for k in sd if k.startswith(lp)})
        for j in range(nblk):
            p = f"{lp}{j}."
            stride = lstride if j == 0 else 1
            bottleneck = (p + "conv3.weight") in sd
            identity = x
            if (p + "downsample.0.weight") in sd:
                identity = _bn(sd, p + "downsample.1.", F.conv2d(x, sd[p + "downsample.0.weight"], None, stride=stride))
            if bottleneck:
                y = F.relu(_bn(sd, p + "bn1.", F.conv2d(x, sd[p + "conv1.weight"])))
                y = F.relu(_bn(sd, p + "bn2.", F.conv2d(y, sd[p + "conv2.weight"], None, stride=stride, padding=1)))
                y = _bn(sd, p + "bn3.", F.conv2d(y, sd[p + "conv3.weight"]))
            else:
                y = F.relu(_bn(sd, p + "bn1.", F.conv2d(x, sd[p + "conv1.weight"], None, stride=stride, padding=1)))
                y = _bn(sd, p + "bn2.", F.conv2d(y, sd[p + "conv2.weight"], None, padding=1))
            x = F.relu(y + identity)
    return x
