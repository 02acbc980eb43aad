"""ConvNeXt-tiny feature extractor with the reference's stride surgery, as a drop-in for
`features/convnext_features.py` of TarasKutsyk/Count_PIPNet (reference file:line cited per symbol).

The module tree reproduces the parameter names of the reference backbone (`features.<i>.<j>.block.<k>.weight`,
`features.<i>.<j>.layer_scale`, ...; torchvision/models/convnext.py:39-151) so reference checkpoints load with
`strict=True` and `util/args.py` can group parameters by name. The modules are ordinary `nn.Module`s: called with
autograd enabled they run as composed PyTorch ops (training, attribution); the inference forward of the whole network
goes through the fused sm_100a kernels in `count_pipnet_b200.engine`, which read these parameters directly.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

# ConvNeXt-tiny geometry (torchvision/models/convnext.py:311-317)
TINY_DIMS = (96, 192, 384, 768)
TINY_DEPTHS = (3, 3, 9, 3)
TINY_SD_PROB = 0.1
LAYER_SCALE_INIT = 1e-6


class LayerNorm2d(nn.LayerNorm):
    """LayerNorm over the channel dim of an NCHW map (torchvision/models/convnext.py:31-36)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.permute(0, 2, 3, 1)
        x = F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps)
        return x.permute(0, 3, 1, 2)


class Permute(nn.Module):
    """Parameter-free dim shuffle; keeps the `block.<k>` numbering of the reference CNBlock."""

    def __init__(self, dims: List[int]):
        super().__init__()
        self.dims = dims

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return torch.permute(x, self.dims)


class StochasticDepth(nn.Module):
    """Row-mode stochastic depth (torchvision/ops/stochastic_depth.py:8-44); identity in eval mode."""

    def __init__(self, p: float):
        super().__init__()
        self.p = p

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self.training or self.p == 0.0:
            return x
        keep = 1.0 - self.p
        noise = torch.empty([x.shape[0]] + [1] * (x.ndim - 1), dtype=x.dtype, device=x.device).bernoulli_(keep)
        if keep > 0.0:
            noise.div_(keep)
        return x * noise


class CNBlock(nn.Module):
    """x + layer_scale * fc2(GELU(fc1(LN(dwconv7x7(x)))))  (torchvision/models/convnext.py:39-67)."""

    def __init__(self, dim: int, layer_scale: float, sd_prob: float):
        super().__init__()
        self.block = nn.Sequential(
            nn.Conv2d(dim, dim, kernel_size=7, padding=3, groups=dim, bias=True),
            Permute([0, 2, 3, 1]),
            nn.LayerNorm(dim, eps=1e-6),
            nn.Linear(dim, 4 * dim, bias=True),
            nn.GELU(),
            nn.Linear(4 * dim, dim, bias=True),
            Permute([0, 3, 1, 2]),
        )
        self.layer_scale = nn.Parameter(torch.ones(dim, 1, 1) * layer_scale)
        self.stochastic_depth = StochasticDepth(sd_prob)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x + self.stochastic_depth(self.layer_scale * self.block(x))


class ConvNeXtTiny(nn.Module):
    """Feature part of ConvNeXt-tiny: `features` = [stem, stage1, down, stage2, down, stage3, down, stage4]
    (torchvision/models/convnext.py:118-151). `avgpool` / `classifier` are identities, as the reference sets them
    (features/convnext_features.py:55-56)."""

    def __init__(self):
        super().__init__()
        layers: List[nn.Module] = []
        layers.append(nn.Sequential(
            nn.Conv2d(3, TINY_DIMS[0], kernel_size=4, stride=4, padding=0, bias=True),
            LayerNorm2d(TINY_DIMS[0], eps=1e-6),
        ))
        total_blocks = sum(TINY_DEPTHS)
        block_id = 0
        for si, (dim, depth) in enumerate(zip(TINY_DIMS, TINY_DEPTHS)):
            stage = []
            for _ in range(depth):
                sd = TINY_SD_PROB * block_id / (total_blocks - 1.0)
                stage.append(CNBlock(dim, LAYER_SCALE_INIT, sd))
                block_id += 1
            layers.append(nn.Sequential(*stage))
            if si + 1 < len(TINY_DIMS):
                layers.append(nn.Sequential(
                    LayerNorm2d(dim, eps=1e-6),
                    nn.Conv2d(dim, TINY_DIMS[si + 1], kernel_size=2, stride=2),
                ))
        self.features = nn.Sequential(*layers)
        self.avgpool = nn.Identity()
        self.classifier = nn.Identity()
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _features_forward(self, x)


def _features_forward(net: nn.Module, x: torch.Tensor) -> torch.Tensor:
    """`net._net(x)` as callers use it (NCHW in, NCHW out). Inference on a CUDA tensor runs the fused kernels;
    with autograd enabled (training / attribution) the composed modules run so gradients exist."""
    from .. import engine  # local import: engine imports this module for isinstance checks
    if engine.use_fused(net, x):
        return engine.backbone_forward_nchw(net, x)
    return net.features(x)


def replace_convlayers_convnext(model: nn.Module, threshold: int) -> nn.Module:
    """Halve the stride of every stride-2 conv whose in_channels > threshold (kernel 2, pad 0 kept):
    >100 -> 26x26 grid at 224^2, >300 -> 13x13 (features/convnext_features.py:5-15)."""
    for m in model.modules():
        if isinstance(m, nn.Conv2d) and m.stride[0] == 2 and m.in_channels > threshold:
            m.stride = tuple(s // 2 for s in m.stride)
    return model


class MidLayerConvNeXt(nn.Module):
    """Keeps `features[0 .. num_stages]` of the full model (features/convnext_features.py:17-36): index 0 is the
    stem, odd indices are block stages, even indices downsample layers."""

    def __init__(self, original_model: nn.Module, num_stages: int = 2):
        super().__init__()
        self.features = nn.Sequential()
        feats = getattr(original_model, "features", None)
        if feats is not None and len(feats) > 0:
            self.features.add_module("0", feats[0])
            for i in range(min(num_stages, len(feats) - 1)):
                self.features.add_module(str(i + 1), feats[i + 1])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _features_forward(self, x)


def _load_pretrained(model: nn.Module) -> None:
    # The reference downloads torchvision's ImageNet weights (features/convnext_features.py:50). Same source here;
    # the state-dict keys of `features.*` are identical, the classifier keys are dropped.
    from torchvision import models as tvm
    sd = tvm.ConvNeXt_Tiny_Weights.DEFAULT.get_state_dict(progress=True)
    sd = {k: v for k, v in sd.items() if k.startswith("features.")}
    model.load_state_dict(sd, strict=True)


def _build(threshold: int, pretrained: bool, use_mid_layers: bool, num_stages: int) -> nn.Module:
    model = ConvNeXtTiny()
    if pretrained:
        _load_pretrained(model)
    with torch.no_grad():
        model = replace_convlayers_convnext(model, threshold)
        if use_mid_layers:
            model = MidLayerConvNeXt(model, num_stages=num_stages)
    return model


def convnext_tiny_26_features(pretrained=False, use_mid_layers=False, num_stages=2, **kwargs):
    """26x26 grid at 224^2 (features/convnext_features.py:38-65)."""
    return _build(100, pretrained, use_mid_layers, num_stages)


def convnext_tiny_13_features(pretrained=False, use_mid_layers=False, num_stages=2, **kwargs):
    """13x13 grid at 224^2 (features/convnext_features.py:67-94)."""
    return _build(300, pretrained, use_mid_layers, num_stages)


def get_feature_dimensions(use_mid_layers=False, num_stages=2, input_size=224):
    """Output shape for a configuration (features/convnext_features.py:97-107), computed from the layer geometry."""
    model = convnext_tiny_26_features(pretrained=False, use_mid_layers=use_mid_layers, num_stages=num_stages)
    c, h = 3, input_size
    for layer in model.features:
        for m in layer.modules():
            if isinstance(m, nn.Conv2d) and m.groups == 1:
                c = m.out_channels
                h = (h - m.kernel_size[0]) // m.stride[0] + 1
    return torch.Size([1, c, h, h])
