"""PIP-Net behind the reference's API (pipnet/pipnet.py of TarasKutsyk/Count_PIPNet):
`get_pipnet(num_classes, args) -> (model, num_prototypes)`, `forward(xs, inference=False) ->
(proto_features, pooled, out)`, identical attribute and parameter names.

In eval mode without autograd, on a CUDA tensor, `forward` runs the fused sm_100a path of
`count_pipnet_b200.engine` (no CPU fallback: it raises if the library or a B200 is missing). With autograd enabled
(training, attribution) the same submodules run composed, so gradients exist; that path is plain PyTorch.
"""
from __future__ import annotations

import argparse

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from ..features.convnext_features import convnext_tiny_13_features, convnext_tiny_26_features
from ..features.resnet_features import (resnet18_features, resnet34_features, resnet50_features,
                                        resnet50_features_inat, resnet101_features, resnet152_features)


class PIPNet(nn.Module):
    """pipnet/pipnet.py:10-41."""

    def __init__(self, num_classes: int, num_prototypes: int, feature_net: nn.Module, args: argparse.Namespace,
                 add_on_layers: nn.Module, pool_layer: nn.Module, classification_layer: nn.Module):
        super().__init__()
        assert num_classes > 0
        self._num_features = args.num_features
        self._num_classes = num_classes
        self._num_prototypes = num_prototypes
        self._net = feature_net
        self._add_on = add_on_layers
        self._pool = pool_layer
        self._classification = classification_layer
        self._multiplier = classification_layer.normalization_multiplier
        # B200 path option: when False the fused forward skips writing the [B,P,H,W] map and returns None for it
        self.materialize_proto_map = True

    def forward(self, xs, inference=False):
        from .. import engine
        if engine.use_fused(self, xs):
            return engine.network_forward(self, xs, inference)
        features = self._net(xs)
        if engine.use_train_head(self, features, want_max=True):
            # autograd forward on a B200: softmax + max pool (and their backward) as one kernel each
            from .train import softmax_pool_head
            z = features
            for layer in list(self._add_on)[:-1]:  # the optional 1x1 add-on conv stays a torch module
                z = layer(z)
            proto_features, pooled = softmax_pool_head(z, "max")
        else:
            proto_features = self._add_on(features)
            pooled = self._pool(proto_features)
        if inference:
            # prototypes with similarity below 0.1 are ignored at inference (pipnet/pipnet.py:36)
            pooled = torch.where(pooled < 0.1, 0., pooled)
        out = self._classification(pooled)
        return proto_features, pooled, out


base_architecture_to_features = {'resnet18': resnet18_features,
                                 'resnet34': resnet34_features,
                                 'resnet50': resnet50_features,
                                 'resnet50_inat': resnet50_features_inat,
                                 'resnet101': resnet101_features,
                                 'resnet152': resnet152_features,
                                 'convnext_tiny_26': convnext_tiny_26_features,
                                 'convnext_tiny_13': convnext_tiny_13_features}


class NonNegLinear(nn.Module):
    """Linear layer whose weights pass through ReLU in forward (pipnet/pipnet.py:54-71). The weight is left
    uninitialised here exactly as in the reference; callers initialise it (main.py:168)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device=None, dtype=None) -> None:
        kw = {'device': device, 'dtype': dtype}
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.weight = nn.Parameter(torch.empty((out_features, in_features), **kw))
        self.normalization_multiplier = nn.Parameter(torch.ones((1,), requires_grad=True))
        if bias:
            self.bias = nn.Parameter(torch.empty(out_features, **kw))
        else:
            self.register_parameter('bias', None)

    def forward(self, input: Tensor) -> Tensor:
        return F.linear(input, torch.relu(self.weight), self.bias)


def get_pip_network(num_classes: int, args: argparse.Namespace):
    """pipnet/pipnet.py:74-115."""
    if 'convnext' in args.net:
        backbone = base_architecture_to_features[args.net](
            pretrained=not args.disable_pretrained,
            use_mid_layers=getattr(args, 'use_mid_layers', False),
            num_stages=getattr(args, 'num_stages', 2))
    elif 'res' in args.net:
        backbone = base_architecture_to_features[args.net](pretrained=not args.disable_pretrained)
    else:
        raise Exception('other base architecture NOT implemented')

    in_channels = [m for m in backbone.modules() if isinstance(m, nn.Conv2d)][-1].out_channels
    if args.num_features == 0:
        num_prototypes = in_channels
        print("Number of prototypes: ", num_prototypes, flush=True)
        add_on = nn.Sequential(nn.Softmax(dim=1))
    else:
        num_prototypes = args.num_features
        print("Number of prototypes set from", in_channels, "to", num_prototypes,
              ". Extra 1x1 conv layer added. Not recommended.", flush=True)
        add_on = nn.Sequential(
            nn.Conv2d(in_channels=in_channels, out_channels=num_prototypes, kernel_size=1, stride=1, padding=0,
                      bias=True),
            nn.Softmax(dim=1))
    pool = nn.Sequential(nn.AdaptiveMaxPool2d(output_size=(1, 1)), nn.Flatten())
    classifier = NonNegLinear(num_prototypes, num_classes, bias=bool(args.bias))
    return backbone, add_on, pool, classifier, num_prototypes


def get_pipnet(num_classes: int, args: argparse.Namespace):
    """pipnet/pipnet.py:117-140."""
    backbone, add_on, pool, classifier, num_prototypes = get_pip_network(num_classes, args)
    model = PIPNet(num_classes=num_classes, num_prototypes=num_prototypes, feature_net=backbone, args=args,
                   add_on_layers=add_on, pool_layer=pool, classification_layer=classifier)
    return model, num_prototypes
