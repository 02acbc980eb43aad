"""Count-PIP-Net behind the reference's API (pipnet/count_pipnet.py of TarasKutsyk/Count_PIPNet):
`get_count_network(num_classes, args, max_count=3, use_ste=True, device=None) -> (model, num_prototypes)`,
`forward(xs, inference=False) -> (proto_features, counts | clamped_counts, out)`, same attributes / parameter names,
and the star re-exports of count_pipnet_utils that reference callers import from here.

Eval-mode, no-grad, CUDA forwards run the fused sm_100a path (count_pipnet_b200.engine); with autograd enabled the
composed PyTorch modules below run instead.
"""
from __future__ import annotations

import argparse
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..features.convnext_features import convnext_tiny_13_features, convnext_tiny_26_features
from .count_pipnet_utils import *  # noqa: F401,F403  (re-exported like the reference, count_pipnet.py:12)
from .count_pipnet_utils import (BilinearIntermediate, ClampSTE, GumbelSoftmax, IdentityIntermediate, LinearFull,
                                 LinearIntermediate, OneHotEncoder, STE_Round)


class CountPIPNet(nn.Module):
    """pipnet/count_pipnet.py:14-168."""

    def __init__(self, num_classes: int, num_prototypes: int, feature_net: nn.Module, args: argparse.Namespace,
                 add_on_layers: nn.Module, intermediate_layer: nn.Module, classification_layer: nn.Module,
                 max_count: int = 3, use_ste: bool = True, backward_clamp_strategy: str = 'Identity'):
        super().__init__()
        assert num_classes > 0
        self._num_features = args.num_features
        self._num_classes = num_classes
        self._num_prototypes = num_prototypes
        self._net = feature_net
        self._add_on = add_on_layers
        self._classification = classification_layer
        self._intermediate = intermediate_layer
        assert backward_clamp_strategy in ['Identity', 'Gated']
        print(f"Using backward clamp strategy: {backward_clamp_strategy}", flush=True)
        self._is_clamp_backward_identity = (backward_clamp_strategy == 'Identity')
        self._max_count = max_count
        self._use_ste = use_ste
        self._multiplier = classification_layer.normalization_multiplier
        self.ste_round = STE_Round.apply
        self.ste_clamp = ClampSTE.apply
        self.materialize_proto_map = True  # see PIPNet.materialize_proto_map

    def forward(self, xs, inference=False):
        from .. import engine
        if engine.use_fused(self, xs):
            return engine.network_forward(self, xs, inference)
        features = self._net(xs)
        if engine.use_train_head(self, features, want_max=False):
            # autograd forward on a B200 with the plain-softmax activation: softmax + spatial sum as one kernel
            from .train import softmax_pool_head
            z = features
            for layer in list(self._add_on)[:-1]:
                z = layer(z)
            proto_features, counts = softmax_pool_head(z, "sum")
        else:
            proto_features = self._add_on(features)
            counts = self._calculate_counts_for_testing(proto_features)
        if self._use_ste:
            rounded = self.ste_round(counts)
            clamped = self.ste_clamp(rounded, 0, self._max_count, self._is_clamp_backward_identity)
        else:
            rounded = counts.round() if inference else counts
            clamped = torch.clamp(rounded, 0, self._max_count)
        out = self._classification(self._intermediate(clamped))
        # inference returns the clamped counts (interpretability), training the raw sums (count_pipnet.py:107-110)
        return proto_features, (clamped if inference else counts), out

    def _calculate_counts_for_testing(self, proto_features):
        """Spatial sum of the prototype map (count_pipnet.py:112-124)."""
        return proto_features.sum(dim=(2, 3))

    def get_prototype_importance_per_class(self, prototype_idx, classifier_input_scalars=None):
        """|intermediate weights of the prototype| . classifier weights, per class (count_pipnet.py:126-147)."""
        w = self._intermediate.prototype_to_classifier_input_weights(prototype_idx)
        if classifier_input_scalars is not None:
            assert classifier_input_scalars.shape == w.shape, \
                "Classifier input scalars must have the same shape as the classifier input weights" + \
                "\nclassifier_input_weights.shape: " + str(w.shape) + \
                "\nclassifier_input_scalars.shape: " + str(classifier_input_scalars.shape)
            w = w * classifier_input_scalars
        return self._classification.weight @ torch.abs(w)

    def get_prototype_importance(self, prototype_idx):
        return self.get_prototype_importance_per_class(prototype_idx).sum().item()

    def update_temperature(self, new_temperature):
        """Set tau of the first GumbelSoftmax in the add-on (count_pipnet.py:156-168)."""
        for module in self._add_on.modules():
            if isinstance(module, GumbelSoftmax):
                module.tau = new_temperature
                break


base_architecture_to_features = {
    'convnext_tiny_26': convnext_tiny_26_features,
    'convnext_tiny_13': convnext_tiny_13_features,
}


class NonNegLinear(nn.Module):
    """ReLU-weight linear layer, Kaiming-uniform initialised (count_pipnet.py:176-224)."""

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
        self.reset_parameters()

    def reset_parameters(self) -> None:
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.in_features)
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return F.linear(input, torch.relu(self.weight), self.bias)


def estimate_mean_intermediate_features(net, dataloader, device, return_full_data=False):
    """Mean intermediate-layer activation over a dataset (count_pipnet.py:226-281). `net` is the wrapped model
    (`net.module` is the CountPIPNet), as in the reference."""
    counts_all, labels_all = [], []
    with torch.no_grad():
        for xs, ys in dataloader:
            try:
                _, clamped, _ = net(xs.to(device), inference=True)
            except Exception as e:  # the reference skips failing batches too
                print(f"\nError during forward pass: {e}. Skipping batch.")
                continue
            counts_all.append(clamped)
            labels_all.append(ys)
    counts_all = torch.cat(counts_all, dim=0)
    labels_all = torch.cat(labels_all, dim=0)
    print("all_clamped_counts.shape: ", counts_all.shape)
    print("all_class_labels.shape: ", labels_all.shape)
    inter = net.module._intermediate(counts_all)
    print("intermediate_features.shape: ", inter.shape)
    if return_full_data:
        return inter, labels_all
    return inter.mean(dim=0)


def calculate_virtual_weights(net, dataloader, device, custom_onehot_scale=False):
    """[num_classes, num_prototypes] importance matrix built from get_prototype_importance_per_class
    (count_pipnet.py:283-321; the reference's optional matplotlib plot of the mean features is omitted)."""
    model = net.module
    scalars = None
    if isinstance(model._intermediate, OneHotEncoder) and custom_onehot_scale:
        print("Intermediate is onehot, computing mean intermediate features...")
        scalars = estimate_mean_intermediate_features(net, dataloader, device)
    weights = torch.zeros((model._num_classes, model._num_prototypes), device=device)
    for i in range(model._num_prototypes):
        weights[:, i] = model.get_prototype_importance_per_class(i, scalars).to(device)
    return weights


def get_count_network(num_classes: int, args: argparse.Namespace, max_count: int = 3, use_ste: bool = True,
                      device=None):
    """pipnet/count_pipnet.py:324-436 (same `getattr` defaults)."""
    if args.net not in base_architecture_to_features:
        supported = list(base_architecture_to_features.keys())
        raise ValueError(f"Network '{args.net}' is not supported. Supported networks: {supported}")
    features = base_architecture_to_features[args.net](
        pretrained=not args.disable_pretrained,
        use_mid_layers=getattr(args, 'use_mid_layers', False),
        num_stages=getattr(args, 'num_stages', 2))
    in_channels = detect_output_channels(features)

    if getattr(args, 'activation', 'gumbel_softmax') == 'softmax':
        activation = nn.Softmax(dim=1)
    else:
        activation = GumbelSoftmax(dim=1, tau=1.0)

    if args.num_features == 0:
        num_prototypes = in_channels
        print(f"Number of prototypes: {num_prototypes}", flush=True)
        add_on = nn.Sequential(activation)
    else:
        num_prototypes = args.num_features
        print(f"Number of prototypes set from {in_channels} to {num_prototypes}. Extra 1x1 conv layer added.",
              flush=True)
        add_on = nn.Sequential(
            nn.Conv2d(in_channels=in_channels, out_channels=num_prototypes, kernel_size=1, stride=1, padding=0,
                      bias=True),
            activation)

    kind = getattr(args, 'intermediate_layer', 'onehot')
    positive_grad_strategy = getattr(args, 'positive_grad_strategy', None)
    print(f"Using positive gradient strategy: {positive_grad_strategy}", flush=True)
    backward_clamp_strategy = getattr(args, 'backward_clamp_strategy', 'Gated')
    print(f"Using backward clamp strategy: {backward_clamp_strategy}", flush=True)

    expanded_dim = num_prototypes * max_count
    if kind == 'linear':
        intermediate = LinearIntermediate(num_prototypes, max_count)
    elif kind == 'linear_full':
        intermediate = LinearFull(num_prototypes, max_count)
    elif kind == 'bilinear':
        intermediate = BilinearIntermediate(num_prototypes, max_count)
    elif kind == 'onehot':
        intermediate = OneHotEncoder(max_count, use_ste=use_ste, respect_active_grad=False,
                                     num_prototypes=num_prototypes, device=device,
                                     positive_grad_strategy=positive_grad_strategy)
    elif kind == 'identity':
        intermediate = IdentityIntermediate(num_prototypes, device=device)
        expanded_dim = num_prototypes
    else:
        raise ValueError(f"Unknown intermediate layer type: {kind}")

    classifier = NonNegLinear(expanded_dim, num_classes, bias=getattr(args, 'bias', False))
    model = CountPIPNet(num_classes=num_classes, num_prototypes=num_prototypes, feature_net=features, args=args,
                        add_on_layers=add_on, classification_layer=classifier, intermediate_layer=intermediate,
                        max_count=max_count, use_ste=use_ste, backward_clamp_strategy=backward_clamp_strategy)
    return model, num_prototypes


def detect_output_channels(features):
    """out_channels of the last Conv2d inside the last feature module (count_pipnet.py:438-464)."""
    feats = getattr(features, 'features', None)
    if feats is not None and len(feats) > 0:
        convs = [m for m in feats[-1].modules() if isinstance(m, nn.Conv2d)]
        if convs:
            channels = convs[-1].out_channels
            print(f"Detected {channels} output channels from last conv layer", flush=True)
            return channels
    raise RuntimeError("Could not detect output channels from the feature extractor.")
