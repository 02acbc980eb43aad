"""Seeded synthetic weights and inputs: there are no datasets or checkpoints on the boxes, so
parity tests, golden fixtures and the benchmark all use this protocol. Values depend only on (state-dict key, shape,
seed), never on module construction order or the global RNG, so the reference model (built in the dev container) and
this package's model (built on the GPU box) get bit-identical weights.
"""
from __future__ import annotations

import argparse
import math
import zlib
from typing import Dict

import torch


def make_args(net: str = "convnext_tiny_26", num_features: int = 0, bias: bool = False, use_mid_layers: bool = False,
              num_stages: int = 2, activation: str = "softmax", intermediate_layer: str = "onehot",
              **extra) -> argparse.Namespace:
    """The fields the reference factories read (pipnet/pipnet.py:74-115, pipnet/count_pipnet.py:324-436)."""
    ns = argparse.Namespace(net=net, num_features=num_features, disable_pretrained=True, bias=bias,
                            use_mid_layers=use_mid_layers, num_stages=num_stages, activation=activation,
                            intermediate_layer=intermediate_layer, positive_grad_strategy=None,
                            backward_clamp_strategy="Identity")
    for k, v in extra.items():
        setattr(ns, k, v)
    return ns


def _gen(key: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def _uniform(shape, lo, hi, g):
    return torch.rand(shape, generator=g) * (hi - lo) + lo


def _normal(shape, std, g, mean=0.0):
    return torch.randn(shape, generator=g) * std + mean


def synth_state_dict(template: Dict[str, torch.Tensor], seed: int = 0, gain: float = 1.0,
                     classes_sparse: int = 8) -> Dict[str, torch.Tensor]:
    """A full state dict for `template` (a model.state_dict(); only keys/shapes are used)."""
    out = {}
    for key, ref in template.items():
        shape = tuple(ref.shape)
        g = _gen(key, seed)
        leaf = key.split(".")[-1]
        if leaf == "num_batches_tracked":
            v = torch.zeros(shape, dtype=ref.dtype)
        elif leaf == "running_mean":
            v = _normal(shape, 0.1, g)
        elif leaf == "running_var":
            v = _uniform(shape, 0.5, 1.5, g)
        elif leaf == "layer_scale":
            v = _uniform(shape, 0.2, 1.0, g)
        elif key.endswith("normalization_multiplier") or key == "_multiplier":
            v = torch.full(shape, 2.0)
        elif key.startswith("_classification.weight"):
            k, d = shape
            v = torch.zeros(shape)
            per = min(classes_sparse, d)
            for c in range(k):
                cols = torch.randperm(d, generator=g)[:per]
                v[c, cols] = _uniform((per,), 1.0, 11.0, g)
        elif key.startswith("_classification.bias"):
            v = _normal(shape, 0.1, g)
        elif key.startswith("_intermediate."):
            if key.endswith("embed.weight") or key.endswith("linear.weight"):
                v = None  # structured constructor init (count ramps); kept
            else:
                bound = 1.0 / math.sqrt(shape[1])
                v = _uniform(shape, -bound, bound, g)
        elif key.startswith("_add_on.") and leaf == "weight":
            fan_in, fan_out = shape[1], shape[0]
            bound = gain * math.sqrt(6.0 / (fan_in + fan_out))  # xavier-uniform (util/func.py:17-19)
            v = _uniform(shape, -bound, bound, g)
        elif key.startswith("_add_on.") and leaf == "bias":
            # one "background" prototype plus a graduated ramp, so counts spread over 0..max_count
            p = shape[0]
            v = torch.cat([torch.tensor([7.0]), torch.linspace(-1.5, 1.5, p - 1)]) if p > 1 else torch.zeros(shape)
        elif len(shape) >= 2:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            v = _normal(shape, gain / math.sqrt(fan_in), g)
        elif leaf == "weight":   # LayerNorm / BatchNorm scale
            v = _uniform(shape, 0.5, 1.5, g)
            # ResNet: damp the BN that closes a residual branch (and the projection shortcut) so that the summed
            # stream keeps |z|max ~ 10 at the softmax instead of growing block over block (measured: 53 -> 13)
            stem = key[: -len("weight")]
            closes_branch = key.endswith("bn3.weight") or \
                (key.endswith("bn2.weight") and (stem[:-4] + "bn3.weight") not in template)
            if "_net.layer" in key and (closes_branch or "downsample.1." in key):
                v = v * 0.8
        elif leaf == "bias":
            v = _normal(shape, 0.1, g)
        else:
            v = _normal(shape, 0.1, g)
        if v is not None:
            out[key] = v.to(ref.dtype)
        else:
            out[key] = ref.detach().clone()
    if "_multiplier" in out and "_classification.normalization_multiplier" in out:
        out["_multiplier"] = out["_classification.normalization_multiplier"]
    return out


def synth_images(batch: int, size: int, seed: int = 1234) -> torch.Tensor:
    """ImageNet-normalised images are roughly N(0,1) per channel."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return torch.randn(batch, 3, size, size, generator=g)


def load_synth(model: torch.nn.Module, seed: int = 0, gain: float = 1.0) -> None:
    sd = synth_state_dict(model.state_dict(), seed=seed, gain=gain)
    model.load_state_dict(sd, strict=True)
