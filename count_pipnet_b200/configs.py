"""The five workloads of BASELINE.json (`configs[0..4]`, SURVEY.md section 8 "cfg1".."cfg5") as data: factory arguments,
image size, per-GPU batch and algorithmic FLOPs per image. bench.py times them, tests/cases.py checks them at small
batch against the reference's golden outputs.

Reference sources of the model settings: configs/count_pipnet_shapes.yaml:1-49 (cfg1), configs/bilinear.yaml:1-48 (cfg4),
pipnet/pipnet.py:74-140 and features/resnet_features.py:199-210 (cfg2, cfg3, cfg5).
"""
from __future__ import annotations

import contextlib
import sys

from . import synth

# flop_per_image: 2 x MAC of the conv / linear layers (torch FlopCounterMode on the reference).
CONFIGS = {
    "cfg1": dict(
        label="Count-PIP-Net shapes config (configs/count_pipnet_shapes.yaml): ConvNeXt-tiny-26 truncated to 3 feature "
              "modules, 16 prototypes, 192x192, 9 classes, one-hot counts, hard Gumbel-softmax, STE",
        model="count_pipnet", classes=9, size=192, batch=64, flop_per_image=2.246e9, max_count=3, use_ste=True,
        materialize_map=False, prototypes=16, cpu_sample=64,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="gumbel_softmax", intermediate_layer="onehot")),
    "cfg2": dict(
        label="PIP-Net ConvNeXt-tiny-26 224x224, 768 prototypes, 200 classes, batch 64 per GPU (BASELINE.json configs[1])",
        model="pipnet", classes=200, size=224, batch=64, flop_per_image=40.094e9, materialize_map=False,
        prototypes=768, cpu_sample=16,
        args=dict(net="convnext_tiny_26", num_features=0)),
    "cfg3": dict(
        label="PIP-Net ResNet-50 224x224, 2048 prototypes, 196 classes, batch 256 over 8 GPUs = 32 per GPU "
              "(BASELINE.json configs[2])",
        model="pipnet", classes=196, size=224, batch=32, flop_per_image=38.155e9, materialize_map=False,
        prototypes=2048, cpu_sample=16,
        args=dict(net="resnet50", num_features=0)),
    "cfg4": dict(
        label="Count-PIP-Net bilinear count encoding (configs/bilinear.yaml), hard Gumbel-softmax + STE, batch 512 over "
              "2/4/8 GPUs (64 per GPU on one) (BASELINE.json configs[3])",
        model="count_pipnet", classes=9, size=192, batch=64, global_batch=512, flop_per_image=2.246e9, max_count=3,
        use_ste=True, materialize_map=False, prototypes=16, cpu_sample=64,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="gumbel_softmax", intermediate_layer="bilinear")),
    "cfg5": dict(
        label="PIP-Net ConvNeXt-tiny-26 448x448 (54x54 patch grid), full prototype map written for visualisation, batch "
              "128 over 8 GPUs = 16 per GPU (BASELINE.json configs[4])",
        model="pipnet", classes=200, size=448, batch=16, flop_per_image=169.343e9, materialize_map=True,
        prototypes=768, cpu_sample=4,
        args=dict(net="convnext_tiny_26", num_features=0)),
}


def batch_per_gpu(name: str, world: int) -> int:
    """cfg4 is defined by its GLOBAL batch (512 over 2/4/8 GPUs: strong scaling); the others by the per-GPU batch."""
    cfg = CONFIGS[name]
    if "global_batch" in cfg and world > 1:
        return max(1, cfg["global_batch"] // world)
    return cfg["batch"]


def build_model(name: str, seed: int = 0):
    """This package's drop-in module for a config, seeded synthetic weights (synth.py), eval mode, on the CPU."""
    from .pipnet import count_pipnet as my_count
    from .pipnet import pipnet as my_pipnet
    cfg = CONFIGS[name]
    args = synth.make_args(**cfg["args"])
    with contextlib.redirect_stdout(sys.stderr):  # the factories print configuration lines like the reference
        if cfg["model"] == "pipnet":
            model, _ = my_pipnet.get_pipnet(cfg["classes"], args)
        else:
            model, _ = my_count.get_count_network(cfg["classes"], args, max_count=cfg["max_count"],
                                                  use_ste=cfg["use_ste"], device=None)
    synth.load_synth(model, seed=seed)
    return model.eval()


def oracle_cfg(name: str) -> dict:
    cfg = CONFIGS[name]
    a = cfg["args"]
    return dict(model=cfg["model"], net=a["net"], activation=a.get("activation", "softmax"),
                intermediate_layer=a.get("intermediate_layer", "onehot"), max_count=cfg.get("max_count", 3),
                use_ste=cfg.get("use_ste", True), tau=1.0)
