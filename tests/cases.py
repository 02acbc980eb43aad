"""Parity cases shared by the golden generator (run against the real reference in the dev container), the oracle
tests (CPU) and the GPU parity tests. Sizes follow BASELINE.json's configs at batch sizes the CPU finishes in seconds."""
from __future__ import annotations

import torch

from count_pipnet_b200 import synth

# name -> dict(model, factory args, classes, image size, batch, inference, ...)
CASES = {
    # configs[0]/cfg1: configs/count_pipnet_shapes.yaml - 3 feature modules, P=16, 192^2, onehot, hard gumbel, STE
    "cfg1_count_shapes_onehot_gumbel": dict(
        model="count_pipnet", classes=9, size=192, batch=4, inference=True, max_count=3, use_ste=True,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="gumbel_softmax", intermediate_layer="onehot")),
    # configs[1]/cfg2: PIP-Net ConvNeXt-tiny 224^2, 768 prototypes, 200 classes
    "cfg2_pipnet_convnext26_224": dict(
        model="pipnet", classes=200, size=224, batch=2, inference=True,
        args=dict(net="convnext_tiny_26", num_features=0)),
    # configs[2]/cfg3: PIP-Net ResNet-50 224^2, 2048 prototypes, 196 classes
    "cfg3_pipnet_resnet50_224": dict(
        model="pipnet", classes=196, size=224, batch=2, inference=True,
        args=dict(net="resnet50", num_features=0)),
    # configs[3]/cfg4: configs/bilinear.yaml
    "cfg4_count_bilinear_gumbel": dict(
        model="count_pipnet", classes=9, size=192, batch=4, inference=True, max_count=3, use_ste=True,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="gumbel_softmax", intermediate_layer="bilinear")),
    # configs[4]/cfg5: 448^2 -> 54x54 grid, full map
    "cfg5_pipnet_convnext26_448_map": dict(
        model="pipnet", classes=200, size=448, batch=1, inference=True,
        args=dict(net="convnext_tiny_26", num_features=0)),
    # further head variants on the truncated backbone
    "count_linear_softmax": dict(
        model="count_pipnet", classes=9, size=96, batch=3, inference=True, max_count=3, use_ste=False,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="softmax", intermediate_layer="linear")),
    "count_linear_full_softmax_train_outputs": dict(
        model="count_pipnet", classes=9, size=96, batch=3, inference=False, max_count=3, use_ste=True,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=3,
                  activation="softmax", intermediate_layer="linear_full")),
    "count_identity_softmax_noste": dict(
        model="count_pipnet", classes=9, size=96, batch=3, inference=False, max_count=4, use_ste=False,
        args=dict(net="convnext_tiny_26", num_features=16, use_mid_layers=True, num_stages=2,
                  activation="softmax", intermediate_layer="identity")),
    "count_onehot_softmax_noste_p192": dict(
        model="count_pipnet", classes=9, size=96, batch=2, inference=True, max_count=3, use_ste=False,
        args=dict(net="convnext_tiny_26", num_features=0, use_mid_layers=True, num_stages=3,
                  activation="softmax", intermediate_layer="onehot")),
    "pipnet_convnext13_addon32_bias": dict(
        model="pipnet", classes=20, size=160, batch=2, inference=True,
        args=dict(net="convnext_tiny_13", num_features=32, bias=True)),
    "pipnet_convnext26_mid5_noinference": dict(
        model="pipnet", classes=20, size=128, batch=2, inference=False,
        args=dict(net="convnext_tiny_26", num_features=0, use_mid_layers=True, num_stages=5)),
    "pipnet_resnet18_64": dict(
        model="pipnet", classes=10, size=64, batch=2, inference=True,
        args=dict(net="resnet18", num_features=0)),
}

WEIGHT_SEED = 0
INPUT_SEED = 1234
GUMBEL_SEED = 7


def oracle_cfg(case: dict) -> dict:
    a = case["args"]
    return dict(model=case["model"], net=a["net"], activation=a.get("activation", "softmax"),
                intermediate_layer=a.get("intermediate_layer", "onehot"), max_count=case.get("max_count", 3),
                use_ste=case.get("use_ste", True), tau=1.0)


def inputs(case: dict) -> torch.Tensor:
    return synth.synth_images(case["batch"], case["size"], seed=INPUT_SEED)


def gumbel_noise(shape) -> torch.Tensor:
    """The draw F.gumbel_softmax makes first thing after torch.manual_seed(GUMBEL_SEED) on the CPU."""
    g = torch.Generator(device="cpu")
    g.manual_seed(GUMBEL_SEED)
    return -torch.empty(shape).exponential_(generator=g).log()


def build(case: dict, pipnet_mod, count_mod):
    """Build a model with either the reference's modules or this package's (same factory API)."""
    args = synth.make_args(**case["args"])
    if case["model"] == "pipnet":
        model, _ = pipnet_mod.get_pipnet(case["classes"], args)
    else:
        model, _ = count_mod.get_count_network(case["classes"], args, max_count=case["max_count"],
                                               use_ste=case["use_ste"], device=None)
    synth.load_synth(model, seed=WEIGHT_SEED)
    return model.eval()
