"""ORACLE - test infrastructure only: a CPU restatement of the reference's forward hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this package; the
product (count_pipnet_b200/) never does. See oracle/reference_forward.py for how it is pinned to the real reference.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import reference_forward as rf
from . import reference_heads as rh


def backbone(sd: Dict[str, torch.Tensor], cfg: dict, x: torch.Tensor) -> torch.Tensor:
    """`net._net(x)`: NCHW fp32 features."""
    if "convnext" in cfg["net"]:
        # convnext_tiny_26 halves strides of convs with in_channels > 100, _13 of those > 300
        thr = 100 if cfg["net"].endswith("26") else 300
        return rf.convnext_features(sd, "_net.", x, thr)
    return rf.resnet_features(sd, "_net.", x)


def forward(sd: Dict[str, torch.Tensor], cfg: dict, x: torch.Tensor, inference: bool = True,
            gumbel_noise: Optional[torch.Tensor] = None, training: bool = False):
    """(proto_features, pooled, out) as the reference's forward(xs, inference) returns them.

    cfg keys: model ('pipnet' | 'count_pipnet'), net, activation ('softmax' | 'gumbel_softmax'), intermediate_layer,
    max_count, use_ste, tau. Eval mode unless `training` (only affects soft vs hard Gumbel)."""
    with torch.no_grad():
        feats = backbone(sd, cfg, x.float())
        act = cfg.get("activation", "softmax") if cfg["model"] == "count_pipnet" else "softmax"
        proto = rh.add_on(sd, feats, act, gumbel_noise, float(cfg.get("tau", 1.0)), hard=not training)
        if cfg["model"] == "pipnet":
            pooled, out = rh.pipnet_head(sd, proto, inference)
        else:
            pooled, out = rh.count_head(sd, proto, inference, cfg.get("intermediate_layer", "onehot"),
                                        int(cfg.get("max_count", 3)), bool(cfg.get("use_ste", True)))
    return proto, pooled, out
