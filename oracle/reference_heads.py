"""ORACLE - test infrastructure only (see oracle/reference_forward.py for the rules).

CPU restatement of the prototype heads: PIPNet.forward (reference pipnet/pipnet.py:31-41) and CountPIPNet.forward
(pipnet/count_pipnet.py:70-110) from the backbone features on. Count encodings are integer/index work and are written
in numpy; softmax / linear layers are torch fp32 on the CPU.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


def add_on(sd: SD, features, activation: str, gumbel_noise: Optional[torch.Tensor], tau: float, hard: bool):
    """`_add_on`: optional Conv2d 1x1 (pipnet/pipnet.py:102, count_pipnet.py:378-379) then Softmax(dim=1)
    (pipnet.py:96,103) or F.gumbel_softmax (count_pipnet_utils.py:33-38) with the noise passed in:
    y_soft = softmax((z + g) / tau); hard: one_hot(argmax) - y_soft + y_soft."""
    z = features
    if "_add_on.0.weight" in sd:
        z = F.conv2d(z, sd["_add_on.0.weight"], sd.get("_add_on.0.bias"))
    if activation == "softmax":
        return torch.softmax(z, dim=1)
    assert gumbel_noise is not None, "gumbel_softmax needs the noise tensor the run used"
    y_soft = torch.softmax((z + gumbel_noise) / tau, dim=1)
    if not hard:
        return y_soft
    index = y_soft.max(1, keepdim=True)[1]
    y_hard = torch.zeros_like(z).scatter_(1, index, 1.0)
    return y_hard - y_soft + y_soft


def non_neg_linear(sd: SD, x):
    """F.linear(x, relu(W), b) (pipnet/pipnet.py:70-71, count_pipnet.py:213-224)."""
    return F.linear(x, torch.relu(sd["_classification.weight"]), sd.get("_classification.bias"))


def pipnet_head(sd: SD, proto_features, inference: bool):
    """AdaptiveMaxPool2d((1,1)) + Flatten, inference threshold, classifier (pipnet/pipnet.py:33-41,105-108)."""
    pooled = proto_features.amax(dim=(2, 3))
    if inference:
        pooled = torch.where(pooled < 0.1, torch.zeros_like(pooled), pooled)
    return pooled, non_neg_linear(sd, pooled)


def onehot_encoding(x: np.ndarray, max_count: int) -> np.ndarray:
    """create_modified_encoding (count_pipnet_utils.py:141-185): mask x > 0.1, index clamp(int(x) - 1, 0, M - 1),
    fp32 output [B, P, M]."""
    b, p = x.shape
    enc = np.zeros((b, p, max_count), dtype=np.float32)
    mask = x > 0.1
    idx = np.clip(np.trunc(x).astype(np.int64) - 1, 0, max_count - 1)
    bi, pi = np.nonzero(mask)
    enc[bi, pi, idx[bi, pi]] = 1.0
    return enc


def intermediate(sd: SD, kind: str, clamped, max_count: int, use_ste: bool):
    """`_intermediate(clamped_counts)` for every encoder of count_pipnet_utils.py."""
    if kind == "identity":        # :461-462
        return clamped
    if kind == "onehot":          # :108-124, :195-217 (STE variant rounds first)
        x = clamped.numpy()
        if use_ste:
            x = np.rint(x)        # torch.round == round-half-to-even
        return torch.from_numpy(onehot_encoding(x, max_count).reshape(x.shape[0], -1))
    if kind == "linear":          # :498-519  out[b, p*E + i] = w[i] * x[b, p]
        w = sd["_intermediate.linear.weight"]  # [E, 1]
        b, p = clamped.shape
        return F.linear(clamped.reshape(b * p, 1), w).reshape(b, p * w.shape[0])
    if kind == "linear_full":     # :429-439
        return F.linear(clamped, sd["_intermediate.linear.weight"])
    if kind == "bilinear":        # :371-385
        e = F.linear(clamped, sd["_intermediate.embed.weight"])
        return F.linear(e, sd["_intermediate.W.weight"]) * F.linear(e, sd["_intermediate.V.weight"])
    raise ValueError(kind)


def count_head(sd: SD, proto_features, inference: bool, kind: str, max_count: int, use_ste: bool):
    """sum over (H,W) -> round (STE or inference) -> clamp [0, max_count] -> intermediate -> classifier
    (pipnet/count_pipnet.py:88-110). Second return value: clamped counts at inference, raw counts otherwise."""
    counts = proto_features.sum(dim=(2, 3))
    if use_ste or inference:
        rounded = counts.round()
    else:
        rounded = counts
    clamped = torch.clamp(rounded, 0, max_count)
    out = non_neg_linear(sd, intermediate(sd, kind, clamped, max_count, use_ste))
    return (clamped if inference else counts), out
