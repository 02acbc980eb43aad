"""Prototype top-k search on the GPU: the part of the reference's visualisation pass that
runs the network (util/vis_pipnet.py:209-271), batched.

The reference walks the projection set one image per forward and, per prototype, reads the score and the arg-max patch
off the full map on the host (`.item()`, `.cpu()` per prototype). Here a forward of B images returns pooled [B, P] and
the arg-max tokens [B, P] from the pooling kernel (no map written), and `cpn_topk_update` merges them into running
per-prototype lists on the device with the reference's list semantics (strictly-greater replacement of the last entry,
descending order, equal scores in arrival order). `get_img_coordinates` / `patch_coordinates` are the reference's
patch-coordinate math (util/vis_pipnet.py:1162-1193). Image loading, cropping and plotting stay with the caller.
"""
from __future__ import annotations

from typing import Iterable, Tuple

import torch

from . import ops


def get_img_coordinates(img_size, softmaxes_shape, patchsize, skip, h_idx, w_idx):
    """Same signature and results as util/vis_pipnet.py:1162-1193 (softmaxes_shape = (P, H, W) or (B, P, H, W) as the
    reference passes it; only [1], [2] and [-1] are read)."""
    if softmaxes_shape[1] == 26 and softmaxes_shape[2] == 26:
        # 26 x 26 latent grid (ConvNeXt with the smaller strides): outer patches have a smaller receptive field
        h_coor_min = max(0, (h_idx - 1) * skip + 4)
        if h_idx >= softmaxes_shape[-1] - 1:
            h_coor_min -= 4
        h_coor_max = h_coor_min + patchsize
        w_coor_min = max(0, (w_idx - 1) * skip + 4)
        if w_idx >= softmaxes_shape[-1] - 1:
            w_coor_min -= 4
        w_coor_max = w_coor_min + patchsize
    else:
        h_coor_min = h_idx * skip
        h_coor_max = min(img_size, h_idx * skip + patchsize)
        w_coor_min = w_idx * skip
        w_coor_max = min(img_size, w_idx * skip + patchsize)
    if h_idx == softmaxes_shape[1] - 1:
        h_coor_max = img_size
    if w_idx == softmaxes_shape[2] - 1:
        w_coor_max = img_size
    if h_coor_max == img_size:
        h_coor_min = img_size - patchsize
    if w_coor_max == img_size:
        w_coor_min = img_size - patchsize
    return h_coor_min, h_coor_max, w_coor_min, w_coor_max


def patch_coordinates(img_size: int, softmaxes_shape, patchsize: int, skip: int, tokens: torch.Tensor) -> torch.Tensor:
    """get_img_coordinates for a whole tensor of patch tokens (h * W + w) at once, on the tokens' device.
    Returns [..., 4] int64 = (h_min, h_max, w_min, w_max)."""
    H, W, last = int(softmaxes_shape[1]), int(softmaxes_shape[2]), int(softmaxes_shape[-1])
    t = tokens.long()
    h, w = t // W, t % W

    def axis(idx: torch.Tensor, n: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if H == 26 and W == 26:
            lo = ((idx - 1) * skip + 4).clamp_min(0)
            lo = torch.where(idx >= last - 1, lo - 4, lo)
            hi = lo + patchsize
        else:
            lo = idx * skip
            hi = (idx * skip + patchsize).clamp_max(img_size)
        hi = torch.where(idx == n - 1, torch.full_like(hi, img_size), hi)
        lo = torch.where(hi == img_size, torch.full_like(lo, img_size - patchsize), lo)
        return lo, hi

    h_lo, h_hi = axis(h, H)
    w_lo, w_hi = axis(w, W)
    return torch.stack([h_lo, h_hi, w_lo, w_hi], dim=-1)


class PrototypeTopK:
    """Running top-k images per prototype, kept on the device. Lists are [k, P]: score (descending), image index into
    the stream of update() calls (-1 = empty slot) and patch token h * W + w."""

    def __init__(self, num_prototypes: int, k: int, device):
        self.k, self.P = int(k), int(num_prototypes)
        self.score = torch.zeros(self.k, self.P, device=device)
        self.image = torch.full((self.k, self.P), -1, dtype=torch.int64, device=device)
        self.token = torch.zeros(self.k, self.P, dtype=torch.int32, device=device)
        self.count = torch.zeros(self.P, dtype=torch.int32, device=device)
        self.seen = 0

    def update(self, pooled: torch.Tensor, argmax: torch.Tensor, image_base: int | None = None) -> None:
        """pooled [B, P] fp32 and argmax [B, P] int32 of images image_base .. image_base + B - 1 (default: continue
        after the images seen so far), in the order the reference's loop would visit them."""
        base = self.seen if image_base is None else int(image_base)
        ops.topk_update(pooled.contiguous(), argmax.contiguous(), base, self.score, self.image, self.token, self.count)
        self.seen = max(self.seen, base + pooled.shape[0])

    def result(self):
        return {"score": self.score, "image": self.image, "token": self.token, "count": self.count}


def search(net, batches: Iterable[torch.Tensor], k: int = 10) -> Tuple[PrototypeTopK, Tuple[int, int]]:
    """Runs `net(xs, inference=True)` (PIP-Net, eval, fused path) over the batches and keeps the top-k images per
    prototype. Returns the lists and the patch grid (H, W) the tokens refer to."""
    keep = (getattr(net, "track_prototype_argmax", False), getattr(net, "materialize_proto_map", True))
    net.track_prototype_argmax, net.materialize_proto_map = True, False
    top, grid = None, None
    try:
        with torch.no_grad():
            for xs in batches:
                _, pooled, _ = net(xs, inference=True)
                if top is None:
                    top = PrototypeTopK(pooled.shape[1], k, pooled.device)
                    grid = tuple(net.cpn_patch_grid)
                top.update(pooled, net.cpn_prototype_argmax)
    finally:
        net.track_prototype_argmax, net.materialize_proto_map = keep
    return top, grid
