"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's per-prototype top-k search
loop, util/vis_pipnet.py:209-271, reduced to what it keeps per prototype (score, image index, arg-max patch).

Parity pinning: the list discipline below is the reference's (append while shorter than k, else replace the last entry
when the score is strictly greater, then a stable descending sort); the arg-max patch follows its two-step
`torch.max(feature_map, dim=0)` then `torch.max(max_h, dim=0)` (:234-238). The patch-coordinate function is not
restated here: tests/golden/search_coordinates.npz holds outputs of the reference's own get_img_coordinates
(tests/golden/make_golden_search.py executes its source from /root/reference)."""
from __future__ import annotations

from typing import Dict, List

import torch


def argmax_patch(feature_map: torch.Tensor):
    """(h_idx, w_idx) of the maximum of a [H, W] map the way util/vis_pipnet.py:234-238 finds it."""
    max_h, h_idx = torch.max(feature_map, dim=0)
    max_w, w_idx = torch.max(max_h, dim=0)
    return int(h_idx[w_idx]), int(w_idx)


def topk_search(pooled: torch.Tensor, proto_maps: torch.Tensor, k: int) -> Dict[int, List[dict]]:
    """pooled [N, P], proto_maps [N, P, H, W], images visited in index order, one at a time (vis_pipnet.py:209-271)."""
    n_img, n_proto = pooled.shape
    storage: Dict[int, List[dict]] = {p: [] for p in range(n_proto)}
    for i in range(n_img):
        for p in range(n_proto):
            score = float(pooled[i, p])
            images = storage[p]
            if len(images) < k or score > images[-1]["score"]:
                h_idx, w_idx = argmax_patch(proto_maps[i, p])
                entry = {"index": i, "score": score, "h_idx": h_idx, "w_idx": w_idx}
                if len(images) < k:
                    images.append(entry)
                else:
                    images[-1] = entry
                images.sort(key=lambda e: e["score"], reverse=True)
    return storage
