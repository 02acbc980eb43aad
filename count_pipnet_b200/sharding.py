"""Batch sharding of the forward path across the GPUs of one box.

The reference's only parallelism is nn.DataParallel (main.py:118): contiguous batch chunks per device, outputs
concatenated in device order. Here every rank (one process per GPU) owns the same contiguous chunk DataParallel would
give it and runs the whole path on it; the one exchange step is the all-gather of the small per-image outputs
(logits [b, K], pooled scores or counts [b, P]). The patch x prototype map stays on the rank that produced it.
Works with any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(batch: int, world: int) -> List[Tuple[int, int]]:
    """[start, end) of every rank's chunk, the split torch.nn.parallel.scatter (tensor.chunk) makes: chunks of
    ceil(batch / world) images, the last ones shorter or empty."""
    per = -(-batch // world) if world > 0 else batch
    out = []
    for r in range(world):
        s = min(r * per, batch)
        out.append((s, min(s + per, batch)))
    return out


def my_shard(x: torch.Tensor, rank: Optional[int] = None, world: Optional[int] = None) -> torch.Tensor:
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    s, e = shard_bounds(x.shape[0], world)[rank]
    return x[s:e]


def gather_outputs(pooled: torch.Tensor, out: torch.Tensor, batch: int, group=None):
    """All-gather the per-image outputs of every rank into full-batch tensors (reference row order). Ragged shards
    are padded to the largest chunk for the collective and trimmed afterwards."""
    world = dist.get_world_size(group)
    bounds = shard_bounds(batch, world)
    per = max(e - s for s, e in bounds)

    even = all(e - s == per for s, e in bounds)

    def gather(t: torch.Tensor) -> torch.Tensor:
        if even and t.is_cuda and t.shape[0] == per and hasattr(dist, "all_gather_into_tensor"):
            # equal shards (the benchmark case): the collective writes the full-batch tensor in reference row order
            # directly -- no padding copy before it, no trimming copies after it
            full = t.new_empty((world * per,) + tuple(t.shape[1:]))
            dist.all_gather_into_tensor(full, t.contiguous(), group=group)
            return full
        pad = t.new_zeros((per,) + tuple(t.shape[1:]))
        pad[: t.shape[0]] = t
        full = t.new_empty((world * per,) + tuple(t.shape[1:]))
        dist.all_gather_into_tensor(full, pad, group=group) if hasattr(dist, "all_gather_into_tensor") and t.is_cuda \
            else dist.all_gather(list(full.split(per)), pad, group=group)
        return torch.cat([full[r * per: r * per + (e - s)] for r, (s, e) in enumerate(bounds)], dim=0)

    if even and pooled.is_cuda and pooled.dtype == out.dtype and pooled.shape[0] == per == out.shape[0] \
            and hasattr(dist, "all_gather_into_tensor"):
        # equal shards (the benchmark case): ONE collective for both outputs. The step is latency-bound (<= 0.3 MB per
        # rank), so a second all-gather costs a second launch + rank-skew wait; rows are packed as [pooled | logits] and
        # the results are column views of the gathered [batch, P + K] tensor (no trimming copies).
        p_cols = pooled.shape[1]
        packed = torch.cat([pooled, out], dim=1)
        full = packed.new_empty((world * per, packed.shape[1]))
        dist.all_gather_into_tensor(full, packed, group=group)
        return full[:, :p_cols], full[:, p_cols:]
    return gather(pooled), gather(out)


def gather_prototype_max(pooled: torch.Tensor, argmax: torch.Tensor, batch: int, group=None):
    """Per-prototype max gather: for every prototype the best-scoring image of the FULL batch, its
    score and the patch token of the maximum -- what the reference's top-k / visualisation search keeps per prototype
    (util/vis_pipnet.py:209-274) when the batch is spread over ranks.

    pooled [b, P] and argmax [b, P] are this rank's shard (rows in reference order). Returns (score [P], image [P]
    int64 index into the full batch, token [P] int64). Ties go to the lowest image index, as torch.max over the
    concatenated batch would. One all-gather of 2 * b * P values per rank; no map leaves its GPU."""
    full_pooled, full_arg = gather_outputs(pooled, argmax.to(pooled.dtype), batch, group=group)
    if full_pooled.shape[0] == 0:
        raise ValueError("gather_prototype_max: empty batch")
    score, image = full_pooled.max(dim=0)
    # torch.max returns the first maximal row on CPU and CUDA for a 2-D reduction only by convention; make it explicit
    is_max = full_pooled == score[None, :]
    idx = torch.arange(full_pooled.shape[0], device=pooled.device)[:, None].expand_as(full_pooled)
    image = torch.where(is_max, idx, torch.full_like(idx, full_pooled.shape[0])).min(dim=0).values
    token = full_arg.gather(0, image[None, :])[0].long()
    return score, image, token
