"""CPU, world size 2 over gloo: the batch split matches DataParallel's chunks and the gathered per-image outputs come
back in reference row order (even and ragged batches). The per-rank "forward" is the oracle head on a feature shard."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from count_pipnet_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, batch, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(5)
    feats = torch.randn(batch, 12, 3, 3, generator=g)
    w = torch.rand(7, 12, generator=g)
    mine = sharding.my_shard(feats)
    proto = torch.softmax(mine, dim=1)
    pooled = proto.amax(dim=(2, 3))
    out = pooled @ w.t()
    full_pooled, full_out = sharding.gather_outputs(pooled, out, batch)
    ref_pooled = torch.softmax(feats, dim=1).amax(dim=(2, 3))
    ok = torch.equal(full_pooled, ref_pooled) and torch.allclose(full_out, ref_pooled @ w.t())
    ret[rank] = bool(ok) and full_out.shape[0] == batch
    dist.destroy_process_group()


def test_shard_bounds_match_dataparallel_chunks():
    for batch, world in [(64, 8), (10, 4), (3, 8), (1, 2)]:
        chunks = torch.arange(batch).chunk(world)
        bounds = [b for b in sharding.shard_bounds(batch, world) if b[1] > b[0]]
        assert [(int(c[0]), int(c[-1]) + 1) for c in chunks] == bounds


def test_gather_two_ranks_gloo():
    for batch in (8, 5):
        with mp.Manager() as m:
            ret = m.dict()
            mp.spawn(_worker, args=(2, _free_port(), batch, ret), nprocs=2, join=True)
            assert ret[0] and ret[1], dict(ret)


def _worker_protomax(rank, world, port, batch, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(11)
    feats = torch.randn(batch, 12, 3, 4, generator=g)
    feats[batch - 1] = feats[0]  # the same image on two ranks: the tie must go to the lower image index
    mine = sharding.my_shard(feats)
    flat = torch.softmax(mine, dim=1).flatten(2)
    pooled, arg = flat.max(dim=2)
    score, image, token = sharding.gather_prototype_max(pooled, arg.int(), batch)
    ref_flat = torch.softmax(feats, dim=1).flatten(2)  # [B, P, HW]
    ref_pooled, ref_arg = ref_flat.max(dim=2)
    ref_score = ref_pooled.max(dim=0).values
    ref_image = torch.stack([(ref_pooled[:, p] == ref_score[p]).nonzero()[0, 0] for p in range(12)])
    ref_token = ref_arg.gather(0, ref_image[None, :])[0]
    ret[rank] = bool(torch.equal(score, ref_score) and torch.equal(image, ref_image) and torch.equal(token, ref_token))
    dist.destroy_process_group()


def test_prototype_max_gather_two_ranks_gloo():
    for batch in (8, 5):
        with mp.Manager() as m:
            ret = m.dict()
            mp.spawn(_worker_protomax, args=(2, _free_port(), batch, ret), nprocs=2, join=True)
            assert ret[0] and ret[1], dict(ret)
