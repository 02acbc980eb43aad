"""Prototype top-k search. CPU: the patch-coordinate math against outputs of the reference's own
function (tests/golden/search_coordinates.npz) and the oracle's list discipline on a hand-checked case. GPU: the
device-side running top-k against the oracle loop, entry for entry (scores bit-exact, image indices and patches exact),
including ties, lists that never fill and batch boundaries; and the end-to-end search through the fused forward."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import reference_search
from count_pipnet_b200 import proto_search

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_img_coordinates_match_reference_function():
    table = np.load(os.path.join(GOLDEN, "search_coordinates.npz"))["table"]
    assert len(table) > 30000
    for row in table[:: 7]:
        img_size, ndim, p, H, W, patch, skip, h, w = (int(v) for v in row[:9])
        shape = (p, H, W) if ndim == 3 else (1, p, H, W)
        assert proto_search.get_img_coordinates(img_size, shape, patch, skip, h, w) == tuple(int(v) for v in row[9:])


def test_patch_coordinates_vectorised_match_reference_function():
    table = np.load(os.path.join(GOLDEN, "search_coordinates.npz"))["table"]
    keys = np.unique(table[:, :7], axis=0)
    for key in keys:
        rows = table[(table[:, :7] == key).all(axis=1)]
        img_size, ndim, p, H, W, patch, skip = (int(v) for v in key)
        shape = (p, H, W) if ndim == 3 else (1, p, H, W)
        if ndim == 4:
            continue  # the vectorised form takes the grid from shape[1], shape[2] exactly like the scalar one; 3-D rows pin it
        tokens = torch.from_numpy(rows[:, 7] * W + rows[:, 8])
        got = proto_search.patch_coordinates(img_size, shape, patch, skip, tokens).numpy()
        assert np.array_equal(got, rows[:, 9:])


def test_oracle_list_discipline():
    pooled = torch.tensor([[0.5], [0.7], [0.7], [0.2], [0.9], [0.5]])
    maps = torch.zeros(6, 1, 2, 3)
    for i in range(6):
        maps[i, 0, i % 2, i % 3] = pooled[i, 0]
    got = reference_search.topk_search(pooled, maps, k=3)[0]
    assert [e["index"] for e in got] == [4, 1, 2]          # equal scores keep arrival order, 0.5 was pushed out
    assert [(e["h_idx"], e["w_idx"]) for e in got] == [(0, 1), (1, 1), (0, 2)]


def _lists_from_oracle(storage, k, P, W):
    score = np.zeros((k, P), np.float32)
    image = np.full((k, P), -1, np.int64)
    token = np.zeros((k, P), np.int32)
    for p, entries in storage.items():
        for i, e in enumerate(entries):
            score[i, p], image[i, p], token[i, p] = e["score"], e["index"], e["h_idx"] * W + e["w_idx"]
    return score, image, token


@pytest.mark.gpu
@pytest.mark.parametrize("N,P,k,batch", [(40, 96, 5, 7), (9, 33, 10, 4), (64, 768, 10, 16), (5, 16, 1, 5)])
def test_topk_update_matches_oracle(cuda_device, N, P, k, batch):
    g = torch.Generator().manual_seed(N + P)
    H, W = 5, 6
    maps = torch.rand(N, P, H, W, generator=g)
    maps = (maps * 16).round() / 16           # many exactly equal scores: the tie rules decide
    maps[:, 0] = 0.0                          # a prototype that never fires: zeros everywhere, token 0
    flat = maps.flatten(2)
    pooled, arg = flat.max(dim=2)
    # the first maximal token in row-major order; where the reference's two-step arg-max picks another cell of the
    # same value (exact ties inside one map) only the value is compared
    storage = reference_search.topk_search(pooled, maps, k)
    o_score, o_image, o_token = _lists_from_oracle(storage, k, P, W)
    top = proto_search.PrototypeTopK(P, k, cuda_device)
    for b0 in range(0, N, batch):
        top.update(pooled[b0:b0 + batch].to(cuda_device), arg[b0:b0 + batch].int().to(cuda_device))
    torch.cuda.synchronize()
    r = top.result()
    assert np.array_equal(r["score"].cpu().numpy(), o_score)
    assert np.array_equal(r["image"].cpu().numpy(), o_image)
    assert np.array_equal(r["count"].cpu().numpy(), np.array([len(storage[p]) for p in range(P)], np.int32))
    got_tok = r["token"].cpu().numpy()
    filled = o_image >= 0
    val_at = lambda tok: np.take_along_axis(flat.numpy()[np.where(filled, o_image, 0), np.arange(P)[None, :]],
                                             tok[..., None].astype(np.int64), axis=2)[..., 0]
    assert np.array_equal(val_at(got_tok)[filled], val_at(o_token)[filled])
    unique_max = (flat == pooled[:, :, None]).sum(2).numpy() == 1
    um = unique_max[np.where(filled, o_image, 0), np.arange(P)[None, :]] & filled
    assert np.array_equal(got_tok[um], o_token[um])


@pytest.mark.gpu
def test_search_through_fused_forward(cuda_device):
    """proto_search.search over three batches of the small PIP-Net case against the oracle loop run on this library's
    own materialised maps (fp32 mode: the search logic is what is compared, not the network numerics)."""
    from count_pipnet_b200.pipnet import count_pipnet as my_count
    from count_pipnet_b200.pipnet import pipnet as my_pipnet
    case = cases.CASES["pipnet_convnext13_addon32_bias"]
    model = cases.build(case, my_pipnet, my_count).to(cuda_device)
    model.cpn_precision = "fp32"
    x = cases.inputs(case).to(cuda_device)
    xs = torch.cat([x, x.flip(0) * 0.5, x * 0.25])  # three different batches
    batches = list(xs.split(x.shape[0]))
    top, grid = proto_search.search(model, batches, k=4)
    maps, pooled = [], []
    model.materialize_proto_map = True
    with torch.no_grad():
        for b in batches:
            pm, pl, _ = model(b, inference=True)
            maps.append(pm.cpu()); pooled.append(pl.cpu())
    maps, pooled = torch.cat(maps), torch.cat(pooled)
    assert grid == tuple(maps.shape[2:])
    P = pooled.shape[1]
    storage = reference_search.topk_search(pooled, maps, 4)
    o_score, o_image, o_token = _lists_from_oracle(storage, 4, P, maps.shape[3])
    r = top.result()
    assert np.array_equal(r["score"].cpu().numpy(), o_score)
    assert np.array_equal(r["image"].cpu().numpy(), o_image)
    # patches: compare where the reference's maximum is unique in its map
    flat = maps.flatten(2)
    filled = o_image >= 0
    uniq = ((flat == flat.amax(2, keepdim=True)).sum(2) == 1).numpy()
    um = uniq[np.where(filled, o_image, 0), np.arange(P)[None, :]] & filled & (o_score > 0)
    assert um.any() and np.array_equal(r["token"].cpu().numpy()[um], o_token[um])
