"""cpn_preprocess_u8 (GPU input pipeline) through the C ABI against oracle/reference_preprocess.py and
the committed Pillow + torchvision outputs: bit-exact (integer resize, fp32 table look-up)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from count_pipnet_b200 import _lib
from count_pipnet_b200.input_pipeline import IMAGENET_MEAN, IMAGENET_STD, EvalTransform
from oracle import reference_preprocess as rp
from test_preprocess_oracle import golden_cases

pytestmark = pytest.mark.gpu


def test_preprocess_golden(cuda_device):
    for src, size, _resized, out in golden_cases():
        y = EvalTransform(size)(torch.from_numpy(src)[None].to(cuda_device))
        torch.cuda.synchronize()
        assert np.array_equal(y[0].cpu().numpy(), out)


@pytest.mark.parametrize("B,h,w,oh,ow", [(3, 37, 53, 24, 24), (2, 100, 80, 224, 224), (2, 224, 224, 224, 224), (1, 500, 333, 448, 448),
                                          (4, 7, 9, 24, 24), (1, 1024, 768, 192, 192), (2, 31, 224, 224, 31), (1, 300, 224, 224, 224),
                                          (2, 4000, 30, 64, 64)])
def test_preprocess_matches_oracle(cuda_device, B, h, w, oh, ow):
    rs = np.random.RandomState(B * 1000 + h + w)
    imgs = (rs.rand(B, h, w, 3) * 255).astype(np.uint8)
    imgs[0, : max(1, h // 3)] = 255  # saturated rows: clipping after each pass
    mean, std = (0.5, 0.4, 0.3), (0.25, 0.2, 0.3)
    y = EvalTransform((oh, ow), mean, std)(torch.from_numpy(imgs).to(cuda_device))
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    for b in range(B):
        ref = rp.to_tensor_normalize(rp.resize_bilinear_u8(imgs[b], oh, ow), mean, std)
        assert np.array_equal(y[b], ref), (b, np.abs(y[b] - ref).max())


def test_preprocess_bench_shape_properties(cuda_device):
    """The benchmark's shape (64 images of 375 x 500 -> 224 x 224): three images against the oracle, and size-independent
    properties on all of them: a constant image stays constant, same-size resize is the identity, and every output value is
    one of the 3 x 256 table entries."""
    rs = np.random.RandomState(11)
    B, h, w, s = 64, 375, 500, 224
    imgs = (rs.rand(B, h, w, 3) * 255).astype(np.uint8)
    imgs[5] = np.array([17, 130, 251], dtype=np.uint8)  # constant image
    tf = EvalTransform(s)
    y = tf(torch.from_numpy(imgs).to(cuda_device))
    torch.cuda.synchronize()
    yc = y.cpu().numpy()
    for b in (0, 31, 63):
        assert np.array_equal(yc[b], rp.preprocess(imgs[b], s))
    table = rp.to_tensor_normalize(np.arange(256, dtype=np.uint8).reshape(256, 1, 1).repeat(3, axis=2), IMAGENET_MEAN, IMAGENET_STD)
    for c, v in enumerate((17, 130, 251)):
        assert (yc[5, c] == table[c, v, 0]).all()
    for c in range(3):
        assert np.isin(yc[:, c], table[c, :, 0]).all()
    same = EvalTransform((h, w))(torch.from_numpy(imgs[:4]).to(cuda_device)).cpu().numpy()
    for b in range(4):
        assert np.array_equal(same[b], rp.to_tensor_normalize(imgs[b], IMAGENET_MEAN, IMAGENET_STD))


def test_preprocess_rejects(cuda_device):
    tf = EvalTransform(8)
    with pytest.raises(_lib.CpnError):
        tf(torch.zeros(1, 16, 16, 3, dtype=torch.uint8))  # CPU tensor: no fallback
    with pytest.raises(ValueError):
        tf(torch.zeros(1, 16, 16, 3, device=cuda_device))  # not uint8
    with pytest.raises(_lib.CpnError, match="down-scaling"):
        EvalTransform(8)(torch.zeros(1, 1024, 16, 3, dtype=torch.uint8, device=cuda_device))
