"""oracle/reference_preprocess.py (numpy restatement of Pillow's BILINEAR resize + torchvision ToTensor / Normalize) against
the committed outputs of Pillow + torchvision (tests/golden/preprocess_golden.npz, written by make_golden_preprocess.py) and,
where those libraries are importable, against the libraries themselves on seeded images."""
from __future__ import annotations

import os

import numpy as np
import pytest

from oracle import reference_preprocess as rp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "preprocess_golden.npz")


def golden_cases():
    g = np.load(GOLDEN)
    n = len([k for k in g.files if k.startswith("src")])
    return [(g[f"src{i}"], int(g[f"size{i}"]), g[f"resized{i}"], g[f"out{i}"]) for i in range(n)]


def test_oracle_matches_golden():
    cases = golden_cases()
    assert len(cases) >= 6
    for src, size, resized, out in cases:
        assert np.array_equal(rp.resize_bilinear_u8(src, size, size), resized)
        assert np.array_equal(rp.preprocess(src, size), out)  # bit-exact fp32


def test_coefficients_known_answers():
    """Resample.c by hand: same size -> identity taps; 2x down-scaling -> triangle (1, 3, 3, 1) / 8 in the interior."""
    b, k = rp.bilinear_coeffs(8, 8)
    for i in range(8):
        taps = {int(b[i, 0]) + j: int(k[i, j]) for j in range(int(b[i, 1])) if k[i, j] != 0}
        assert taps == {i: 1 << 22}
    b, k = rp.bilinear_coeffs(16, 8)
    assert tuple(b[3]) == (5, 4)
    assert [int(v) for v in k[3, :4]] == [1 << 19, 3 << 19, 3 << 19, 1 << 19]


@pytest.mark.parametrize("h,w,oh,ow", [(375, 500, 224, 224), (100, 80, 224, 224), (224, 224, 224, 224), (500, 333, 448, 448),
                                        (7, 9, 24, 24), (1024, 768, 192, 192), (31, 224, 224, 31), (300, 224, 224, 224)])
def test_oracle_matches_pillow_live(h, w, oh, ow):
    Image = pytest.importorskip("PIL.Image")
    rs = np.random.RandomState(h * 7 + w)
    img = (rs.rand(h, w, 3) * 255).astype(np.uint8)
    ref = np.asarray(Image.fromarray(img).resize((ow, oh), Image.BILINEAR))
    assert np.array_equal(rp.resize_bilinear_u8(img, oh, ow), ref)


def test_oracle_matches_torchvision_live():
    Image = pytest.importorskip("PIL.Image")
    transforms = pytest.importorskip("torchvision.transforms")
    rs = np.random.RandomState(5)
    img = (rs.rand(120, 90, 3) * 255).astype(np.uint8)
    mean, std = (0.5, 0.4, 0.3), (0.25, 0.2, 0.3)
    tf = transforms.Compose([transforms.Resize(size=(64, 64)), transforms.ToTensor(), transforms.Normalize(mean=mean, std=std)])
    assert np.array_equal(rp.preprocess(img, 64, mean, std), tf(Image.fromarray(img)).numpy())
