"""CPU restatement of the reference's evaluation-time input pipeline -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py):
only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it; the product never does.

The reference builds, per dataset, `transform_no_augment = Compose([Resize(size=(S, S)), ToTensor(), Normalize(mean, std)])`
(util/data.py:381-385 shapes, :265-275 pets / :296-307 partimagenet / :465-480 CUB / :522-531 cars: the same three steps)
and applies it to PIL RGB images in the DataLoader workers (util/data.py:596-617).

The arithmetic lives in two pinned third-party dependencies that are not part of /root/reference:
  * Pillow (12.2 in this image), `Image.resize(..., BILINEAR)` = src/libImaging/Resample.c: a separable triangle filter
    whose support widens with the down-scaling factor, coefficients computed in double precision and rounded to 22-bit
    fixed point, a horizontal pass to uint8 and then a vertical pass to uint8 (`precompute_coeffs`, `normalize_coeffs_8bpc`,
    `ImagingResampleHorizontal_8bpc`, `ImagingResampleVertical_8bpc`);
  * torchvision 0.26 `ToTensor` (uint8 HWC -> float32 CHW, true division by 255) and `Normalize` ((x - mean) / std, fp32).
Parity is pinned by tests/test_preprocess_oracle.py against Pillow + torchvision themselves (installed here and on the GPU
box) and by the committed fixture tests/golden/preprocess_golden.npz written by tests/make_golden_preprocess.py.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2  # Resample.c


def bilinear_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the triangle filter (support 1.0), box = whole image.
    Returns (bounds [out_size, 2] int32 = (first input index, count), coeffs [out_size, ksize] int32)."""
    scale = filterscale = float(in_size) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    coeffs = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)  # C cast: truncation (the argument is never below -0.5 + eps here)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = np.zeros(ksize, dtype=np.float64)
        ww = 0.0
        for x in range(xmax):
            a = (x + xmin - center + 0.5) * ss
            if a < 0.0:
                a = -a
            w = 1.0 - a if a < 1.0 else 0.0
            k[x] = w
            ww += w
        for x in range(xmax):
            if ww != 0.0:
                k[x] /= ww
        for x in range(ksize):
            v = k[x] * (1 << PRECISION_BITS)
            coeffs[xx, x] = int(-0.5 + v) if k[x] < 0 else int(0.5 + v)
        bounds[xx] = (xmin, xmax)
    return bounds, coeffs


def _clip8(acc: np.ndarray) -> np.ndarray:
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def resize_bilinear_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """Image.resize((out_w, out_h), BILINEAR) of a uint8 [H, W, C] image: horizontal pass, then vertical pass."""
    h, w, _ = img.shape
    cur = img
    if out_w != w:
        b, k = bilinear_coeffs(w, out_w)
        out = np.empty((h, out_w, img.shape[2]), dtype=np.uint8)
        src = cur.astype(np.int64)
        for xx in range(out_w):
            x0, n = int(b[xx, 0]), int(b[xx, 1])
            acc = (src[:, x0:x0 + n, :] * k[xx, :n].astype(np.int64)[None, :, None]).sum(axis=1) + (1 << (PRECISION_BITS - 1))
            out[:, xx, :] = _clip8(acc)
        cur = out
    if out_h != h:
        b, k = bilinear_coeffs(h, out_h)
        out = np.empty((out_h, cur.shape[1], img.shape[2]), dtype=np.uint8)
        src = cur.astype(np.int64)
        for yy in range(out_h):
            y0, n = int(b[yy, 0]), int(b[yy, 1])
            acc = (src[y0:y0 + n, :, :] * k[yy, :n].astype(np.int64)[:, None, None]).sum(axis=0) + (1 << (PRECISION_BITS - 1))
            out[yy] = _clip8(acc)
        cur = out
    return cur


def to_tensor_normalize(img_u8: np.ndarray, mean, std) -> np.ndarray:
    """ToTensor + Normalize: uint8 [H, W, 3] -> float32 [3, H, W] = ((x / 255) - mean) / std, every step in fp32."""
    x = img_u8.transpose(2, 0, 1).astype(np.float32) / np.float32(255.0)
    m = np.asarray(mean, dtype=np.float32)[:, None, None]
    s = np.asarray(std, dtype=np.float32)[:, None, None]
    return ((x - m) / s).astype(np.float32)


def preprocess(img_u8: np.ndarray, size: int, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)) -> np.ndarray:
    """transform_no_augment of util/data.py:381-385 on one decoded RGB image."""
    return to_tensor_normalize(resize_bilinear_u8(img_u8, size, size), mean, std)
