"""Writes tests/golden/preprocess_golden.npz: outputs of the reference's own evaluation transform
(util/data.py:381-385: torchvision Compose([Resize((S, S)), ToTensor(), Normalize(mean, std)]) on PIL images) for a few
seeded images, produced by Pillow + torchvision themselves in this container. Run from the repo root:

    python tests/golden/make_golden_preprocess.py
"""
import os

import numpy as np
import PIL
import torchvision
import torchvision.transforms as transforms
from PIL import Image

MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)   # util/data.py:377-378
CASES = [  # (source height, source width, target size)
    (37, 53, 24), (20, 16, 24), (64, 48, 32), (32, 32, 32), (75, 100, 48), (9, 120, 16),
]


def main():
    out = {"pillow": np.array(PIL.__version__), "torchvision": np.array(torchvision.__version__)}
    rs = np.random.RandomState(20260)
    for i, (h, w, s) in enumerate(CASES):
        # smooth structure + noise + saturated patches: exercises rounding, clipping and both filter regimes
        yy, xx = np.mgrid[0:h, 0:w]
        base = 127 + 100 * np.sin(yy[..., None] / 5.0 + np.arange(3)) * np.cos(xx[..., None] / 7.0)
        img = np.clip(base + rs.randn(h, w, 3) * 40, 0, 255).astype(np.uint8)
        img[: h // 4, : w // 4] = 255
        img[-(h // 5) - 1:, -(w // 5) - 1:] = 0
        tf = transforms.Compose([transforms.Resize(size=(s, s)), transforms.ToTensor(), transforms.Normalize(mean=MEAN, std=STD)])
        y = tf(Image.fromarray(img)).numpy()
        r = np.asarray(Image.fromarray(img).resize((s, s), Image.BILINEAR))
        out[f"src{i}"] = img
        out[f"resized{i}"] = r
        out[f"out{i}"] = y
        out[f"size{i}"] = np.array(s)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "preprocess_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
