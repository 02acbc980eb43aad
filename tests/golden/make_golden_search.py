"""Generates tests/golden/search_coordinates.npz from the reference's OWN get_img_coordinates
(util/vis_pipnet.py:1162-1193). The module itself does not import here (matplotlib / kornia are missing), so the
function's source is cut out of the file with `ast` and executed as is. Run in the dev container only:
    python tests/golden/make_golden_search.py
"""
import ast
import os

import numpy as np

REF = "/root/reference/util/vis_pipnet.py"


def reference_function():
    src = open(REF).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == "get_img_coordinates":
            ns = {}
            exec(compile(ast.Module(body=[node], type_ignores=[]), REF, "exec"), ns)
            return ns["get_img_coordinates"]
    raise RuntimeError("get_img_coordinates not found")


def main():
    fn = reference_function()
    rows = []
    # (img_size, grid H, grid W): the 26x26 special case, the plain case at several sizes, non-square grids
    for img_size, H, W in [(224, 26, 26), (224, 28, 28), (224, 7, 7), (192, 24, 24), (448, 54, 54), (224, 13, 13),
                           (224, 26, 28), (256, 10, 12)]:
        for patchsize in (32, 16):
            skip = round((img_size - patchsize) / (W - 1))  # util/func.py:14
            for h in range(H):
                for w in range(W):
                    # 3-D shape (P, H, W) as at util/vis_pipnet.py:1095-1147, and the 4-D proto_features.shape
                    # (1, P, H, W) that the top-k search passes (:241) -- the function only reads [1], [2], [-1]
                    for shape in ((768, H, W), (1, 26, H, W), (1, 768, H, W)):
                        rows.append((img_size, len(shape), shape[-3], H, W, patchsize, skip, h, w)
                                    + tuple(fn(img_size, shape, patchsize, skip, h, w)))
    arr = np.asarray(rows, dtype=np.int64)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "search_coordinates.npz")
    np.savez_compressed(out, table=arr)
    print(out, arr.shape)


if __name__ == "__main__":
    main()
