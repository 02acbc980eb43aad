"""The parity rules of the GPU tests, in one place (used by tests/test_parity_gpu.py and tests/parity_report.py).

North star: predicted class, arg-max prototype per image and count values EXACT; pooled scores and logits within 1e-4
relative (fp32 mode) / 2e-2 (bf16 mode), relative = max|delta| / max|reference| per tensor.

fp32 mode (the drop-in default) is held to exactly that, on every image, with no exception.

bf16 mode (opt-in) cannot be exact at the two discontinuities of the head -- `round(sum)` of the count head
(pipnet/count_pipnet.py:88-104) and `where(pooled < 0.1, 0, pooled)` of PIP-Net inference (pipnet/pipnet.py:36-37) -- so
its rules are:
  * an entry of `pooled` may differ discontinuously ("flip") ONLY where the reference's raw value sits within the
    tolerance of the boundary; flips are counted, printed and bounded; every other entry is within the tolerance;
  * logits are checked on EVERY image against the reference logits plus the exact effect of that image's permitted
    flips through the reference classifier (relu(W) @ (enc(mixed) - enc(reference))); no image is dropped;
  * the predicted class must equal the arg-max of those expected logits on every image whose top-2 margin exceeds
    twice the tolerance -- for an image without flips that IS the reference's class;
  * the arg-max prototype must equal the reference's wherever the reference's top-2 pooled margin exceeds twice the
    tolerance (integer counts: wherever the image has no flipped entry).
"""
from __future__ import annotations

import numpy as np
import torch

import oracle.reference_heads as rh

TOL = {"fp32": 1e-4, "bf16": 2e-2}
MAX_FLIP_FRACTION = 0.01


def rel(a, b) -> float:
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))


def _encode(sd, case, pooled: np.ndarray) -> np.ndarray:
    """Classifier input for a given `pooled` (second return value of forward): identity for PIP-Net, the count
    encoding for Count-PIP-Net at inference (pooled = clamped counts)."""
    if case["model"] != "count_pipnet":
        return pooled
    kind = case["args"].get("intermediate_layer", "onehot")
    enc = rh.intermediate(sd, kind, torch.from_numpy(np.ascontiguousarray(pooled, dtype=np.float32)),
                          int(case.get("max_count", 3)), bool(case.get("use_ste", True)))
    return enc.numpy()


def check_outputs(case: dict, sd: dict, precision: str, pooled: np.ndarray, out: np.ndarray, ref_pooled: np.ndarray,
                  ref_out: np.ndarray, ref_raw: np.ndarray | None = None) -> dict:
    """Asserts the rules above; returns the measured numbers (for the report / the test's printout).

    ref_raw: the reference's value BEFORE the discontinuity for inference cases (spatial sum for counts, spatial max for
    PIP-Net); needed in bf16 mode only."""
    tol = TOL[precision]
    pooled = np.asarray(pooled, dtype=np.float32)
    out = np.asarray(out, dtype=np.float32)
    inference = bool(case["inference"])
    is_count = case["model"] == "count_pipnet"
    p_scale = max(float(np.abs(ref_pooled).max()), 1e-12)
    o_scale = max(float(np.abs(ref_out).max()), 1e-12)
    ref_pred = ref_out.argmax(1)
    stats = dict(pooled=rel(pooled, ref_pooled), out=rel(out, ref_out), flips=0, flip_images=0,
                 class_equal=int((out.argmax(1) == ref_pred).sum()), images=int(out.shape[0]))

    if precision == "fp32":
        if is_count and inference:
            assert np.array_equal(pooled, ref_pooled), ("counts differ", np.argwhere(pooled != ref_pooled)[:8])
        else:
            assert stats["pooled"] <= tol, stats
        assert stats["out"] <= tol, stats
        assert np.array_equal(out.argmax(1), ref_pred), ("predicted class", out.argmax(1), ref_pred)
        assert np.array_equal(pooled.argmax(1), ref_pooled.argmax(1)), "arg-max prototype"
        stats["class_checked"] = stats["images"]
        return stats

    # ---- bf16 mode
    # c = what the classifier path consumes: clamped integer counts where the head rounds, `pooled` otherwise
    max_count = float(case.get("max_count", 3))
    rounds = is_count and (inference or bool(case.get("use_ste", True)))
    if rounds:
        raw = ref_raw if inference else ref_pooled  # the reference's spatial sums
        assert raw is not None
        c_ours = pooled if inference else np.clip(np.rint(pooled), 0.0, max_count)
        c_ref = ref_pooled if inference else np.clip(np.rint(ref_pooled), 0.0, max_count)
        rawc = np.clip(raw, 0.0, max_count + 0.49)
        near = np.abs(rawc - np.floor(rawc) - 0.5) <= tol * np.maximum(1.0, rawc)
        flipped = c_ours != c_ref
        assert not (flipped & ~near).any(), ("count differs away from a rounding boundary", np.argwhere(flipped & ~near)[:8])
        assert np.abs(c_ours - c_ref)[flipped].max(initial=0.0) <= 1.0, "a permitted flip moves a count by one"
        if not inference:  # the returned pooled values are the raw sums: continuous
            assert np.abs(pooled - ref_pooled).max() <= tol * p_scale, stats
    elif inference and not is_count:
        assert ref_raw is not None
        c_ours, c_ref = pooled, ref_pooled
        near = np.abs(ref_raw - 0.1) <= tol * p_scale
        flipped = (pooled == 0.0) != (ref_pooled == 0.0)
        assert not (flipped & ~near).any(), ("thresholded score flipped away from 0.1", np.argwhere(flipped & ~near)[:8])
        # a flipped entry still carries the right value on the side that kept it
        kept = np.where(pooled == 0.0, ref_pooled, pooled)
        assert np.abs(kept - ref_raw)[flipped].max(initial=0.0) <= tol * p_scale
        assert np.abs(pooled - ref_pooled)[~flipped].max(initial=0.0) <= tol * p_scale, stats
    else:
        c_ours, c_ref = pooled, ref_pooled
        flipped = np.zeros(pooled.shape, dtype=bool)
        assert stats["pooled"] <= tol, stats
    stats["flips"] = int(flipped.sum())
    stats["flip_images"] = int(flipped.any(axis=1).sum())
    assert stats["flips"] <= max(1, int(MAX_FLIP_FRACTION * flipped.size)), stats

    # expected logits = reference logits + the exact effect of the permitted flips through the reference classifier
    w = torch.relu(sd["_classification.weight"].detach().float().cpu()).numpy()
    expected = ref_out
    if flipped.any():
        mixed = np.where(flipped, c_ours, c_ref)
        expected = ref_out + (_encode(sd, case, mixed) - _encode(sd, case, c_ref)) @ w.T
    stats["out_vs_expected"] = float(np.abs(out - expected).max() / o_scale)
    assert stats["out_vs_expected"] <= tol, stats
    srt = np.sort(expected, axis=1)
    clear = (srt[:, -1] - srt[:, -2]) > 2 * tol * o_scale
    stats["class_checked"] = int(clear.sum())
    assert np.array_equal(out.argmax(1)[clear], expected.argmax(1)[clear]), "predicted class"
    if rounds and inference:
        same = ~flipped.any(axis=1)
        assert np.array_equal(pooled.argmax(1)[same], ref_pooled.argmax(1)[same]), "arg-max prototype"
    else:
        ps = np.sort(ref_pooled, axis=1)
        pclear = (ps[:, -1] - ps[:, -2]) > 2 * tol * p_scale
        assert np.array_equal(pooled.argmax(1)[pclear], ref_pooled.argmax(1)[pclear]), "arg-max prototype"
    return stats
