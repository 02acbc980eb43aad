"""Dev tool (CPU): numerics of folding the CNBlock LayerNorm into the fc1 GEMM.

direct : a = bf16(LN(y));            h = a @ bf16(W1)^T + b1
folded : a = bf16(y); stats of a;    h = rstd * (a @ bf16(W1*gamma)^T - mean * s) + (b1 + W1 @ beta),  s = rowsum(bf16(W1*gamma))

Both are compared against the fp32 oracle on the cfg2 synthetic weights (pooled scores, logits, feature map).
"""
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, "tests")
sys.path.insert(0, ".")
import cases  # noqa: E402
from count_pipnet_b200.pipnet import count_pipnet as my_count  # noqa: E402
from count_pipnet_b200.pipnet import pipnet as my_pipnet  # noqa: E402


def bf(x):
    return x.to(torch.bfloat16).float()


def block(sd, p, x, mode):
    c = x.shape[1]
    y = F.conv2d(x, sd[p + "block.0.weight"], sd[p + "block.0.bias"], padding=3, groups=c).permute(0, 2, 3, 1)
    g, b = sd[p + "block.2.weight"], sd[p + "block.2.bias"]
    w1, b1 = sd[p + "block.3.weight"], sd[p + "block.3.bias"]
    if mode == "fp32":
        h = F.gelu(F.linear(F.layer_norm(y, (c,), g, b, 1e-6), w1, b1))
        o = F.linear(h, sd[p + "block.5.weight"], sd[p + "block.5.bias"])
    else:
        if mode == "direct":
            a = bf(F.layer_norm(y, (c,), g, b, 1e-6))
            h = F.linear(a, bf(w1), b1)
        else:
            a = bf(y)
            mean = a.mean(-1, keepdim=True)
            var = (a * a).mean(-1, keepdim=True) - mean * mean
            rstd = torch.rsqrt(var.clamp_min(0) + 1e-6)
            w1g = bf(w1 * g[None, :])
            s = w1g.sum(1)
            b1f = b1 + w1 @ b
            h = rstd * (F.linear(a, w1g) - mean * s) + b1f
        h = bf(F.gelu(h))
        o = F.linear(h, bf(sd[p + "block.5.weight"]), sd[p + "block.5.bias"])
    return x + sd[p + "layer_scale"] * o.permute(0, 3, 1, 2)


def features(sd, x, mode):
    prefix = "_net."
    idx = sorted({int(k[len(prefix):].split(".")[1]) for k in sd if k.startswith(prefix + "features.")})
    for i in idx:
        p = f"{prefix}features.{i}."
        if i == 0:
            x = F.conv2d(x, sd[p + "0.weight"], sd[p + "0.bias"], stride=4)
            x = F.layer_norm(x.permute(0, 2, 3, 1), (x.shape[1],), sd[p + "1.weight"], sd[p + "1.bias"], 1e-6).permute(0, 3, 1, 2)
        elif i % 2 == 1:
            nblk = len({k[len(p):].split(".")[0] for k in sd if k.startswith(p)})
            for j in range(nblk):
                x = block(sd, f"{p}{j}.", x, mode)
        else:
            c = x.shape[1]
            xn = F.layer_norm(x.permute(0, 2, 3, 1), (c,), sd[p + "0.weight"], sd[p + "0.bias"], 1e-6).permute(0, 3, 1, 2)
            stride = 1 if c > 100 else 2
            w = sd[p + "1.weight"]
            if mode != "fp32":
                xn, w = bf(xn), bf(w)
            x = F.conv2d(xn, w, sd[p + "1.bias"], stride=stride)
    return x


def main():
    torch.set_num_threads(8)
    case = cases.CASES["cfg2_pipnet_convnext26_224"]
    model = cases.build(case, my_pipnet, my_count)
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    x = cases.inputs(case)
    outs = {}
    with torch.no_grad():
        for mode in ("fp32", "direct", "folded"):
            f = features(sd, x, mode)
            pm = torch.softmax(f, 1)
            pooled = pm.flatten(2).amax(2)
            logits = F.linear(pooled, torch.relu(sd["_classification.weight"]), sd.get("_classification.bias"))
            outs[mode] = (f, pooled, logits)
    ref = outs["fp32"]
    for mode in ("direct", "folded"):
        o = outs[mode]
        print(mode, *[f"{n}={float((a - b).abs().max() / b.abs().max()):.3e}" for n, a, b in
                      zip(("feat", "pooled", "logits"), o, ref)],
              "argmax_equal", bool((o[2].argmax(1) == ref[2].argmax(1)).all()),
              "protoargmax_equal", bool((o[1].argmax(1) == ref[1].argmax(1)).all()))


if __name__ == "__main__":
    main()
