#!/usr/bin/env python
"""Runs one C-ABI kernel alone at a benchmark shape (for ncu captures and CUDA-event timing).
    python tools/run_op.py dwconv --B 64 --H 27 --W 27 --C 384 [--iters 5]
    python tools/run_op.py gemm --M 46656 --N 1536 --K 384 --act 1 [--res 1]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from count_pipnet_b200 import _lib, ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("op")
    for k, d in dict(B=64, H=27, W=27, C=384, M=46656, N=1536, K=384, act=0, res=0, iters=5).items():
        ap.add_argument("--" + k, type=int, default=d)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    if a.op == "dwconv":
        x = torch.randn(a.B, a.H, a.W, a.C, generator=g).to(dev)
        w49 = (torch.randn(49, a.C, generator=g) * 0.2).to(dev)
        b, lw, lb = (torch.randn(a.C, generator=g).to(dev) for _ in range(3))
        out = torch.empty(a.B * a.H * a.W, a.C, device=dev, dtype=torch.bfloat16)
        fn = lambda: ops.dwconv7_ln(x, w49, b, lw, lb, out)
        nbytes = x.numel() * 4 + out.numel() * 2
        flops = 2.0 * 49 * x.numel()
    elif a.op == "dwstats":
        x = torch.randn(a.B, a.H, a.W, a.C, generator=g).to(dev)
        w49 = (torch.randn(49, a.C, generator=g) * 0.2).to(dev)
        b = torch.randn(a.C, generator=g).to(dev)
        out = torch.empty(a.B * a.H * a.W, a.C, device=dev, dtype=torch.bfloat16)
        stats = torch.empty(a.B * a.H * a.W, a.C // 32, 2, device=dev)
        fn = lambda: ops.dwconv7_stats(x, w49, b, out, stats)
        nbytes = x.numel() * 4 + out.numel() * 2 + stats.numel() * 4
        flops = 2.0 * 49 * x.numel()
    elif a.op == "dwsplit":
        x = torch.randn(a.B, a.H, a.W, a.C, generator=g).to(dev)
        w49 = (torch.randn(49, a.C, generator=g) * 0.2).to(dev)
        b = torch.randn(a.C, generator=g).to(dev)
        hi = torch.empty(a.B * a.H * a.W, a.C, device=dev, dtype=torch.bfloat16)
        lo = torch.empty_like(hi)
        stats = torch.empty(a.B * a.H * a.W, a.C // 32, 2, device=dev)
        fn = lambda: ops.dwconv7_stats_split(x, w49, b, hi, lo, stats)
        nbytes = x.numel() * 4 + hi.numel() * 4 + stats.numel() * 4
        flops = 2.0 * 49 * x.numel()
    elif a.op == "preprocess":
        # B decoded images of H x W (uint8 RGB) -> [B, 3, C, C] fp32 (C = target size): the reference's eval transform
        from count_pipnet_b200.input_pipeline import EvalTransform
        u8 = torch.randint(0, 256, (a.B, a.H, a.W, 3), generator=g, dtype=torch.uint8).to(dev)
        tf = EvalTransform(a.C)
        outp = torch.empty(a.B, 3, a.C, a.C, device=dev)
        fn = lambda: tf(u8, outp)
        nbytes = u8.numel() + outp.numel() * 4
        flops = 0.0
    elif a.op == "gemm_ln":
        A = torch.randn(a.M, a.K, generator=g).to(dev).to(torch.bfloat16)
        Wt = (torch.randn(a.N, a.K, generator=g) / a.K ** 0.5).to(dev).to(torch.bfloat16)
        bias = torch.randn(a.N, generator=g).to(dev)
        colsum = Wt.float().sum(1).contiguous()
        af = A.float().reshape(a.M, a.K // 32, 32)
        stats = torch.stack([af.sum(-1), (af * af).sum(-1)], dim=-1).contiguous()
        out = torch.empty(a.M, a.N, device=dev, dtype=torch.bfloat16)
        fn = lambda: ops.gemm_ln_gelu(A, Wt, a.M, a.N, a.K, stats, a.K // 32, colsum, bias, out)
        nbytes = A.numel() * 2 + Wt.numel() * 2 + out.numel() * 2
        flops = 2.0 * a.M * a.N * a.K
    elif a.op == "gemm":
        A = torch.randn(a.M, a.K, generator=g).to(dev).to(torch.bfloat16)
        Wt = (torch.randn(a.N, a.K, generator=g) / a.K ** 0.5).to(dev).to(torch.bfloat16)
        bias = torch.randn(a.N, generator=g).to(dev)
        if a.res:
            res = torch.randn(a.M, a.N, generator=g).to(dev)
            scale = torch.rand(a.N, generator=g).to(dev)
            fn = lambda: ops.gemm_bf16(A, Wt, a.M, a.N, a.K, bias=bias, scale=scale, res_f32=res, out_f32=res)
            nbytes = A.numel() * 2 + Wt.numel() * 2 + res.numel() * 8
        else:
            out = torch.empty(a.M, a.N, device=dev, dtype=torch.bfloat16)
            fn = lambda: ops.gemm_bf16(A, Wt, a.M, a.N, a.K, bias=bias, out_bf16=out, act=a.act)
            nbytes = A.numel() * 2 + Wt.numel() * 2 + out.numel() * 2
        flops = 2.0 * a.M * a.N * a.K
    elif a.op == "mlp":
        C = a.C
        y = (torch.randn(a.M, C, generator=g)).to(dev).to(torch.bfloat16)
        w1g = (torch.randn(4 * C, C, generator=g) / C ** 0.5).to(dev).to(torch.bfloat16)
        w2 = (torch.randn(C, 4 * C, generator=g) / (4 * C) ** 0.5).to(dev).to(torch.bfloat16)
        b1 = torch.randn(4 * C, generator=g).to(dev)
        b2 = torch.randn(C, generator=g).to(dev)
        scale = (torch.rand(C, generator=g) * 0.01).to(dev)
        colsum = w1g.float().sum(1).contiguous()
        yf = y.float().reshape(a.M, C // 32, 32)
        stats = torch.stack([yf.sum(-1), (yf * yf).sum(-1)], dim=-1).contiguous()
        x = torch.randn(a.M, C, generator=g).to(dev)
        fn = lambda: ops.mlp_fused(y, w1g, w2, a.M, C, stats, colsum, b1, b2, scale, x)
        nbytes = y.numel() * 2 + x.numel() * 8 + w1g.numel() * 4
        flops = 16.0 * a.M * C * C
    elif a.op in ("pipnet_loss", "head_fwd", "head_bwd"):
        # training side of the head: --B image PAIRS, --C prototypes, --H x --W map, --N classes
        z = (torch.randn(2 * a.B, a.C, a.H, a.W, generator=g) * 3).to(dev)
        pf = torch.softmax(z, dim=1)
        pooled = pf.amax(dim=(2, 3))
        K = min(a.N, 1000)
        out = pooled @ torch.relu(torch.randn(K, a.C, generator=g)).to(dev).t()
        ys = torch.randint(0, K, (a.B,), generator=g).to(dev)
        mult = torch.tensor([1.3], device=dev)
        gpf, gpl, go = torch.empty_like(pf), torch.empty_like(pooled), torch.empty_like(out)
        flops = 0.0
        if a.op == "pipnet_loss":
            res = torch.empty(8, device=dev)
            ws = torch.empty(ops.pipnet_loss_ws_floats(a.B, a.C, a.H * a.W), device=dev)
            fn = lambda: ops.pipnet_loss(pf, pooled, out, ys, mult, a.B, a.C, a.H * a.W, K, 5., 2., 2., 1., 1e-8, False,
                                         False, True, gpf, gpl, go, res, ws)
            nbytes = 8 * pf.numel()
        else:
            arg = torch.empty(2 * a.B, a.C, device=dev, dtype=torch.int32)
            ws = torch.empty(ops.softmax_pool_nchw_ws_floats(2 * a.B, a.C, a.H * a.W), device=dev)
            if a.op == "head_fwd":
                fn = lambda: ops.softmax_pool_nchw(z, _lib.POOL_MAX, pf, pooled, arg, ws)
                nbytes = 8 * pf.numel()
            else:
                ops.softmax_pool_nchw(z, _lib.POOL_MAX, pf, pooled, arg, ws)
                gpf.normal_()
                gpl.normal_()
                fn = lambda: ops.softmax_pool_nchw_bwd(pf, gpf, gpl, arg, _lib.POOL_MAX, z)
                nbytes = 12 * pf.numel()
    else:
        raise SystemExit("unknown op")
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    print(f"{a.op} {vars(a)}: {ms:.4f} ms  {flops / ms / 1e9:.1f} TFLOP/s  {nbytes / ms / 1e6:.1f} GB/s")


if __name__ == "__main__":
    main()
