"""Per-kernel numerics on the GPU: every C-ABI kernel against a plain PyTorch fp32 computation of the same op."""
import pytest
import torch
import torch.nn.functional as F

from count_pipnet_b200 import _lib, ops

pytestmark = pytest.mark.gpu


def _bf16_split(t):
    hi = t.to(torch.bfloat16)
    lo = (t - hi.float()).to(torch.bfloat16)
    return hi, lo


@pytest.mark.parametrize("M,N,K", [(128, 96, 64), (1000, 384, 96), (4096, 768, 192), (5000, 1536, 384),
                                   (777, 192, 768), (3000, 3072, 768), (2048, 64, 152), (300, 32, 64)])
def test_gemm_plain(cuda_device, M, N, K):
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(N, generator=g).to(cuda_device)
    out = torch.empty(M, N, device=cuda_device)
    ops.gemm_bf16(a, w, M, N, K, bias=bias, out_f32=out)
    ref = a.float() @ w.float().t() + bias
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("M,N,K", [(3000, 768, 192), (777, 384, 96), (2500, 1536, 384), (1300, 3072, 768), (200, 64, 64)])
def test_gemm_gelu_bf16_out(cuda_device, M, N, K):
    g = torch.Generator().manual_seed(1)
    a = torch.randn(M, K, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(N, generator=g).to(cuda_device)
    out = torch.empty(M, N, device=cuda_device, dtype=torch.bfloat16)
    ops.gemm_bf16(a, w, M, N, K, bias=bias, out_bf16=out, act=_lib.ACT_GELU)
    ref = F.gelu(a.float() @ w.float().t() + bias)
    torch.cuda.synchronize()
    assert (out.float() - ref).abs().max().item() < 0.03


@pytest.mark.parametrize("M,N,K", [(2500, 192, 768), (5000, 96, 384), (1111, 384, 1536), (900, 768, 3072), (130, 32, 64)])
def test_gemm_scale_residual_inplace(cuda_device, M, N, K):
    g = torch.Generator().manual_seed(2)
    a = torch.randn(M, K, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(N, generator=g).to(cuda_device)
    ls = torch.rand(N, generator=g).to(cuda_device)
    res = torch.randn(M, N, generator=g).to(cuda_device)
    ref = res + ls * (a.float() @ w.float().t() + bias)
    ops.gemm_bf16(a, w, M, N, K, bias=bias, scale=ls, res_f32=res, out_f32=res)
    torch.cuda.synchronize()
    assert (res - ref).abs().max().item() < 5e-3


def test_gemm_relu_bf16_residual(cuda_device):
    M, N, K = 1568, 256, 64
    g = torch.Generator().manual_seed(3)
    a = torch.randn(M, K, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(N, generator=g).to(cuda_device)
    res = torch.randn(M, N, generator=g).to(cuda_device).to(torch.bfloat16)
    out = torch.empty(M, N, device=cuda_device, dtype=torch.bfloat16)
    out32 = torch.empty(M, N, device=cuda_device)
    ops.gemm_bf16(a, w, M, N, K, bias=bias, res_bf16=res, out_bf16=out, out_f32=out32, act=_lib.ACT_RELU,
                  relu_after_res=True)
    ref = torch.relu(a.float() @ w.float().t() + bias + res.float())
    torch.cuda.synchronize()
    assert (out32 - ref).abs().max().item() < 5e-3
    assert (out.float() - ref).abs().max().item() < 0.05


def test_gemm_split_fp32_mode(cuda_device):
    """bf16x3 (hi/lo) GEMM must be ~fp32 accurate."""
    M, N, K = 2000, 384, 1536
    g = torch.Generator().manual_seed(4)
    a = torch.randn(M, K, generator=g).to(cuda_device)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device)
    a_hi, a_lo = _bf16_split(a)
    w_hi, w_lo = _bf16_split(w)
    out = torch.empty(M, N, device=cuda_device)
    ops.gemm_bf16(a_hi, w_hi, M, N, K, a_lo=a_lo, bw_lo=w_lo, out_f32=out)
    ref = (a.double() @ w.double().t()).float()
    torch.cuda.synchronize()
    rel = (out - ref).abs().max().item() / ref.abs().max().item()
    assert rel < 5e-5, rel


@pytest.mark.parametrize("B,H,W", [(3, 64, 96), (2, 224, 224), (1, 448, 448), (2, 36, 52), (5, 8, 8), (2, 30, 41)])
def test_stem_ln(cuda_device, B, H, W):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, 3, H, W, generator=g).to(cuda_device)
    w = (torch.randn(96, 3, 4, 4, generator=g) * 0.2).to(cuda_device)
    b = torch.randn(96, generator=g).to(cuda_device)
    lw = torch.rand(96, generator=g).to(cuda_device) + 0.5
    lb = torch.randn(96, generator=g).to(cuda_device)
    out = torch.full((B, H // 4, W // 4, 96), float("nan"), device=cuda_device)
    ops.stem_ln(x, w.reshape(96, 48).contiguous(), b, lw, lb, out)
    torch.backends.cudnn.allow_tf32 = False
    y = F.conv2d(x, w, b, stride=4).permute(0, 2, 3, 1)
    ref = F.layer_norm(y, (96,), lw, lb, 1e-6)
    torch.cuda.synchronize()
    assert (out - ref).abs().max().item() < 2e-4


@pytest.mark.parametrize("B,H,W,C", [(2, 12, 12, 96), (3, 28, 28, 192), (2, 27, 27, 384), (2, 26, 26, 768), (1, 5, 9, 192),
                                     (2, 56, 56, 96), (1, 54, 55, 384), (70, 7, 6, 96), (1, 3, 2, 64), (2, 13, 13, 40)])
def test_dwconv7_ln(cuda_device, B, H, W, C):
    g = torch.Generator().manual_seed(C + H)
    x = torch.randn(B, H, W, C, generator=g).to(cuda_device)
    w = (torch.randn(C, 1, 7, 7, generator=g) * 0.2).to(cuda_device)
    b = torch.randn(C, generator=g).to(cuda_device)
    lw = torch.rand(C, generator=g).to(cuda_device) + 0.5
    lb = torch.randn(C, generator=g).to(cuda_device)
    out = torch.empty(B * H * W, C, device=cuda_device, dtype=torch.bfloat16)
    lo = torch.empty_like(out)
    w49 = w.reshape(C, 49).t().contiguous()
    ops.dwconv7_ln(x, w49, b, lw, lb, out, lo)
    torch.backends.cudnn.allow_tf32 = False
    y = F.conv2d(x.permute(0, 3, 1, 2), w, b, padding=3, groups=C).permute(0, 2, 3, 1)
    ref = F.layer_norm(y, (C,), lw, lb, 1e-6).reshape(B * H * W, C)
    torch.cuda.synchronize()
    assert (out.float() - ref).abs().max().item() < 0.03
    assert ((out.float() + lo.float()) - ref).abs().max().item() < 5e-4


DW_SHAPES = [(2, 12, 12, 96), (3, 28, 28, 192), (2, 27, 27, 384), (2, 26, 26, 768), (1, 5, 9, 192), (1, 56, 56, 96),
             (2, 3, 2, 96), (1, 1, 1, 384), (1, 54, 54, 768),
             # second-generation kernel: odd image count with two images per job, column blocks (W > 56), several
             # images per job on narrow maps, the benchmark's own shapes
             (3, 27, 27, 384), (1, 112, 112, 96), (2, 55, 55, 384), (5, 6, 6, 192), (9, 13, 13, 96), (7, 24, 24, 192),
             (64, 27, 27, 384), (64, 56, 56, 96), (16, 26, 26, 768),
             # widths that are not ConvNeXt-tiny's: the generic kernel (dwconv_v2.cuh)
             (3, 27, 27, 128), (2, 56, 56, 64), (1, 112, 112, 32), (5, 6, 6, 160)]


@pytest.mark.parametrize("B,H,W,C", DW_SHAPES)
def test_dwconv7_stats(cuda_device, B, H, W, C):
    """Un-normalised depthwise conv (bf16) + per-32-channel partial LayerNorm sums, vs F.conv2d in fp32."""
    g = torch.Generator().manual_seed(C + H)
    x = torch.randn(B, H, W, C, generator=g).to(cuda_device)
    w = (torch.randn(C, 1, 7, 7, generator=g) * 0.2).to(cuda_device)
    b = torch.randn(C, generator=g).to(cuda_device)
    T = B * H * W
    out = torch.full((T, C), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    stats = torch.full((T, C // 32, 2), float("nan"), device=cuda_device)
    w49 = w.reshape(C, 49).t().contiguous()
    ops.dwconv7_stats(x, w49, b, out, stats)
    out2, stats2 = torch.empty_like(out), torch.empty_like(stats)
    ops.dwconv7_stats(x, w49, b, out2, stats2)
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(x.permute(0, 3, 1, 2), w, b, padding=3, groups=C).permute(0, 2, 3, 1).reshape(T, C)
    torch.cuda.synchronize()
    assert torch.equal(out, out2) and torch.equal(stats, stats2)  # fixed summation order
    assert (out.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-5
    # the sums are those of the ROUNDED values, per 32-channel group
    o = out.float().reshape(T, C // 32, 32)
    assert torch.allclose(stats[..., 0], o.sum(-1), rtol=1e-5, atol=1e-4)
    assert torch.allclose(stats[..., 1], (o * o).sum(-1), rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("B,H,W,C", DW_SHAPES[:14] + DW_SHAPES[-4:])
def test_dwconv7_stats_split(cuda_device, B, H, W, C):
    """fp32 mode: hi + lo planes carry conv + bias to ~2^-17; the partial sums are those of hi + lo."""
    g = torch.Generator().manual_seed(C + H)
    x = torch.randn(B, H, W, C, generator=g).to(cuda_device)
    w = (torch.randn(C, 1, 7, 7, generator=g) * 0.2).to(cuda_device)
    b = torch.randn(C, generator=g).to(cuda_device)
    T = B * H * W
    hi = torch.full((T, C), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    lo = torch.full((T, C), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    stats = torch.full((T, C // 32, 2), float("nan"), device=cuda_device)
    w49 = w.reshape(C, 49).t().contiguous()
    ops.dwconv7_stats_split(x, w49, b, hi, lo, stats)
    hi2, lo2, stats2 = torch.empty_like(hi), torch.empty_like(lo), torch.empty_like(stats)
    ops.dwconv7_stats_split(x, w49, b, hi2, lo2, stats2)
    ref = F.conv2d(x.double().permute(0, 3, 1, 2), w.double(), b.double(), padding=3, groups=C).permute(0, 2, 3, 1)
    ref = ref.reshape(T, C)
    torch.cuda.synchronize()
    assert torch.equal(hi, hi2) and torch.equal(lo, lo2) and torch.equal(stats, stats2)  # fixed summation order
    v = hi.float() + lo.float()
    assert (v.double() - ref).abs().max().item() <= 3e-5 * max(1.0, ref.abs().max().item())
    o = v.reshape(T, C // 32, 32)
    assert torch.allclose(stats[..., 0], o.sum(-1), rtol=1e-5, atol=1e-4)
    assert torch.allclose(stats[..., 1], (o * o).sum(-1), rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("M,C", [(3000, 192), (777, 96), (2500, 384), (1300, 768), (128, 96), (46656, 384), (20000, 96)])
def test_gemm_ln_gelu_split(cuda_device, M, C):
    """fp32 mode: LayerNorm folded into the three-pass fc1 with the exact GELU, hi/lo output planes, vs fp64."""
    g = torch.Generator().manual_seed(C + 1)
    N = 4 * C
    a = (torch.randn(M, C, generator=g) * 1.5 + 0.7).to(cuda_device)
    w1 = (torch.randn(N, C, generator=g) / C ** 0.5).to(cuda_device)
    b1 = torch.randn(N, generator=g).to(cuda_device)
    gamma = (torch.rand(C, generator=g) + 0.5).to(cuda_device)
    beta = torch.randn(C, generator=g).to(cuda_device)
    a_hi, a_lo = _bf16_split(a)
    av = (a_hi.float() + a_lo.float()).reshape(M, C // 32, 32)
    stats = torch.stack([av.sum(-1), (av * av).sum(-1)], dim=-1).contiguous()
    w_hi, w_lo = _bf16_split(w1 * gamma[None, :])
    colsum = (w_hi.double() + w_lo.double()).sum(1).float().contiguous()
    b1f = (b1.double() + w1.double() @ beta.double()).float().contiguous()
    o_hi = torch.full((M, N), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    o_lo = torch.full((M, N), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    ops.gemm_ln_gelu_split(a_hi, a_lo, w_hi, w_lo, M, N, C, stats, C // 32, colsum, b1f, o_hi, o_lo)
    ref = F.gelu(F.linear(F.layer_norm(a.double(), (C,), gamma.double(), beta.double(), 1e-6), w1.double(), b1.double()))
    torch.cuda.synchronize()
    out = o_hi.double() + o_lo.double()
    rel = (out - ref).abs().max().item() / ref.abs().max().item()
    assert rel < 5e-5, rel
    # the planes are a proper split: lo is the rounding remainder of hi
    assert (o_lo.float().abs() <= o_hi.float().abs() * 2.0 ** -8 + 1e-30).all()


@pytest.mark.parametrize("M,N,K,res", [(2500, 192, 768, True), (5000, 96, 384, True), (1111, 384, 1536, True),
                                       (900, 768, 3072, True), (20000, 768, 1536, False), (20000, 384, 768, False),
                                       (30000, 192, 384, False), (46656, 384, 1536, True)])
def test_gemm_split_tma_epilogues(cuda_device, M, N, K, res):
    """fp32 mode, fp32 outputs through the TMA epilogue: x += scale * (h @ W2^T + b) in place, and plain bias (downsample)."""
    g = torch.Generator().manual_seed(N + K)
    a = torch.randn(M, K, generator=g).to(cuda_device)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device)
    bias = torch.randn(N, generator=g).to(cuda_device)
    a_hi, a_lo = _bf16_split(a)
    w_hi, w_lo = _bf16_split(w)
    ref = a.double() @ w.double().t() + bias.double()
    if res:
        x = torch.randn(M, N, generator=g).to(cuda_device)
        scale = (torch.rand(N, generator=g) + 0.25).to(cuda_device)
        ref = x.double() + scale.double() * ref
        ops.gemm_bf16(a_hi, w_hi, M, N, K, a_lo=a_lo, bw_lo=w_lo, bias=bias, scale=scale, res_f32=x, out_f32=x)
        out = x
    else:
        out = torch.full((M, N), float("nan"), device=cuda_device)
        ops.gemm_bf16(a_hi, w_hi, M, N, K, a_lo=a_lo, bw_lo=w_lo, bias=bias, out_f32=out)
    torch.cuda.synchronize()
    rel = (out.double() - ref).abs().max().item() / ref.abs().max().item()
    assert rel < 5e-5, rel


@pytest.mark.parametrize("M,C", [(3000, 192), (777, 96), (2500, 384), (1300, 768), (128, 96)])
def test_gemm_ln_gelu(cuda_device, M, C):
    """LayerNorm folded into fc1: gelu(fc1(LN(a))) from the un-normalised a and its partial sums."""
    g = torch.Generator().manual_seed(C)
    N = 4 * C
    a = (torch.randn(M, C, generator=g) * 1.5 + 0.7).to(cuda_device).to(torch.bfloat16)
    w1 = (torch.randn(N, C, generator=g) / C ** 0.5).to(cuda_device)
    b1 = torch.randn(N, generator=g).to(cuda_device)
    gamma = (torch.rand(C, generator=g) + 0.5).to(cuda_device)
    beta = torch.randn(C, generator=g).to(cuda_device)
    af = a.float().reshape(M, C // 32, 32)
    stats = torch.stack([af.sum(-1), (af * af).sum(-1)], dim=-1).contiguous()
    w1g = (w1 * gamma[None, :]).to(torch.bfloat16).contiguous()
    colsum = w1g.float().sum(1).contiguous()
    b1f = (b1 + w1 @ beta).contiguous()
    out = torch.empty(M, N, device=cuda_device, dtype=torch.bfloat16)
    ops.gemm_ln_gelu(a, w1g, M, N, C, stats, C // 32, colsum, b1f, out)
    ref = F.gelu(F.linear(F.layer_norm(a.float(), (C,), gamma, beta, 1e-6), w1, b1))
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err < 0.05 * max(1.0, ref.abs().max().item()) * 0.5, err


def _mlp_operands(M, C, device, seed):
    g = torch.Generator().manual_seed(seed)
    N = 4 * C
    y = (torch.randn(M, C, generator=g) * 1.5 + 0.7).to(device).to(torch.bfloat16)
    w1 = (torch.randn(N, C, generator=g) / C ** 0.5).to(device)
    b1 = torch.randn(N, generator=g).to(device)
    gamma = (torch.rand(C, generator=g) + 0.5).to(device)
    beta = torch.randn(C, generator=g).to(device)
    w2 = (torch.randn(C, N, generator=g) / N ** 0.5).to(device)
    b2 = torch.randn(C, generator=g).to(device)
    scale = (torch.rand(C, generator=g) + 0.25).to(device)
    x = torch.randn(M, C, generator=g).to(device)
    yf = y.float().reshape(M, C // 32, 32)
    stats = torch.stack([yf.sum(-1), (yf * yf).sum(-1)], dim=-1).contiguous()
    w1g = (w1 * gamma[None, :]).to(torch.bfloat16).contiguous()
    colsum = w1g.float().sum(1).contiguous()
    b1f = (b1 + w1 @ beta).contiguous()
    return dict(y=y, w1=w1, b1=b1, gamma=gamma, beta=beta, w2=w2, b2=b2, scale=scale, x=x, stats=stats, w1g=w1g,
                colsum=colsum, b1f=b1f, w2b=w2.to(torch.bfloat16).contiguous())


@pytest.mark.parametrize("cg", ["1", "2"])
@pytest.mark.parametrize("M,C", [(3000, 192), (777, 96), (2500, 384), (128, 96), (129, 192), (46656, 384), (50176, 192),
                                 (1, 96), (130, 384), (257, 192)])
def test_mlp_fused(cuda_device, M, C, cg, monkeypatch):
    """One-kernel CNBlock MLP against (a) the two-GEMM path of this library (same arithmetic: bf16 hidden activations,
    fp32 accumulation) and (b) torch fp32. Several tiles per CTA at the large sizes; ragged last tiles."""
    import subprocess, sys, os, json
    # CPN_MLP_CG is read once per process: run the comparison in a child so that both settings are covered
    code = f"""
import sys, json, torch
sys.path.insert(0, {os.path.dirname(os.path.abspath(__file__))!r})
import torch.nn.functional as F
from count_pipnet_b200 import ops
from test_kernels_gpu import _mlp_operands
M, C = {M}, {C}
d = torch.device('cuda:0')
o = _mlp_operands(M, C, d, C + M)
x1 = o['x'].clone()
ops.mlp_fused(o['y'], o['w1g'], o['w2b'], M, C, o['stats'], o['colsum'], o['b1f'], o['b2'], o['scale'], x1)
x1b = o['x'].clone()
ops.mlp_fused(o['y'], o['w1g'], o['w2b'], M, C, o['stats'], o['colsum'], o['b1f'], o['b2'], o['scale'], x1b)
h = torch.empty(M, 4 * C, device=d, dtype=torch.bfloat16)
ops.gemm_ln_gelu(o['y'], o['w1g'], M, 4 * C, C, o['stats'], C // 32, o['colsum'], o['b1f'], h)
x2 = o['x'].clone()
ops.gemm_bf16(h, o['w2b'], M, C, 4 * C, bias=o['b2'], scale=o['scale'], res_f32=x2, out_f32=x2)
ref = o['x'] + o['scale'] * F.linear(F.gelu(F.linear(F.layer_norm(o['y'].float(), (C,), o['gamma'], o['beta'], 1e-6),
                                                         o['w1'], o['b1'])), o['w2'], o['b2'])
torch.cuda.synchronize()
print(json.dumps(dict(vs_two_gemm=(x1 - x2).abs().max().item(), vs_torch=(x1 - ref).abs().max().item(),
                      two_gemm_vs_torch=(x2 - ref).abs().max().item(), repeat=bool(torch.equal(x1, x1b)),
                      scale=ref.abs().max().item())))
"""
    env = dict(os.environ, CPN_MLP_CG=cg, PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    print(res)
    assert res["repeat"]
    assert res["vs_two_gemm"] <= 2e-3 * max(1.0, res["scale"]), res
    assert res["vs_torch"] <= max(2.0 * res["two_gemm_vs_torch"], 1e-2 * max(1.0, res["scale"])), res


@pytest.mark.parametrize("stride,H,W,C", [(2, 16, 16, 96), (1, 14, 14, 192), (1, 9, 11, 384), (2, 7, 9, 96), (1, 3, 5, 192),
                                            (1, 2, 2, 384), (2, 10, 10, 64), (1, 6, 7, 768)])
def test_ln_im2col2(cuda_device, stride, H, W, C):
    g = torch.Generator().manual_seed(C)
    B = 2
    x = torch.randn(B, H, W, C, generator=g).to(cuda_device)
    lw = torch.rand(C, generator=g).to(cuda_device) + 0.5
    lb = torch.randn(C, generator=g).to(cuda_device)
    Ho, Wo = (H - 2) // stride + 1, (W - 2) // stride + 1
    out = torch.zeros(B * Ho * Wo, 4 * C, device=cuda_device, dtype=torch.bfloat16)
    ops.ln_im2col2(x, lw, lb, out, stride=stride)
    y = F.layer_norm(x, (C,), lw, lb, 1e-6)
    cols = []
    for kh in range(2):
        for kw in range(2):
            cols.append(y[:, kh:kh + stride * (Ho - 1) + 1:stride, kw:kw + stride * (Wo - 1) + 1:stride, :])
    ref = torch.cat(cols, dim=-1).reshape(B * Ho * Wo, 4 * C)
    torch.cuda.synchronize()
    assert (out.float() - ref).abs().max().item() < 0.03
    # fp32 mode: hi + lo planes
    hi = torch.zeros_like(out)
    lo = torch.zeros_like(out)
    ops.ln_im2col2(x, lw, lb, hi, lo, stride=stride)
    torch.cuda.synchronize()
    assert (hi.float() - ref).abs().max().item() < 0.03
    assert ((hi.float() + lo.float()) - ref).abs().max().item() < 1e-4


@pytest.mark.parametrize("mode", [_lib.POOL_MAX, _lib.POOL_SUM])
@pytest.mark.parametrize("P,HW,chunks", [(16, 100, 4), (768, 100, 4), (768, 333, 3), (128, 70, 2), (1024, 65, 1)])
def test_softmax_pool_and_map(cuda_device, mode, P, HW, chunks):
    g = torch.Generator().manual_seed(P)
    B = 3
    z = (torch.randn(B, HW, P, generator=g) * 3).to(cuda_device)
    partial = torch.empty(B, chunks, P, device=cuda_device)
    pmap = torch.empty(B, P, HW, device=cuda_device)
    ops.softmax_pool(z, partial, B, HW, P, chunks, mode=mode, map_out=pmap)
    sm = torch.softmax(z, dim=2)
    torch.cuda.synchronize()
    assert (pmap - sm.permute(0, 2, 1)).abs().max().item() < 1e-6
    if mode == _lib.POOL_MAX:
        assert (partial.amax(1) - sm.amax(1)).abs().max().item() < 1e-6
    else:
        assert (partial.sum(1) - sm.sum(1)).abs().max().item() < 1e-4


@pytest.mark.parametrize("mode", [_lib.POOL_MAX, _lib.POOL_SUM])
@pytest.mark.parametrize("P,HW,chunks", [(768, 101, 4), (128, 37, 3), (1024, 64, 5), (768, 676, 5), (2048, 50, 3)])
def test_softmax_pool_register_path(cuda_device, mode, P, HW, chunks):
    """Plain softmax without the map (register-resident kernel), odd token counts and ragged chunks."""
    g = torch.Generator().manual_seed(P + HW)
    B = 3
    z = (torch.randn(B, HW, P, generator=g) * 3).to(cuda_device)
    partial = torch.full((B, chunks, P), float("nan"), device=cuda_device)
    ops.softmax_pool(z, partial, B, HW, P, chunks, mode=mode)
    sm = torch.softmax(z, dim=2)
    torch.cuda.synchronize()
    if mode == _lib.POOL_MAX:
        assert (partial.amax(1) - sm.amax(1)).abs().max().item() < 1e-6
    else:
        assert (partial.sum(1) - sm.sum(1)).abs().max().item() < 1e-4


@pytest.mark.parametrize("with_map", [False, True])
@pytest.mark.parametrize("P,HW,chunks", [(768, 101, 4), (128, 37, 3), (1024, 64, 5), (768, 676, 5), (16, 49, 2), (2048, 50, 3)])
def test_softmax_pool_argmax(cuda_device, with_map, P, HW, chunks):
    """Token of each prototype's spatial maximum (both softmax kernels), bit-exact incl. ties -> lowest token."""
    g = torch.Generator().manual_seed(P + HW + 1)
    B = 3
    z = (torch.randn(B, HW, P, generator=g) * 3)
    z[0, HW - 1] = z[0, 0]          # duplicated tokens: exact ties across warps and chunks
    z[1, HW // 2] = z[1, HW // 2 - 1]
    z[2, :, 0] = -200.0             # a prototype whose probability underflows to 0 everywhere -> token 0
    z = z.to(cuda_device)
    partial = torch.full((B, chunks, P), float("nan"), device=cuda_device)
    parg = torch.full((B, chunks, P), -1, dtype=torch.int32, device=cuda_device)
    arg = torch.full((B, P), -1, dtype=torch.int32, device=cuda_device)
    pmap = torch.empty(B, P, HW, device=cuda_device) if with_map else None
    ops.softmax_pool(z, partial, B, HW, P, chunks, mode=_lib.POOL_MAX, map_out=pmap, partial_arg=parg)
    ops.pool_argmax_reduce(partial, parg, chunks, arg, B, P)
    torch.cuda.synchronize()
    # reference on the kernel's own probabilities where the map exists, else on torch's softmax away from near-ties
    sm = pmap if with_map else torch.softmax(z, dim=2).permute(0, 2, 1)
    ref_val, ref_arg = sm.max(dim=2)
    is_max = sm == ref_val[:, :, None]
    first = torch.where(is_max, torch.arange(HW, device=cuda_device)[None, None, :], HW).amin(2)
    if with_map:
        assert torch.equal(arg.long(), first)
    else:
        srt = sm.sort(dim=2).values
        clear = (srt[:, :, -1] - srt[:, :, -2]) > 1e-5 * srt[:, :, -1].clamp_min(1e-30)
        assert torch.equal(arg.long()[clear], first[clear])
        # the planted exact ties must resolve to the lower token
        picked = torch.softmax(z, dim=2).permute(0, 2, 1).gather(2, arg.long()[:, :, None])[:, :, 0]
        assert (picked - ref_val).abs().max().item() < 1e-6
        assert not (arg[0] == HW - 1).any() and not (arg[1] == HW // 2).any()
    assert int(arg[2, 0]) == 0
    assert (partial.amax(1) - ref_val).abs().max().item() < 1e-6


def test_softmax_pool_gumbel_hard(cuda_device):
    g = torch.Generator().manual_seed(9)
    B, HW, P, chunks = 2, 64, 16, 2
    z = torch.randn(B, HW, P, generator=g).to(cuda_device)
    noise = -torch.empty(B, P, HW).exponential_(generator=g).log()
    noise = noise.to(cuda_device)
    partial = torch.empty(B, chunks, P, device=cuda_device)
    pmap = torch.empty(B, P, HW, device=cuda_device)
    ops.softmax_pool(z, partial, B, HW, P, chunks, gumbel=noise, tau=0.7, hard=True, mode=_lib.POOL_SUM, map_out=pmap)
    logits = (z.permute(0, 2, 1) + noise) / 0.7
    y_soft = logits.softmax(1)
    idx = y_soft.max(1, keepdim=True)[1]
    y_hard = torch.zeros_like(y_soft).scatter_(1, idx, 1.0)
    ref = y_hard - y_soft + y_soft
    torch.cuda.synchronize()
    assert (pmap - ref).abs().max().item() < 1e-6
    assert (partial.sum(1) - ref.sum(2)).abs().max().item() < 1e-4


def test_addon_conv(cuda_device):
    g = torch.Generator().manual_seed(10)
    T, C, P = 1000, 192, 16
    f = torch.randn(T, C, generator=g).to(cuda_device)
    w = torch.randn(P, C, generator=g).to(cuda_device) * 0.1
    b = torch.randn(P, generator=g).to(cuda_device)
    z = torch.empty(T, P, device=cuda_device)
    ops.addon_conv1x1(f, w, b, z, T, C, P)
    torch.cuda.synchronize()
    assert (z - (f @ w.t() + b)).abs().max().item() < 1e-4


@pytest.mark.parametrize("mode", [_lib.POOL_MAX, _lib.POOL_SUM])
@pytest.mark.parametrize("B,HW,C,P,chunks,gum", [(3, 576, 192, 16, 7, True), (2, 100, 768, 32, 3, False),
                                                  (2, 37, 384, 16, 5, False), (1, 1000, 192, 16, 4, True)])
def test_addon_head_fused(cuda_device, mode, B, HW, C, P, chunks, gum):
    """Add-on conv + softmax / hard Gumbel + pooling in one kernel == conv, softmax, pool in torch (map, partial sums /
    maxima reduced over chunks, arg-max token with the lowest token on ties)."""
    g = torch.Generator().manual_seed(HW + P)
    f = torch.randn(B * HW, C, generator=g).to(cuda_device)
    w = (torch.randn(P, C, generator=g) * 0.2).to(cuda_device)
    b = torch.randn(P, generator=g).to(cuda_device)
    noise = (-torch.empty(B, P, HW).exponential_(generator=g).log()).to(cuda_device) if gum else None
    partial = torch.full((B, chunks, P), float("nan"), device=cuda_device)
    parg = torch.full((B, chunks, P), -1, device=cuda_device, dtype=torch.int32)
    pmap = torch.full((B, P, HW), float("nan"), device=cuda_device)
    ops.addon_softmax_pool(f, w, b, partial, B, HW, C, P, chunks, gumbel=noise, tau=1.0, hard=gum, mode=mode, map_out=pmap,
                           partial_arg=parg if mode == _lib.POOL_MAX else None)
    z = (f.double() @ w.double().t() + b.double()).reshape(B, HW, P).permute(0, 2, 1)  # [B, P, HW]
    if gum:
        y_soft = torch.softmax(z + noise.double(), dim=1)
        idx = y_soft.argmax(dim=1, keepdim=True)
        y = torch.zeros_like(y_soft).scatter_(1, idx, 1.0)
    else:
        y = torch.softmax(z, dim=1)
    torch.cuda.synchronize()
    assert (pmap.double() - y).abs().max().item() < 2e-5
    if mode == _lib.POOL_SUM:
        assert (partial.double().sum(1) - y.sum(2)).abs().max().item() < 1e-3
    else:
        assert (partial.double().amax(1) - y.amax(2)).abs().max().item() < 2e-5
        # the arg-max token of the best chunk holds the maximum of this library's own map
        best = partial.argmax(1, keepdim=True)
        tok = parg.gather(1, best)[:, 0, :].long()
        picked = pmap.gather(2, tok[:, :, None])[:, :, 0]
        assert (picked - pmap.amax(2)).abs().max().item() <= 1e-6


@pytest.mark.parametrize("B,H,W,Cin,Cout,k,kind", [
    (2, 28, 28, 128, 128, 3, "relu"), (3, 56, 56, 64, 64, 3, "relu"), (2, 28, 28, 512, 512, 3, "relu"),
    (32, 28, 28, 256, 256, 3, "relu"), (2, 28, 28, 256, 1024, 1, "res"), (32, 28, 28, 512, 2048, 1, "res"),
    (2, 56, 56, 64, 256, 1, "bias"), (5, 28, 28, 1024, 512, 1, "relu"), (3, 56, 56, 64, 256, 1, "res"),
    (1, 28, 28, 64, 64, 3, "res"), (7, 8, 16, 64, 128, 3, "bias")])
def test_conv_bn_act(cuda_device, B, H, W, Cin, Cout, k, kind):
    """ResNet conv + folded BN as one GEMM (implicit 3x3 taps, TMA epilogue) vs F.conv2d on the same bf16 operands."""
    g = torch.Generator().manual_seed(Cin + Cout + k)
    x = torch.randn(B, H, W, Cin, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(cuda_device)
    wk = w.permute(0, 2, 3, 1).reshape(Cout, k * k * Cin).contiguous()  # K ordered (kh, kw, cin)
    M = B * H * W
    if k == 3:
        assert ops.conv3x3_tile_rows(H, W) > 0
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=k // 2).permute(0, 2, 3, 1).reshape(M, Cout)
    out32 = torch.full((M, Cout), float("nan"), device=cuda_device)
    out16 = torch.full((M, Cout), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    if kind == "relu":
        ops.conv_bn_act(x, wk, bias, B, H, W, Cin, Cout, k, relu=True, out_bf16=out16)
        ref = torch.relu(ref)
    elif kind == "res":
        res = torch.randn(M, Cout, generator=g).to(cuda_device)
        ops.conv_bn_act(x, wk, bias, B, H, W, Cin, Cout, k, relu=True, res_f32=res, out_f32=out32, out_bf16=out16)
        ref = torch.relu(ref + res)
    else:
        ops.conv_bn_act(x, wk, bias, B, H, W, Cin, Cout, k, relu=False, out_f32=out32)
    torch.cuda.synchronize()
    scale = ref.abs().max().item()
    if kind != "relu":
        assert (out32 - ref).abs().max().item() <= 2e-3 * scale  # fp32 accumulation order only
    if kind != "bias":
        assert (out16.float() - ref).abs().max().item() <= 2.0 ** -8 * scale + 2e-3 * scale
    if kind == "res":
        assert torch.equal(out16, out32.to(torch.bfloat16))


@pytest.mark.parametrize("B,H,W,Cin,Cout,k,kind", [
    (2, 28, 28, 256, 1024, 1, "resb"), (32, 28, 28, 512, 2048, 1, "resb"), (3, 56, 56, 64, 256, 1, "resb"),
    (1, 28, 28, 64, 64, 3, "resb"), (2, 28, 28, 512, 2048, 1, "resb_f32"), (3, 8, 16, 64, 128, 3, "resb_f32"),
    (2, 28, 28, 256, 512, 1, "bias_bf16"), (5, 28, 28, 1024, 2048, 1, "bias_bf16"), (2, 7, 7, 512, 512, 1, "resb")])
def test_conv_bn_act_bf16_stream(cuda_device, B, H, W, Cin, Cout, k, kind):
    """Block end on the bf16 residual stream (identity read as bf16, bf16 or fp32 result) and the bf16 projection
    shortcut, vs F.conv2d on the same bf16 operands."""
    g = torch.Generator().manual_seed(Cin + Cout + k + 1)
    x = torch.randn(B, H, W, Cin, generator=g).to(cuda_device).to(torch.bfloat16)
    w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(cuda_device).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(cuda_device)
    wk = w.permute(0, 2, 3, 1).reshape(Cout, k * k * Cin).contiguous()
    M = B * H * W
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=k // 2).permute(0, 2, 3, 1).reshape(M, Cout)
    out32 = torch.full((M, Cout), float("nan"), device=cuda_device)
    out16 = torch.full((M, Cout), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    if kind == "bias_bf16":
        ops.conv_bn_act(x, wk, bias, B, H, W, Cin, Cout, k, relu=False, out_bf16=out16)
        assert (ref < 0).any()  # the un-activated result keeps its negative values
    else:
        res = torch.randn(M, Cout, generator=g).to(cuda_device).to(torch.bfloat16)
        ref = torch.relu(ref + res.float())
        if kind == "resb":
            ops.conv_bn_act_resb(x, wk, bias, res, B, H, W, Cin, Cout, k, out_bf16=out16)
        else:
            ops.conv_bn_act_resb(x, wk, bias, res, B, H, W, Cin, Cout, k, out_f32=out32)
    torch.cuda.synchronize()
    scale = ref.abs().max().item()
    if kind == "resb_f32":
        assert (out32 - ref).abs().max().item() <= 2e-3 * scale
        assert torch.isnan(out16.float()).all()
    else:
        assert (out16.float() - ref).abs().max().item() <= 2.0 ** -8 * scale + 2e-3 * scale
        assert torch.isnan(out32).all()


@pytest.mark.parametrize("B,H,W,C", [(2, 112, 112, 64), (1, 37, 51, 64), (3, 8, 8, 128)])
def test_maxpool3s2_bf16(cuda_device, B, H, W, C):
    """bf16 pooling == bf16(fp32 pooling): rounding is monotonic (ResNet stem on the bf16 stream)."""
    g = torch.Generator().manual_seed(H + C)
    x32 = torch.randn(B, H, W, C, generator=g).to(cuda_device)
    xb = x32.to(torch.bfloat16)
    Ho, Wo = (H + 2 - 3) // 2 + 1, (W + 2 - 3) // 2 + 1
    out = torch.full((B * Ho * Wo, C), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    ops.maxpool3s2_bf16(xb, out, B, H, W, C)
    ref = F.max_pool2d(x32.permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1).reshape(B * Ho * Wo, C)
    torch.cuda.synchronize()
    assert torch.equal(out, ref.to(torch.bfloat16))


def test_conv3x3_tile_rows():
    assert ops.conv3x3_tile_rows(28, 28) == 4 and ops.conv3x3_tile_rows(56, 56) == 2
    assert ops.conv3x3_tile_rows(14, 14) == 0 and ops.conv3x3_tile_rows(8, 16) == 8


@pytest.mark.parametrize("B,H,W", [(2, 64, 64), (1, 37, 51), (3, 224, 224)])
def test_im2col_nchw_stem(cuda_device, B, H, W):
    """ResNet stem gather (conv 7x7, stride 2, pad 3 on the fp32 NCHW input) == F.unfold, K = 147 padded to 152."""
    g = torch.Generator().manual_seed(H)
    x = torch.randn(B, 3, H, W, generator=g).to(cuda_device)
    Ho, Wo = (H + 6 - 7) // 2 + 1, (W + 6 - 7) // 2 + 1
    hi = torch.full((B * Ho * Wo, 152), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    lo = torch.full((B * Ho * Wo, 152), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    ops.im2col_nchw_f32(x, hi, B, 3, H, W, 7, 2, 3, 152, out_lo=lo)
    ref = F.unfold(x, kernel_size=7, stride=2, padding=3).transpose(1, 2).reshape(B * Ho * Wo, 147)
    torch.cuda.synchronize()
    assert torch.equal(hi[:, :147], ref.to(torch.bfloat16))
    assert ((hi[:, :147].float() + lo[:, :147].float()) - ref).abs().max().item() < 1e-4
    assert (hi[:, 147:] == 0).all() and (lo[:, 147:] == 0).all()


def test_transposes(cuda_device):
    x = torch.randn(2, 50, 70, device=cuda_device)
    out = torch.empty(2, 70, 50, device=cuda_device)
    ops.nhwc_to_nchw(x, out, 2, 50, 70)
    torch.cuda.synchronize()
    assert torch.equal(out, x.permute(0, 2, 1).contiguous())
    back = torch.empty_like(x)
    ops.nchw_to_nhwc(out, back, 2, 50, 70)
    torch.cuda.synchronize()
    assert torch.equal(back, x)
