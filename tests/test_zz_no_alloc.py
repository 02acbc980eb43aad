"""cpn_softmax_pool_static (include/cpn_b200.h): the allocation-free softmax + pool, checked through the C-ABI against
cpn_softmax_pool with caller-owned scratch, on two calls in a row (the static scratch keeps the first call's contents).

Collected last on purpose (the file name sorts after every other test file)."""
from __future__ import annotations

import ctypes as C

import pytest
import torch

from count_pipnet_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_BYTES = 4 << 20  # CPN_SOFTMAX_POOL_STATIC_MAX_BYTES


def _existing(z, mode, B, HW, P):
    """cpn_softmax_pool with one chunk: partial [B,1,P] is the pooled result."""
    partial = torch.full((B, 1, P), float("nan"), device=z.device)
    _lib.call("cpn_softmax_pool", None, z.data_ptr(), None, C.c_float(1.0), 0, mode, None, partial.data_ptr(), B, HW, P,
              1, None)
    return partial[:, 0]


def _static(z, mode, B, HW, P, pooled):
    _lib.call("cpn_softmax_pool_static", z.data_ptr(), mode, pooled.data_ptr(), B, HW, P, None)


@pytest.mark.parametrize("mode", [_lib.POOL_MAX, _lib.POOL_SUM])
def test_softmax_pool_static_two_calls(cuda_device, mode):
    g = torch.Generator().manual_seed(1234 + mode)
    tol = 1e-6 if mode == _lib.POOL_MAX else 1e-4  # the tolerances of tests/test_kernels_gpu.py::test_softmax_pool_*
    # second call smaller and with another chunk count: stale entries of the first call stay in the scratch around it
    shapes = [(5, 729, 768), (3, 50, 2048), (2, 37, 16)]
    outs = []
    for B, HW, P in shapes:
        z = (torch.randn(B, HW, P, generator=g) * 3).to(cuda_device)
        pooled = torch.full((B, P), float("nan"), device=cuda_device)
        mem_before = torch.cuda.memory_allocated()
        _static(z, mode, B, HW, P, pooled)
        assert torch.cuda.memory_allocated() == mem_before
        outs.append((z, pooled, B, HW, P))
    torch.cuda.synchronize()
    for z, pooled, B, HW, P in outs:
        ref = _existing(z, mode, B, HW, P)
        torch.cuda.synchronize()
        assert torch.isfinite(pooled).all()
        assert (pooled - ref).abs().max().item() < tol
        sm = torch.softmax(z, dim=2)
        want = sm.amax(1) if mode == _lib.POOL_MAX else sm.sum(1)
        assert (pooled - want).abs().max().item() < tol


def test_softmax_pool_static_same_input_twice(cuda_device):
    """Two calls in a row on the same input give the same bits (deterministic fold, nothing left over in the scratch)."""
    g = torch.Generator().manual_seed(77)
    B, HW, P = 4, 196, 768
    z = (torch.randn(B, HW, P, generator=g) * 3).to(cuda_device)
    a = torch.empty(B, P, device=cuda_device)
    b = torch.empty(B, P, device=cuda_device)
    _static(z, _lib.POOL_SUM, B, HW, P, a)
    _static(z, _lib.POOL_SUM, B, HW, P, b)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert (a - _existing(z, _lib.POOL_SUM, B, HW, P)).abs().max().item() < 1e-4


def test_softmax_pool_static_rejects_oversize(cuda_device):
    z = torch.zeros(8, device=cuda_device)
    out = torch.zeros(8, device=cuda_device)
    B, P = 1024, 2048  # B * P * 4 bytes = 8 MiB > the 4 MiB scratch
    assert B * P * 4 > MAX_BYTES
    with pytest.raises(_lib.CpnError, match="exceeds the static scratch"):
        _static(z, _lib.POOL_MAX, B, 1, P, out)
    with pytest.raises(_lib.CpnError, match="3072"):
        _static(z, _lib.POOL_MAX, 1, 1, 4096, out)
