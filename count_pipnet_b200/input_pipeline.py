"""GPU version of the reference's evaluation-time input transform.

The reference decodes an image with PIL and applies `transform_no_augment = Compose([Resize(size=(S, S)), ToTensor(),
Normalize(mean, std)])` in DataLoader worker processes (util/data.py:381-385 for the shapes dataset; the same three
steps at :265-275, :296-307, :465-480, :522-531; applied in TwoAugSupervisedDataset / ImageFolder, :596-617). This module
runs those three steps for a batch of decoded uint8 images on the device, bit-exact with Pillow + torchvision
(csrc/api_preprocess.cu), so that the forward's input can be produced at the forward's rate.

    tf = EvalTransform(224)                       # mean / std default to the ImageNet values every dataset getter uses
    x = tf(batch_u8)                              # [B, Hs, Ws, 3] uint8 cuda -> [B, 3, 224, 224] float32 cuda
    proto, pooled, out = net(x, inference=True)

There is no CPU fallback: a CPU tensor raises (the CPU path is the reference's own torchvision Compose).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib

IMAGENET_MEAN = (0.485, 0.456, 0.406)  # util/data.py:377-378
IMAGENET_STD = (0.229, 0.224, 0.225)


class EvalTransform:
    """Resize((size, size)) + ToTensor() + Normalize(mean, std) on uint8 channels-last batches."""

    def __init__(self, size, mean: Sequence[float] = IMAGENET_MEAN, std: Sequence[float] = IMAGENET_STD):
        self.size: Tuple[int, int] = (int(size), int(size)) if isinstance(size, int) else (int(size[0]), int(size[1]))
        if len(mean) != 3 or len(std) != 3:
            raise ValueError("mean / std need three values (RGB)")
        # fp32 values exactly as torchvision's Normalize builds them (torch.as_tensor(mean, dtype=float32))
        self._mean = (C.c_float * 3)(*[float(torch.tensor(m, dtype=torch.float32)) for m in mean])
        self._std = (C.c_float * 3)(*[float(torch.tensor(s, dtype=torch.float32)) for s in std])
        self._ws: Dict[Tuple[int, int, int], torch.Tensor] = {}

    def _workspace(self, device: torch.device, hs: int, ws: int) -> torch.Tensor:
        key = (device.index if device.index is not None else torch.cuda.current_device(), hs, ws)
        t = self._ws.get(key)
        if t is None:
            n = int(_lib.load().cpn_preprocess_ws_ints(hs, ws, self.size[0], self.size[1]))
            if n <= 0:
                raise _lib.CpnError(f"cpn_preprocess_ws_ints({hs}, {ws}, {self.size}) failed")
            t = torch.empty(n, dtype=torch.int32, device=device)
            self._ws[key] = t
        return t

    def __call__(self, batch_u8: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if batch_u8.dtype != torch.uint8 or batch_u8.dim() != 4 or batch_u8.shape[-1] != 3:
            raise ValueError("expected a [B, H, W, 3] uint8 batch (decoded RGB, channels last)")
        if not batch_u8.is_cuda:
            raise _lib.CpnError("EvalTransform runs on the B200 only; on the CPU use the reference's torchvision Compose")
        x = batch_u8.contiguous()
        B, hs, ws = int(x.shape[0]), int(x.shape[1]), int(x.shape[2])
        ho, wo = self.size
        if out is None:
            out = torch.empty(B, 3, ho, wo, dtype=torch.float32, device=x.device)
        elif out.shape != (B, 3, ho, wo) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != x.device:
            raise ValueError("out must be a contiguous float32 [B, 3, H, W] tensor on the batch's device")
        with torch.cuda.device(x.device):
            work = self._workspace(x.device, hs, ws)
            _lib.call("cpn_preprocess_u8", None, x.data_ptr(), out.data_ptr(), work.data_ptr(), B, hs, ws, ho, wo, self._mean,
                      self._std, torch.cuda.current_stream().cuda_stream)
        return out
