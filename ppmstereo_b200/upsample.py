"""Convex upsampling of the flow field on the GPU path.

Same contract as ``PPMStereo.convex_upsample(flow, mask, rate=4)`` (reference models/core/ppmstereo.py:185-197):
flow (N, 2, H, W), mask (N, 9*rate*rate, H, W) -> (N, 2, rate*H, rate*W) fp32.  One kernel
(csrc/convex_upsample.cu) instead of view/softmax/unfold/mul/sum/permute.  CUDA only.
"""
from __future__ import annotations

import torch

from . import _lib


def convex_upsample(flow: torch.Tensor, mask: torch.Tensor, rate: int = 4) -> torch.Tensor:
    if not (flow.is_cuda and mask.is_cuda):
        raise RuntimeError("convex_upsample needs CUDA tensors (ppmstereo_b200 has no CPU path)")
    if flow.dim() != 4 or flow.shape[1] != 2:
        raise RuntimeError(f"flow must be (N, 2, H, W), got {tuple(flow.shape)}")
    N, _, H, W = flow.shape
    if tuple(mask.shape) != (N, 9 * rate * rate, H, W):
        raise RuntimeError(f"mask must be ({N}, {9 * rate * rate}, {H}, {W}), got {tuple(mask.shape)}")
    flow = flow.float().contiguous()
    if mask.dtype not in (torch.float32, torch.bfloat16):
        mask = mask.float()
    mask = mask.contiguous()
    out = torch.empty(N, 2, rate * H, rate * W, dtype=torch.float32, device=flow.device)
    lib = _lib.load()
    _lib.check(lib.ppm_convex_upsample(flow.data_ptr(), mask.data_ptr(), _lib.dtype_code(mask), N, H, W, rate,
                                       out.data_ptr(), _lib.stream_ptr()), "ppm_convex_upsample")
    return out
