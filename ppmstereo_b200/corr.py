"""Drop-in for the reference's ``models.core.corr.CorrBlock1D`` (models/core/corr.py:55-104).

Same constructor, attributes (``num_levels``, ``radius``, ``corr_pyramid``, ``coords``),
``__call__(flow)`` and static ``corr(fmap1, fmap2)`` surface; the arithmetic runs in
libppm_sm100.so (tcgen05 + TMA pyramid build, HBM-bound lookup).  CUDA tensors only.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn.functional as F

from . import _lib


def _check_fmap(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (ppmstereo_b200 has no CPU path)")
    if t.dim() != 4:
        raise RuntimeError(f"{name} must be (B, D, H, W), got {tuple(t.shape)}")
    if t.dtype not in (torch.float32, torch.bfloat16):
        raise RuntimeError(f"{name}: dtype {t.dtype} not supported (float32 or bfloat16)")


def _pitched(t: torch.Tensor, mult: int) -> torch.Tensor:
    """Contiguous NCHW with the row pitch padded to a multiple of ``mult`` (TMA needs 16-byte pitches)."""
    t = t.contiguous()
    pad = (-t.shape[-1]) % mult
    if pad or t.data_ptr() % 16:
        t = F.pad(t, (0, pad)) if pad else t.clone()
    return t


def build_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4,
                  out_dtype: Optional[torch.dtype] = None,
                  levels: Optional[List[torch.Tensor]] = None) -> List[torch.Tensor]:
    """ppm_corr_build: returns ``num_levels + 1`` tensors (B*H*W1, 1, 1, W2 >> l).  ``levels`` may pass
    buffers of exactly those shapes to be overwritten (no allocation on the call path)."""
    _check_fmap(fmap1, "fmap1")
    _check_fmap(fmap2, "fmap2")
    if fmap1.dtype != fmap2.dtype or fmap1.shape[:3] != fmap2.shape[:3]:
        raise RuntimeError("fmap1 and fmap2 must agree in dtype and (B, D, H)")
    B, D, H, W1 = fmap1.shape
    W2 = fmap2.shape[3]
    if (W2 >> num_levels) < 1:
        raise RuntimeError("Given input size: W2 too small for the pyramid: Output size is too small")
    out_dtype = out_dtype or fmap1.dtype
    mult = 8 if fmap1.dtype == torch.bfloat16 else 4
    f1, f2 = _pitched(fmap1, mult), _pitched(fmap2, mult)
    shapes = [(B * H * W1, 1, 1, W2 >> l) for l in range(num_levels + 1)]
    if levels is None:
        levels = [torch.empty(s, dtype=out_dtype, device=fmap1.device) for s in shapes]
    elif [tuple(t.shape) for t in levels] != shapes or any(
            t.dtype != out_dtype or not t.is_contiguous() or t.device != fmap1.device for t in levels):
        raise RuntimeError("levels must be contiguous tensors of shapes %s and dtype %s" % (shapes, out_dtype))
    lib = _lib.load()
    code_out = _lib.PPM_BF16 if out_dtype == torch.bfloat16 else _lib.PPM_F32
    _lib.check(lib.ppm_corr_build(f1.data_ptr(), f2.data_ptr(), _lib.dtype_code(f1), B, D, H, W1, W2,
                                  f1.shape[3], f2.shape[3], _lib.ptr_array(levels), num_levels, code_out,
                                  _lib.stream_ptr()), "ppm_corr_build")
    return levels


def lookup(levels: List[torch.Tensor], flow: torch.Tensor, W2: int, radius: int = 4,
           num_levels: int = 4) -> torch.Tensor:
    """ppm_corr_lookup: (B, num_levels*(2r+1), H, W1) fp32."""
    if not flow.is_cuda or flow.dim() != 4 or flow.shape[1] != 2:
        raise RuntimeError(f"flow must be a CUDA tensor (B, 2, H, W), got {tuple(flow.shape)}")
    flow = flow.float().contiguous()
    B, _, H, W1 = flow.shape
    out = torch.empty(B, num_levels * (2 * radius + 1), H, W1, dtype=torch.float32, device=flow.device)
    lib = _lib.load()
    _lib.check(lib.ppm_corr_lookup(_lib.ptr_array(levels[:num_levels]), _lib.dtype_code(levels[0]),
                                   flow.data_ptr(), B, H, W1, W2, radius, num_levels, out.data_ptr(),
                                   _lib.stream_ptr()), "ppm_corr_lookup")
    return out


def lookup_l0(level0: torch.Tensor, flow: torch.Tensor, W2: int) -> torch.Tensor:
    """ppm_corr_lookup_l0: radius-4, 4-level lookup of a bf16 pyramid sampled from its level 0 only (the
    pooled taps are rebuilt in registers; equal to ``lookup`` on a pyramid written by ``build_pyramid``)."""
    if not flow.is_cuda or flow.dim() != 4 or flow.shape[1] != 2:
        raise RuntimeError(f"flow must be a CUDA tensor (B, 2, H, W), got {tuple(flow.shape)}")
    if level0.dtype != torch.bfloat16 or W2 % 8:
        raise RuntimeError("lookup_l0 needs a bf16 level 0 and W2 % 8 == 0")
    flow = flow.float().contiguous()
    B, _, H, W1 = flow.shape
    out = torch.empty(B, 36, H, W1, dtype=torch.float32, device=flow.device)
    lib = _lib.load()
    _lib.check(lib.ppm_corr_lookup_l0(level0.data_ptr(), flow.data_ptr(), B, H, W1, W2, out.data_ptr(),
                                      _lib.stream_ptr()), "ppm_corr_lookup_l0")
    return out


class CorrBlock1D:
    """Same surface as the reference class (models/core/corr.py:55)."""

    def __init__(self, fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4, radius: int = 4):
        self.num_levels = num_levels
        self.radius = radius
        B, _, H, W1 = fmap1.shape
        self._w2 = fmap2.shape[3]
        self._grid = (B, H, W1, fmap1.device)
        self._coords = None
        # The lookup the model performs (4 levels, r = 4, bf16 pyramid) reads level 0 only (K2b rebuilds the pooled
        # taps in registers, bit-identically).  Levels 1-4 -- 48 % of the pyramid's bytes -- are then written for
        # nobody: build level 0 now and the pooled levels only if someone asks for `corr_pyramid`.
        self._l0_only = (num_levels == 4 and radius == 4 and fmap1.dtype == torch.bfloat16 and self._w2 % 8 == 0)
        if self._l0_only:
            self._level0 = build_pyramid(fmap1.detach(), fmap2.detach(), 0)[0]
            self._l0_only = self._level0.data_ptr() % 32 == 0
        self._pyramid = None if self._l0_only else self._full_pyramid(fmap1, fmap2)
        self._fmaps = (fmap1.detach(), fmap2.detach()) if self._l0_only else None

    def _full_pyramid(self, fmap1, fmap2) -> List[torch.Tensor]:
        return build_pyramid(fmap1.detach(), fmap2.detach(), self.num_levels)

    @property
    def corr_pyramid(self) -> List[torch.Tensor]:
        """The reference's attribute: num_levels + 1 tensors (B*H*W1, 1, 1, W2 >> l); pooled levels are materialised
        on first access (same kernel, same rounding: level 0 is rewritten with identical bits)."""
        if self._pyramid is None:
            self._pyramid = self._full_pyramid(*self._fmaps)
            self._level0 = self._pyramid[0]
        return self._pyramid

    @property
    def coords(self) -> torch.Tensor:
        """(B, 2, H, W1) pixel (x, y) grid, as coords_grid at corr.py:47-52.  Nothing on the path reads it
        (the lookup kernel derives x from the thread index), so it is built on first access only."""
        if self._coords is None:
            B, H, W1, dev = self._grid
            ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W1, device=dev), indexing="ij")
            self._coords = torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1)
        return self._coords

    def __call__(self, flow: torch.Tensor) -> torch.Tensor:
        if self._l0_only:
            return lookup_l0(self._level0, flow.detach(), self._w2)    # K2b: all levels from one level-0 span
        return lookup(self.corr_pyramid, flow.detach(), self._w2, self.radius, self.num_levels)

    @staticmethod
    def corr(fmap1: torch.Tensor, fmap2: torch.Tensor) -> torch.Tensor:
        B, _, H, W1 = fmap1.shape
        return build_pyramid(fmap1, fmap2, 0)[0].reshape(B, H, W1, 1, fmap2.shape[3])
