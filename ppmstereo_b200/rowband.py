"""Optional row-band split of ONE clip over several GPUs (north_star: "NCCL over
NVLink used only for an optional row-band split of a single high-resolution clip").

Every rank owns a contiguous band of image rows at each cascade level:
  * correlation build and lookup are strictly per image row (models/core/corr.py:85-89, :102): a rank
    simply builds and samples the rows of its band -- no halo, no exchange;
  * the memory read needs the keys and values of ALL pixels of the picked frames, so per iteration the
    value bands (and the uncertainty bands that feed the frame score) are all-gathered; the QAM pick is
    then computed redundantly and deterministically on every rank (bit-identical frame sets), and each
    rank aggregates only the queries of its own band (ppm_memory_attend_band).
The exchange is T*hw*128 values per iteration (301 MB in bf16 at 720p 1/4 res) against ~150 ms of
attention per GPU, i.e. well under 1 % of the step over NVLink.
``gather_bands`` is backend-agnostic (NCCL on GPUs; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from .memory import PPMMemory


def row_bands(h: int, world: int) -> List[Tuple[int, int]]:
    """[r0, r1) per rank: contiguous, sizes differ by at most one row, larger bands first."""
    if world <= 0 or h < 0:
        raise ValueError("world must be > 0 and h >= 0")
    base, extra = divmod(h, world)
    out, r = [], 0
    for k in range(world):
        n = base + (1 if k < extra else 0)
        out.append((r, r + n))
        r += n
    return out


def gather_bands(band: torch.Tensor, h: int, dim: int, group=None) -> torch.Tensor:
    """All-gather row bands along ``dim`` into the full-height tensor.  ``band`` is this rank's slice
    (rows row_bands(h, world)[rank]); bands may differ by one row (padded for the collective)."""
    if not (dist.is_available() and dist.is_initialized()):
        if band.shape[dim] != h:
            raise RuntimeError("no process group: the band must be the whole tensor")
        return band
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    bands = row_bands(h, world)
    r0, r1 = bands[rank]
    if band.shape[dim] != r1 - r0:
        raise RuntimeError(f"rank {rank} owns rows [{r0}, {r1}) but got {band.shape[dim]} rows")
    hb = max(b1 - b0 for b0, b1 in bands)
    moved = band.movedim(dim, 0).contiguous()                       # rows first: one contiguous block per rank
    if moved.shape[0] < hb:
        moved = torch.cat([moved, moved.new_zeros((hb - moved.shape[0],) + tuple(moved.shape[1:]))], 0)
    gathered = moved.new_empty((world * hb,) + tuple(moved.shape[1:]))
    dist.all_gather_into_tensor(gathered, moved, group=group)
    parts = [gathered[k * hb: k * hb + (b1 - b0)] for k, (b0, b1) in enumerate(bands)]
    return torch.cat(parts, 0).movedim(0, dim).contiguous()


class RowBandMemory:
    """PPM read of one rank's row band.  ``query`` / ``key`` are full-frame (b, c, T, h, w) tensors -- they
    are produced once per level (gather them with ``gather_bands`` if the projections ran per band)."""

    def __init__(self, query: torch.Tensor, key: torch.Tensor, pe: Optional[torch.Tensor] = None,
                 top_k: int = 5, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.mem = PPMMemory(query, key, pe, top_k)
        self.h, self.w = self.mem.h, self.mem.w
        self.r0, self.r1 = row_bands(self.h, self.world)[self.rank]
        self._mf_full = None

    def read(self, value_band, motion_features_band, sim, strive, uncertainty_band, beta, scale=None):
        """value_band, motion_features_band: (b, c, T, rows, w); uncertainty_band: (b*T, 1, rows, w).
        Returns (motion_features_global of the band (b*T, c, rows, w) fp32, topk mask)."""
        m = self.mem
        value = gather_bands(value_band, self.h, 3, self.group)
        unc = gather_bands(uncertainty_band, self.h, 2, self.group)
        mask = m.pick(sim, strive, unc)                    # replicated and deterministic on every rank
        m.pack_kv(value)
        if self._mf_full is None or self._mf_full.dtype != motion_features_band.dtype:
            self._mf_full = motion_features_band.new_zeros(value.shape)   # only the band rows are ever read
        full_mf = self._mf_full
        full_mf[:, :, :, self.r0:self.r1] = motion_features_band
        out = m.aggregate(full_mf, beta, scale, band=(self.r0 * self.w, (self.r1 - self.r0) * self.w))
        return out[:, :, self.r0:self.r1], mask
