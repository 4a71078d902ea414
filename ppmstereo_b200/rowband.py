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


def verify_against_full_read(device, shape=(1, 7, 21, 40), iterations: int = 3, seed: int = 5, group=None) -> bool:
    """Self-check on an initialised process group: every rank computes the full-frame read AND its own row band
    through ``RowBandMemory`` on the same seeded inputs; True iff, on every rank and iteration, the band equals the
    matching rows of the full result bit for bit, masks / strive_time agree and all ranks picked the same frames.
    Used by tests/mgpu_rowband_check.py and by ``bench.py --gpus N`` (key ``rowband_bit_exact``)."""
    b, T, h, w = shape
    c = 128
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    g = torch.Generator().manual_seed(seed)                    # same inputs on every rank
    rn = lambda *s_: torch.randn(*s_, generator=g)
    q, k, v, mf = rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w)
    unc = torch.rand(b * T, 1, h, w, generator=g)
    sim = torch.rand(b, 1, T, T, generator=g) * 2 - 1
    q, k, v, mf, unc, sim = (t.to(device) for t in (q, k, v, mf, unc, sim))
    beta = torch.tensor([0.5], device=device)
    ok = True
    full = PPMMemory(q, k, None, top_k=5)
    band = RowBandMemory(q, k, None, top_k=5, group=group)
    s_full, s_band = torch.ones_like(sim), torch.ones_like(sim)
    r0, r1 = band.r0, band.r1
    for it in range(iterations):                               # strive_time evolves: picks change
        vv = v + 0.1 * it
        want, mask_f = full.read(vv, mf, sim, s_full, unc, beta)
        got, mask_b = band.read(vv[:, :, :, r0:r1].contiguous(), mf[:, :, :, r0:r1].contiguous(), sim, s_band,
                                unc[:, :, r0:r1].contiguous(), beta)
        ok &= bool(torch.equal(got, want[:, :, r0:r1])) and bool(torch.equal(mask_f, mask_b))
        ok &= bool(torch.equal(s_full, s_band))
        if world > 1:
            picks = [torch.empty_like(band.mem.picked) for _ in range(world)]
            dist.all_gather(picks, band.mem.picked, group=group)
            ok &= all(bool(torch.equal(p_, picks[0])) for p_ in picks)
    flag = torch.tensor([1 if ok else 0], device=device)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return int(flag) == 1
