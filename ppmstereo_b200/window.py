"""Temporal-window scheduler behind ``PPMStereo.forward_batch_test``.

The reference (models/core/ppmstereo.py:238-320) walks a clip in windows of ``kernel_size`` frames with stride
``kernel_size // 2`` and, per window, uploads the frames synchronously (``.cuda()``), runs ``forward`` -- which
encodes every frame of the window again, although half of them were encoded for the previous window -- and
downloads the result synchronously (``.cpu()``).  ``forward_batch_test`` below keeps the reference's windows,
padding, stitching rule and return value and changes only the schedule:

  * ``FrameFeatureCache`` wraps the two per-frame encoders (``fnet``: BasicEncoder with InstanceNorm,
    ``cnet``: ConvNeXt with LayerNorm -- no statistic crosses frames), so a frame is encoded once per clip
    (40-frame clip: 40 instead of 70 frame encodings);
  * the next window's frames are uploaded from pinned memory on a copy stream while this window computes, and
    results go back through pinned buffers on a second stream; the host blocks once, at the end.

``windows`` / ``stitch_slices`` restate the reference's loop arithmetic (:277-307) and are shared with
``shard.clip_windows``.
"""
from __future__ import annotations

from collections import defaultdict
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn


def windows(num_frames: int, kernel_size: int) -> List[Tuple[int, int]]:
    """[start, stop) of every window the reference runs (ppmstereo.py:277-278); one window when the clip is
    shorter than ``kernel_size`` (:252)."""
    if kernel_size > num_frames:
        return [(0, num_frames)]
    stride = kernel_size // 2
    return [(i, min(i + kernel_size, num_frames)) for i in range(0, num_frames, stride)]


def stitch_slices(num_frames: int, kernel_size: int) -> List[Optional[slice]]:
    """Which frames of each window's output the reference keeps (ppmstereo.py:296-307); None = dropped."""
    if kernel_size > num_frames:
        return [slice(None)]
    stride = kernel_size // 2
    keep: List[Optional[slice]] = []
    kept_any = False
    for start, stop in windows(num_frames, kernel_size):
        n = stop - start
        if kept_any and n >= stride:
            keep.append(slice(stride // 2, None) if n < kernel_size else slice(stride // 2, -stride // 2))
        elif not kept_any:
            keep.append(slice(None, -stride // 2))
            kept_any = True
        else:
            keep.append(None)
    return keep


class FrameFeatureCache(nn.Module):
    """Per-frame memo around an encoder whose output rows depend only on the matching input rows.

    ``begin(frame_ids)`` names the clip frames the next call will be given (row r of every input tensor is
    frame ``frame_ids[r]``); ``forward`` runs the wrapped module on the rows not seen before and assembles the
    result in row order.  Inputs / outputs may be a tensor or a list / tuple of tensors (``fnet([l, r])`` ->
    ``(fmap_l, fmap_r)``; ``cnet(l)`` -> three scales).  ``keep(ids)`` drops everything else.
    """

    def __init__(self, inner: nn.Module):
        super().__init__()
        self.inner = inner
        self._ids: Optional[List[int]] = None
        self._store: Dict[int, Tuple[torch.Tensor, ...]] = {}
        self._single = False
        self.encoded = 0                                   # frames actually run through the encoder

    def begin(self, frame_ids: Sequence[int]) -> None:
        self._ids = [int(i) for i in frame_ids]

    def keep(self, frame_ids: Sequence[int]) -> None:
        wanted = {int(i) for i in frame_ids}
        for k in [k for k in self._store if k not in wanted]:
            del self._store[k]

    def clear(self) -> None:
        self._store.clear()
        self._ids = None

    def forward(self, x):
        if self._ids is None:
            return self.inner(x)
        is_list = isinstance(x, (tuple, list))
        rows = (x[0] if is_list else x).shape[0]
        if rows != len(self._ids):
            raise RuntimeError(f"FrameFeatureCache: {rows} input rows for {len(self._ids)} announced frames")
        missing = [r for r, f in enumerate(self._ids) if f not in self._store]
        if missing:
            if missing == list(range(missing[0], missing[-1] + 1)):
                # the usual case (the new frames are the tail of the window): a view, no index upload
                pick = lambda t: t[missing[0]:missing[-1] + 1]
            else:
                dev = (x[0] if is_list else x).device
                idx = torch.tensor(missing, dtype=torch.long)
                if dev.type == "cuda":                      # pinned + non_blocking: no host sync on the schedule
                    idx = idx.pin_memory().to(dev, non_blocking=True)
                pick = lambda t: t.index_select(0, idx)
            sub = [pick(t) for t in x] if is_list else pick(x)
            out = self.inner(sub)
            self._single = torch.is_tensor(out)
            outs = (out,) if self._single else tuple(out)
            for n, r in enumerate(missing):
                self._store[self._ids[r]] = tuple(o[n:n + 1] for o in outs)
            self.encoded += len(missing)
        parts = [self._store[f] for f in self._ids]
        outs = tuple(torch.cat([p[k] for p in parts], dim=0) for k in range(len(parts[0])))
        return outs[0] if self._single else outs


def _pad_to(t: torch.Tensor, divis_by: int = 32) -> Tuple[torch.Tensor, Tuple[int, int, int, int]]:
    """InputPadder(mode 'sintel') of the reference (models/core/utils/utils.py:22-38): replicate padding,
    split evenly between the two sides."""
    ht, wd = t.shape[-2:]
    pad_ht = (((ht // divis_by) + 1) * divis_by - ht) % divis_by
    pad_wd = (((wd // divis_by) + 1) * divis_by - wd) % divis_by
    pad = (pad_wd // 2, pad_wd - pad_wd // 2, pad_ht // 2, pad_ht - pad_ht // 2)
    if not any(pad):
        return t, pad
    return torch.nn.functional.pad(t, pad, mode="replicate"), pad


def _unpad(t: torch.Tensor, pad: Tuple[int, int, int, int]) -> torch.Tensor:
    ht, wd = t.shape[-2:]
    return t[..., pad[2]:ht - pad[3], pad[0]:wd - pad[1]]


def forward_batch_test(self, batch_dict: Dict, kernel_size: int = 20, iters: int = 20, device=None):
    """Drop-in for ``PPMStereo.forward_batch_test`` (models/core/ppmstereo.py:238-320): same arguments and
    return value (``{"disparity": (T,1,H,W), "uncertainties": (T,1,H,W)}`` on the host).  ``device`` defaults to
    the model's device (the reference hard-codes ``.cuda()``)."""
    video = batch_dict["stereo_video"]
    num_ims = len(video)
    dev = torch.device(device) if device is not None else next(self.parameters()).device
    on_gpu = dev.type == "cuda"
    wins = windows(num_ims, kernel_size)
    keep = stitch_slices(num_ims, kernel_size)

    caches = []
    for name in ("fnet", "cnet"):
        enc = getattr(self, name, None)
        if isinstance(enc, nn.Module) and not isinstance(enc, FrameFeatureCache):
            setattr(self, name, FrameFeatureCache(enc))
        if isinstance(getattr(self, name, None), FrameFeatureCache):
            caches.append((name, getattr(self, name)))

    copy_in = torch.cuda.Stream(dev) if on_gpu else None
    copy_out = torch.cuda.Stream(dev) if on_gpu else None
    resident: Dict[int, torch.Tensor] = {}                 # frame id -> (2, 3, H, W) on the device

    def upload(frames: Sequence[int]):
        """Start the copies of the frames not resident yet; returns the event to wait for."""
        todo = [f for f in frames if f not in resident]
        if not on_gpu:
            for f in todo:
                resident[f] = video[f].to(dev)
            return None
        with torch.cuda.stream(copy_in):
            for f in todo:
                src = video[f]
                if not src.is_cuda and not src.is_pinned():
                    src = src.pin_memory()
                resident[f] = src.to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_in)
        return ev

    results: List[Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]] = []
    done = []
    try:
        ready = upload(range(*wins[0]))
        for n, (start, stop) in enumerate(wins):
            ids = list(range(start, stop))
            if ready is not None:
                cur = torch.cuda.current_stream(dev)
                cur.wait_event(ready)
                # the frame buffers were allocated on copy_in: tell the caching allocator that the compute stream
                # reads them, or a later upload could reuse a freed block before this stack has executed
                for f in ids:
                    resident[f].record_stream(cur)
            clip = torch.stack([resident[f] for f in ids])              # (t, 2, 3, H, W)
            if n + 1 < len(wins):
                ready = upload(range(*wins[n + 1]))                     # overlaps this window's kernels
            left, pad = _pad_to(clip[:, 0])
            right, _ = _pad_to(clip[:, 1])
            for _, c in caches:
                c.begin(ids)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bool(self.mixed_precision) and on_gpu):
                disp, unc = self.forward(left[None], right[None], iters=iters, test_mode=True)
            disp, unc = _unpad(disp[0], pad)[:, None], _unpad(unc[0], pad)[:, None]
            nxt = set(range(*wins[n + 1])) if n + 1 < len(wins) else set()
            for _, c in caches:
                c.keep(nxt)
            for f in [f for f in resident if f not in nxt]:
                del resident[f]
            if keep[n] is None:
                results.append((None, None))
                continue
            disp, unc = disp[keep[n]], unc[keep[n]]
            if on_gpu:
                finished = torch.cuda.Event()
                finished.record(torch.cuda.current_stream(dev))
                with torch.cuda.stream(copy_out):
                    copy_out.wait_event(finished)
                    hd = torch.empty(disp.shape, dtype=disp.dtype, pin_memory=True)
                    hu = torch.empty(unc.shape, dtype=unc.dtype, pin_memory=True)
                    disp.record_stream(copy_out)
                    unc.record_stream(copy_out)
                    hd.copy_(disp, non_blocking=True)
                    hu.copy_(unc, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(copy_out)
                done.append(ev)
                results.append((hd, hu))
            else:
                results.append((disp.cpu(), unc.cpu()))
        for ev in done:
            ev.synchronize()
    finally:
        for name, c in caches:
            c.clear()
            setattr(self, name, c.inner)                                # leave the module tree as it was

    predictions = defaultdict(list)
    disps = [d for d, _ in results if d is not None]
    uncs = [u for _, u in results if u is not None]
    predictions["disparity"] = torch.cat(disps).squeeze(1).abs()[:, :1]
    predictions["uncertainties"] = torch.cat(uncs).squeeze(1).abs()[:, :1]
    return predictions
