"""Window scheduler (ppmstereo_b200/window.py) against the reference's forward_batch_test semantics
(models/core/ppmstereo.py:238-320).  The arithmetic tests run anywhere; the model tests need the reference
tree (build container only) and run it on CPU -- what is under test is host logic: window cuts, the stitching
rule, the per-frame encoder cache and that the module tree is left untouched."""
import os

import pytest
import torch

from oracle import ref_import
from ppmstereo_b200 import shard, window

needs_reference = pytest.mark.skipif(not os.path.isdir(ref_import.REFERENCE_ROOT),
                                     reason="reference tree not present (GPU box)")


def reference_loop(num_ims, kernel_size):
    """The loop of ppmstereo.py:277-307 on frame indices instead of tensors."""
    if kernel_size > num_ims:
        return list(range(num_ims))
    stride = kernel_size // 2
    preds = []
    for i in range(0, num_ims, stride):
        out = list(range(i, min(i + kernel_size, num_ims)))
        if len(preds) > 0 and len(out) >= stride:
            preds.append(out[stride // 2:] if len(out) < kernel_size else out[stride // 2: -stride // 2])
        elif len(preds) == 0:
            preds.append(out[: -stride // 2])
    return [f for p in preds for f in p]


@pytest.mark.parametrize("num_ims,kernel_size", [(40, 20), (41, 20), (50, 20), (20, 20), (5, 20), (150, 20), (23, 10),
                                                 (6, 4), (7, 6), (31, 10)])
def test_windows_and_stitching_match_the_reference_loop(num_ims, kernel_size):
    wins = window.windows(num_ims, kernel_size)
    keep = window.stitch_slices(num_ims, kernel_size)
    assert len(wins) == len(keep)
    got = []
    for (a, b), k in zip(wins, keep):
        if k is not None:
            got += list(range(a, b))[k]
    assert got == reference_loop(num_ims, kernel_size)
    if kernel_size <= num_ims:
        assert wins == shard.clip_windows(num_ims, kernel_size)


def test_40_frame_clip_keeps_every_frame_once():
    assert reference_loop(40, 20) == list(range(40))
    assert [b - a for a, b in window.windows(40, 20)] == [20, 20, 20, 10]


class CountingEncoder(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.rows = 0

    def forward(self, x):
        if isinstance(x, (list, tuple)):
            self.rows += x[0].shape[0]
            return tuple(t * 2.0 + 1.0 for t in x)
        self.rows += x.shape[0]
        return x * 3.0, x[:, :1] - 1.0, x.mean(dim=(2, 3), keepdim=True)


def test_frame_cache_encodes_each_frame_once():
    g = torch.Generator().manual_seed(0)
    frames = torch.randn(8, 3, 4, 5, generator=g)
    for as_list in (True, False):
        inner = CountingEncoder()
        cache = window.FrameFeatureCache(inner)
        for ids in ([0, 1, 2, 3], [2, 3, 4, 5], [4, 5, 6, 7], [6, 7]):
            cache.begin(ids)
            x = frames[ids]
            got = cache([x, -x]) if as_list else cache(x)
            want = CountingEncoder()([x, -x]) if as_list else CountingEncoder()(x)
            assert len(got) == len(want)
            for a, b in zip(got, want):
                assert torch.equal(a, b)
            cache.keep(range(ids[0] + 2, ids[0] + 6))
        assert inner.rows == 8 and cache.encoded == 8
    with pytest.raises(RuntimeError):
        cache.begin([0, 1])
        cache(frames[:3])


@needs_reference
def test_forward_batch_test_matches_per_window_forward():
    """Ours (cached encoders, staged copies disabled on CPU) against the reference rule applied to plain
    ``forward`` calls per window; T = 6, kernel 4 -> windows [0:4] [2:6] [4:6]."""
    model, ppm = ref_import.build_reference_model(seed=0, beta=0.5)
    T, H, W, ks, iters = 6, 64, 256, 4, 2
    video = torch.rand(T, 2, 3, H, W, generator=torch.Generator().manual_seed(3)) * 255.0
    fnet, cnet = model.fnet, getattr(model, "cnet", None)
    with torch.no_grad():
        got = window.forward_batch_test(model, {"stereo_video": video}, kernel_size=ks, iters=iters, device="cpu")
        assert model.fnet is fnet and getattr(model, "cnet", None) is cnet       # wrappers removed again
        disps, uncs = [], []
        for (a, b), k in zip(window.windows(T, ks), window.stitch_slices(T, ks)):
            d, u = model(video[None, a:b, 0], video[None, a:b, 1], iters=iters, test_mode=True)
            if k is not None:
                disps.append(d[0][:, None][k])
                uncs.append(u[0][:, None][k])
    want_d = torch.cat(disps).squeeze(1).abs()[:, :1]
    want_u = torch.cat(uncs).squeeze(1).abs()[:, :1]
    assert got["disparity"].shape == want_d.shape == (T, 1, H, W)
    assert float((got["disparity"] - want_d).abs().mean()) <= 1e-3
    assert float((got["uncertainties"] - want_u).abs().mean()) <= 1e-4


@needs_reference
def test_reference_encoders_are_per_frame():
    """The premise of the cache: fnet / cnet rows depend only on their own frame."""
    model, _ = ref_import.build_reference_model(seed=0, beta=0.5)
    x = torch.rand(4, 3, 64, 256, generator=torch.Generator().manual_seed(4)) * 2 - 1
    with torch.no_grad():
        full = model.fnet([x, x.flip(0)])
        part = model.fnet([x[1:3], x.flip(0)[1:3]])
        for f, p in zip(full, part):
            assert float((f[1:3] - p).abs().max()) <= 1e-4 * float(f.abs().max())
        if getattr(model, "use_cnet", False):
            full, part = model.cnet(x), model.cnet(x[2:3])
            for f, p in zip(full, part):
                assert float((f[2:3] - p).abs().max()) <= 1e-4 * float(f.abs().max())
