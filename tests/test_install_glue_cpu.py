"""Drop-in glue check (CPU, build container only): the reference PPMStereo with
``ppmstereo_b200.install`` applied must reproduce the reference's own disparity.

The CUDA kernels cannot run here, so the kernel-backed pieces the glue calls (CorrBlock1D,
PPMMemory, convex_upsample) are replaced IN THIS TEST by stand-ins built on the CPU oracle -- the oracle the GPU parity
tests hold the kernels to.  What is under test is the host-side mirror: the replacement
``forward_update_block`` (same signature, side effects and tensor plumbing as models/core/ppmstereo.py:
426-594) and the ``CorrBlock1D`` rebinding.  Bar (north_star): final disparity within 0.01 px mean
absolute difference of the unmodified reference (fixture tests/golden/model_T6_64x256_it4.npz).
Skipped where /root/reference does not exist (the GPU box)."""
import os

import numpy as np
import pytest
import torch

from oracle import ppm_oracle as O
from oracle import ref_import

pytestmark = pytest.mark.skipif(not os.path.isdir(ref_import.REFERENCE_ROOT),
                                reason="reference tree not present (GPU box)")


class OracleCorrBlock1D:
    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels, self.radius = num_levels, radius
        self.levels = O.corr_pyramid(fmap1.float(), fmap2.float(), num_levels)
        self.corr_pyramid = [l.reshape(l.shape[0], 1, 1, l.shape[-1]) for l in self.levels]

    def __call__(self, flow):
        return O.corr_lookup(self.levels, flow, self.radius)


class OraclePPMMemory:
    def __init__(self, query, key, pe=None, top_k=5):
        self.query, self.key, self.top_k = query.float(), key.float(), top_k
        b, c, T, h, w = query.shape
        self.pe = O.temporal_pe(T, c) if pe is None else pe

    def similarity(self):
        return O.qk_similarity(self.query, self.key)

    def read(self, value, motion_features, sim, strive, uncertainty, beta, scale=None):
        r = O.ppm_memory_read(self.query, self.key, self.pe, value, motion_features, sim, strive, uncertainty,
                              float(beta), scale, self.top_k)
        strive.copy_(r["strive"])                                # the CUDA path updates strive_time in place
        return r["motion_features_global"], r["mask"]


def test_installed_model_reproduces_reference_disparity(golden_dir, monkeypatch):
    g = np.load(os.path.join(golden_dir, "model_T6_64x256_it4.npz"))
    T, H, W, iters, seed = (int(x) for x in g["meta"])
    model, ppm = ref_import.build_reference_model(seed=0, beta=0.5)
    original = (ppm.CorrBlock1D, ppm.PPMStereo.forward_update_block)
    import importlib
    inst = importlib.import_module("ppmstereo_b200.install")
    monkeypatch.setattr(inst, "CorrBlock1D", OracleCorrBlock1D)
    monkeypatch.setattr(inst, "PPMMemory", OraclePPMMemory)
    monkeypatch.setattr(inst, "convex_upsample", O.convex_upsample)
    try:
        inst.install(ppm)
        assert ppm.CorrBlock1D is OracleCorrBlock1D
        assert ppm.PPMStereo.forward_update_block is inst.forward_update_block
        gen = torch.Generator().manual_seed(seed)
        video = torch.rand(T, 2, 3, H, W, generator=gen) * 255.0
        with torch.no_grad():
            disp, unc = model(video[None, :, 0], video[None, :, 1], iters=iters, test_mode=True)
    finally:
        ppm.CorrBlock1D, ppm.PPMStereo.forward_update_block = original
    ref_d, ref_u = torch.from_numpy(g["disparity"]), torch.from_numpy(g["uncertainty"])
    assert disp.shape == ref_d.shape and unc.shape == ref_u.shape
    mad = float((disp - ref_d).abs().mean())
    assert mad <= 0.01, f"disparity mean abs diff {mad} px"
    assert float((unc - ref_u).abs().mean()) <= 1e-3
