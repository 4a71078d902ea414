"""Rewire a loaded reference PPMStereo onto the B200 kernels without touching its signatures.

``install(module)`` takes the imported ``models.core.ppmstereo`` module of the reference and
  * rebinds the name ``CorrBlock1D`` used at models/core/ppmstereo.py:707, :743, :776, and
  * replaces ``PPMStereo.forward_update_block`` (models/core/ppmstereo.py:426-594) by the
    function below: same arguments, same return value, same side effects on ``predictions`` /
    ``uncertainties``; the block :501-555 becomes one PPMMemory.read call, and
  * replaces ``PPMStereo.forward_batch_test`` (:238-320) by ``window.forward_batch_test``: same windows,
    stitching and return value, each frame encoded once per clip, copies overlapped with the kernels.
``PPMStereoModel.forward`` / ``PPMStereo.forward`` and all state-dict keys are unchanged.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

import sys

from .corr import CorrBlock1D
from .gru import flow_head_forward, gru_forward
from .memory import C, PPMMemory, softmax_scale
from .upsample import convex_upsample
from .window import forward_batch_test


def _inference_only(inp, *params):
    """The kernels have no backward: refuse to run where autograd would silently drop gradients."""
    if torch.is_grad_enabled() and (inp.requires_grad or any(p.requires_grad for p in params)):
        raise RuntimeError("ppmstereo_b200 is inference only (no autograd through the sm_100a kernels): call the "
                           "installed model under torch.no_grad(), or ppmstereo_b200.uninstall() for training")


def forward_update_block(self, image1, update_block, corr_fn, flow, net, inp, motion_hidden_state,
                         attn_block, predictions, uncertainties, iters, interp_scale, t):
    _inference_only(inp, attn_block.to_qk.weight, update_block.aggregator.beta)
    bt, c, h, w = inp.shape
    b = bt // t
    if c != C:
        raise RuntimeError(f"context dim {c} != {C}")
    # once per level: q/k projections, QAM similarity, pick counter (ppmstereo.py:447-476)
    qk = attn_block.to_qk(inp)
    query = qk[:, :c].reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
    key = qk[:, c:].reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
    memory = PPMMemory(query, key, None, top_k=5)
    sim_score = memory.similarity()                       # compute_qk_similarity (ppmstereo.py:397-423)
    strive_time = torch.ones_like(sim_score)
    scale = softmax_scale(c)
    autocast = torch.autocast("cuda", dtype=torch.bfloat16, enabled=bool(self.mixed_precision))
    flow_out = flow
    for _ in range(iters):
        flow = flow.detach()
        out_corrs = corr_fn(flow)
        motion_features, motion_hidden_state, value = update_block.get_motion_and_value(
            flow, out_corrs, motion_hidden_state, inp)
        uncertainty_score = update_block.get_uncertainty(torch.cat([net, value], dim=1))
        mf5 = motion_features.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
        v5 = value.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
        motion_features_global, _ = memory.read(v5, mf5, sim_score, strive_time, uncertainty_score,
                                                update_block.aggregator.beta, scale)
        with autocast:
            net, up_mask, delta_flow = update_block(net, inp, motion_features, motion_features_global, t=t)
        flow = flow + delta_flow
        if self.use_convex_3d:
            flow_up = flow_out = self.convex_upsample_3d(flow, up_mask, rate=4, T=t)
        else:
            flow_up = flow_out = convex_upsample(flow, up_mask, rate=4)     # K4 (ppmstereo.py:185-197)
        uncertainty_up = F.interpolate(uncertainty_score, scale_factor=4 * interp_scale, mode="bilinear")
        if interp_scale > 1:
            size = (interp_scale * flow_out.shape[2], interp_scale * flow_out.shape[3])
            flow_up = interp_scale * F.interpolate(flow_out, size=size, mode="bilinear", align_corners=True)
        predictions.append(flow_up[:, :1])
        uncertainties.append(uncertainty_up)
    return flow_out, net, motion_hidden_state


_ORIGINALS = "_ppmstereo_b200_originals"


def _update_classes(ppmstereo_module):
    """The reference's SKSepConvGRU3D (models/core/ppmtereo_update.py:254) and FlowHead3D (:670) classes with their
    replacement forwards, found through the module that defines SequenceUpdateBlock3D (imported by name at
    models/core/ppmstereo.py:17-24)."""
    upd = sys.modules.get(ppmstereo_module.SequenceUpdateBlock3D.__module__)
    pairs = [(getattr(upd, "SKSepConvGRU3D", None), gru_forward), (getattr(upd, "FlowHead3D", None), flow_head_forward)]
    return [(c, f) for c, f in pairs if c is not None]


def install(ppmstereo_module, gru: bool = True):
    """Patch the reference module in place; returns it.  Idempotent; ``uninstall`` restores the originals.
    ``gru=False`` leaves SKSepConvGRU3D and FlowHead3D on the reference's cuDNN path (SURVEY 8a rows only)."""
    for cls, fwd in _update_classes(ppmstereo_module):                      # K5 / K5b (ppmtereo_update.py:291-312, :677)
        if gru and not hasattr(cls, "_ppm_reference_forward"):
            cls._ppm_reference_forward = cls.forward
            cls.forward = fwd
        elif not gru and hasattr(cls, "_ppm_reference_forward"):
            cls.forward = cls._ppm_reference_forward
            del cls._ppm_reference_forward
    if not hasattr(ppmstereo_module, _ORIGINALS):
        cls = ppmstereo_module.PPMStereo
        setattr(ppmstereo_module, _ORIGINALS, {
            "CorrBlock1D": ppmstereo_module.CorrBlock1D,
            "forward_update_block": cls.forward_update_block,
            "forward_batch_test": cls.forward_batch_test,
        })
    ppmstereo_module.CorrBlock1D = CorrBlock1D
    ppmstereo_module.PPMStereo.forward_update_block = forward_update_block
    ppmstereo_module.PPMStereo.forward_batch_test = forward_batch_test      # window scheduler (ppmstereo.py:238-320)
    return ppmstereo_module


def uninstall(ppmstereo_module):
    """Put the reference's own CorrBlock1D / forward_update_block / forward_batch_test back."""
    for cls, _ in _update_classes(ppmstereo_module):
        if hasattr(cls, "_ppm_reference_forward"):
            cls.forward = cls._ppm_reference_forward
            del cls._ppm_reference_forward
    orig = getattr(ppmstereo_module, _ORIGINALS, None)
    if orig is not None:
        ppmstereo_module.CorrBlock1D = orig["CorrBlock1D"]
        ppmstereo_module.PPMStereo.forward_update_block = orig["forward_update_block"]
        ppmstereo_module.PPMStereo.forward_batch_test = orig["forward_batch_test"]
        delattr(ppmstereo_module, _ORIGINALS)
    return ppmstereo_module
