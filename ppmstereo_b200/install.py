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
from .gru import flow_head_forward, gru_forward, update_block_forward
from .motion import LookupRequest, encoder_forward, value_and_uncertainty
from .memory import C, PPMMemory, softmax_scale
from .upsample import convex_upsample
from .window import forward_batch_test


def _inference_only(inp, *params):
    """The kernels have no backward: refuse to run where autograd would silently drop gradients."""
    if torch.is_grad_enabled() and (inp.requires_grad or any(p.requires_grad for p in params)):
        raise RuntimeError("ppmstereo_b200 is inference only (no autograd through the sm_100a kernels): call the "
                           "installed model under torch.no_grad(), or ppmstereo_b200.uninstall() for training")


# Replay of the update iteration from a CUDA graph -- built, measured, and left OFF by default.
# Every kernel of the library launches on the caller's stream, allocates nothing and never synchronises, and the
# reference's glue is plain torch modules, so one iteration can be captured per (window, level) and replayed: iteration 0
# runs eagerly (it warms up cuDNN plans, function attributes, weight packs, and takes the `motion_hidden_state is None`
# branch), iteration 1 is captured on static input buffers, iterations 1..n-1 are replays.  Measured on the B200
# (scripts/small_clip_probe.py, BASELINE configs[0] size, 5 x 320 x 512, iters 10): eager 76 ms per forward, graph 91 ms
# (12 frames: 130 vs 157 ms), with spikes to 2x when the allocator recycles the capture pools -- capturing and
# instantiating a ~400-node graph per (window, level) costs more than 4-19 replays save, and at 720p the device is the
# bound anyway (device time = wall time).  Replayed results are bit-identical to the eager loop's (the small deviations
# seen earlier in round 2 were a descriptor-cache bug in ppm_memory_attend, fixed).
GRAPH = {"mode": "never"}           # "never" (default) | "auto" | "always"   (ppmstereo_b200.install(..., graph=...))
GRAPH_MAX_POSITIONS = 400_000
GRAPH_STATS = {"captured": 0, "replayed": 0, "eager": 0}


def _use_graph(iters: int, positions: int, device) -> bool:
    if device.type != "cuda" or GRAPH["mode"] == "never" or iters < 3 or torch.cuda.is_current_stream_capturing():
        return False
    return GRAPH["mode"] == "always" or (iters >= 4 and positions <= GRAPH_MAX_POSITIONS)


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
    beta = update_block.aggregator.beta

    def step(flow, net, motion_hidden_state):
        """One update iteration (ppmstereo.py:482-591)."""
        flow = flow.detach()
        # K6: when the encoder's front end is fused, the lookup runs inside it (the (B,36,h,w) tensor is never written)
        fused = hasattr(type(update_block.encoder), "_ppm_reference_forward") and getattr(corr_fn, "_l0_only", False)
        out_corrs = LookupRequest(corr_fn, flow) if fused else corr_fn(flow)
        motion_features, motion_hidden_state, value, uncertainty_score = value_and_uncertainty(
            update_block, flow, out_corrs, motion_hidden_state, inp, net)               # K7 + K8
        mf5 = motion_features.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
        v5 = value.reshape(b, t, c, h, w).permute(0, 2, 1, 3, 4)
        motion_features_global, _ = memory.read(v5, mf5, sim_score, strive_time, uncertainty_score, beta, scale)
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
        return flow, net, motion_hidden_state, flow_out, flow_up[:, :1], uncertainty_up

    flow, net, motion_hidden_state, flow_out, pred, unc_up = step(flow, net, motion_hidden_state)    # iteration 0
    predictions.append(pred)
    uncertainties.append(unc_up)
    done = 1
    if _use_graph(iters, bt * h * w, inp.device):
        static_in = [flow.clone(), net.clone(), motion_hidden_state.clone()]
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            outs = step(*static_in)
        if all(o.shape == i.shape and o.dtype == i.dtype for o, i in zip(outs[:3], static_in)):
            GRAPH_STATS["captured"] += 1
            while done < iters:
                graph.replay()
                GRAPH_STATS["replayed"] += 1
                predictions.append(outs[4].clone())
                uncertainties.append(outs[5].clone())
                done += 1
                if done < iters:
                    for dst, src in zip(static_in, outs[:3]):
                        dst.copy_(src)
            flow, net, motion_hidden_state, flow_out = (o.clone() for o in outs[:4])
        del graph, outs, static_in                        # returns the capture's private pool
    while done < iters:
        flow, net, motion_hidden_state, flow_out, pred, unc_up = step(flow, net, motion_hidden_state)
        GRAPH_STATS["eager"] += 1
        predictions.append(pred)
        uncertainties.append(unc_up)
        done += 1
    return flow_out, net, motion_hidden_state


_ORIGINALS = "_ppmstereo_b200_originals"


def _update_classes(ppmstereo_module):
    """The reference's SKSepConvGRU3D (models/core/ppmtereo_update.py:254), FlowHead3D (:670) and BasicMotionEncoder_v2
    (:445) classes with their replacement forwards, found through the module that defines SequenceUpdateBlock3D (imported by name at
    models/core/ppmstereo.py:17-24)."""
    upd = sys.modules.get(ppmstereo_module.SequenceUpdateBlock3D.__module__)
    pairs = [(getattr(upd, "SKSepConvGRU3D", None), gru_forward), (getattr(upd, "FlowHead3D", None), flow_head_forward),
             (getattr(upd, "BasicMotionEncoder_v2", None), encoder_forward),
             (getattr(upd, "SequenceUpdateBlock3D", None), update_block_forward)]
    return [(c, f) for c, f in pairs if c is not None]


def install(ppmstereo_module, gru: bool = True, graph: str = "never"):
    """Patch the reference module in place; returns it.  Idempotent; ``uninstall`` restores the originals.
    ``gru=False`` leaves the SURVEY 8f kernels off: SKSepConvGRU3D, FlowHead3D and the motion encoder's front end stay
    on the reference's cuDNN path (SURVEY 8a rows only).
    ``graph``: "never" (default: eager iterations, measured faster -- see GRAPH above); "always" replays the update
    iteration of every level from a CUDA graph, "auto" only on levels of <= GRAPH_MAX_POSITIONS pixels x frames."""
    if graph not in ("auto", "always", "never"):
        raise ValueError("graph must be 'auto', 'always' or 'never'")
    GRAPH["mode"] = graph
    for cls, fwd in _update_classes(ppmstereo_module):          # K5 / K5b / K6 (ppmtereo_update.py:291-312, :677, :458-482)
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
