"""Whole-model measurement harness (TEST / BENCH INFRASTRUCTURE ONLY -- imports the reference through ref_import).

Runs ``PPMStereo.forward_batch_test`` (what ``PPMStereoModel.forward`` calls, models/ppm_stereo_model.py:47-50) on a
synthetic clip on the GPU, bf16 autocast, (a) with the reference's own stack (einsum / grid_sample / flash_attn),
(b) after ``ppmstereo_b200.install()``, and splits the device time of either into phases with CUDA events recorded
by forward hooks (modules) and thin wrappers (functions).  Used by ``bench.py --full-model`` and
``scripts/full_model_probe.py``; nothing in ``ppmstereo_b200`` imports it.

Phases (non-overlapping; "other" = whole-clip device time minus their sum: interpolations, cats, the reference's
inline memory-block glue, copies on the compute stream):
  encoders        fnet (BasicEncoder) + cnet (ConvNeXt)                       ppmstereo.py:618-632
  sst             self / cross / temporal attention blocks at 1/16           ppmstereo.py:322-395
  to_qk           Attention_qk.to_qk, once per level                         ppmtereo_update.py:118-133
  corr_build      CorrBlock1D.__init__                                       corr.py:56-72
  corr_lookup     CorrBlock1D.__call__                                       corr.py:74-94
  motion_encoder  BasicMotionEncoder_v2                                      ppmtereo_update.py:445-482
  to_v            Aggregate.to_v                                             ppmtereo_update.py:646
  uncertainty     the uncertainty head (installed: K8 = to_v + the head)     ppmtereo_update.py:889-893
  memory_read     installed: PPMMemory.read (+ per-level prep); reference: flash_attn_func calls only
  upd_attn        TimeAttnBlock + SpaceAttnBlock of update_block16           ppmtereo_update.py:593-631
  gru3d           SKSepConvGRU3D                                             ppmtereo_update.py:254-312
  flow_head       FlowHead3D                                                 ppmtereo_update.py:670-678
  mask_head       mask_2d                                                    ppmtereo_update.py:~905
  update_block    installed arm only: SequenceUpdateBlock3D.forward as a whole (= gru3d + flow_head + mask_head + upd_attn)
  upsample        convex_upsample                                            ppmstereo.py:185-197
"""
from __future__ import annotations

import time
from collections import defaultdict
from typing import Dict, Optional

import torch

from . import ref_import


class PhaseTimer:
    def __init__(self):
        self.pairs = defaultdict(list)
        self._undo = []

    def _rec(self, name):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.pairs[name].append((s, e))
        return s, e

    def module(self, mod, name):
        stack = []

        def pre(m, args):
            s, e = self._rec(name)
            s.record()
            stack.append(e)

        def post(m, args, out):
            stack.pop().record()

        h1, h2 = mod.register_forward_pre_hook(pre), mod.register_forward_hook(post)
        self._undo += [h1.remove, h2.remove]

    def function(self, owner, attr, name):
        fn = getattr(owner, attr)
        timer = self

        def wrapped(*a, **k):
            s, e = timer._rec(name)
            s.record()
            try:
                return fn(*a, **k)
            finally:
                e.record()

        wrapped.__wrapped_by_phase_timer__ = True
        setattr(owner, attr, wrapped)
        self._undo.append(lambda: setattr(owner, attr, fn))

    def close(self):
        for u in reversed(self._undo):
            u()
        self._undo.clear()

    def totals_ms(self) -> Dict[str, float]:
        torch.cuda.synchronize()
        return {k: sum(s.elapsed_time(e) for s, e in v) for k, v in self.pairs.items()}

    def counts(self) -> Dict[str, int]:
        return {k: len(v) for k, v in self.pairs.items()}


def _hook_model(tm: PhaseTimer, model):
    tm.module(model.fnet, "encoders")
    if getattr(model, "use_cnet", False):
        tm.module(model.cnet, "encoders")
    for name in ("self_attn_blocks", "cross_attn_blocks", "time_attn_blocks"):
        for blk in getattr(model, name, []):
            tm.module(blk, "sst")
    for att in model.att:
        tm.module(att.to_qk, "to_qk")
    for ub in (model.update_block16, model.update_block08, model.update_block04):
        tm.module(ub.encoder, "motion_encoder")
        tm.module(ub.aggregator.to_v, "to_v")
        tm.module(ub.uncertainty, "uncertainty")
        if hasattr(type(ub), "_ppm_reference_forward"):
            # installed: SequenceUpdateBlock3D.forward is one call (K5c) on the levels without update attention, so its
            # sub-modules never run there -- time the block as a whole (= gru3d + flow_head + mask_head + upd_attn)
            tm.module(ub, "update_block")
            continue
        tm.module(ub.gru, "gru3d")
        tm.module(ub.flow_head, "flow_head")
        tm.module(ub.mask_2d, "mask_head")
        for nm in ("time_attn", "space_attn"):
            if hasattr(ub, nm):
                tm.module(getattr(ub, nm), "upd_attn")


def run_clip(model, ppm, video, kernel_size: int, iters: int, installed: bool, split: bool = True,
             repeats: int = 1) -> Dict:
    """One (or ``repeats``) forward_batch_test over ``video`` (T,2,3,H,W host tensor).  Returns frames/s by device
    events and by wall clock, the phase split in ms per clip, and the result (host)."""
    import importlib
    import ppmstereo_b200 as P
    inst = importlib.import_module("ppmstereo_b200.install")       # (the package attribute `install` is the function)
    mem_mod = importlib.import_module("ppmstereo_b200.memory")
    if installed:
        P.install(ppm)                                   # defaults: K5 on, eager iterations
    else:
        P.uninstall(ppm)
    tm = PhaseTimer()
    try:
        if split:
            _hook_model(tm, model)
            if installed:
                tm.function(P.CorrBlock1D, "__init__", "corr_build")
                tm.function(P.CorrBlock1D, "__call__", "corr_lookup")
                tm.function(mem_mod.PPMMemory, "__init__", "memory_read")
                tm.function(mem_mod.PPMMemory, "similarity", "memory_read")
                tm.function(mem_mod.PPMMemory, "read", "memory_read")
                tm.function(inst, "convex_upsample", "upsample")
                motion_mod = importlib.import_module("ppmstereo_b200.motion")
                tm.function(motion_mod.ValueUncertainty, "__call__", "uncertainty")      # K8: to_v + uncertainty head
            else:
                tm.function(ppm.CorrBlock1D, "__init__", "corr_build")
                tm.function(ppm.CorrBlock1D, "__call__", "corr_lookup")
                tm.function(ppm, "flash_attn_func", "memory_read")
                tm.function(ppm.PPMStereo, "convex_upsample", "upsample")
        batch = {"stereo_video": video}
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        with torch.no_grad():
            for _ in range(repeats):
                out = model.forward_batch_test(batch, kernel_size=kernel_size, iters=iters)
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / repeats
        dev_ms = e0.elapsed_time(e1) / repeats
        phases = {k: v / repeats for k, v in tm.totals_ms().items()} if split else {}
        counts = tm.counts() if split else {}
    finally:
        tm.close()
        P.uninstall(ppm)
    if split:
        phases["other"] = dev_ms - sum(phases.values())
    n = len(video)
    return {"frames_per_s": n / (dev_ms * 1e-3), "frames_per_s_wall": n / wall, "ms_per_clip": dev_ms,
            "wall_s_per_clip": wall, "phases_ms": {k: round(v, 2) for k, v in sorted(phases.items(), key=lambda kv: -kv[1])},
            "phase_calls": counts, "disparity": out["disparity"], "uncertainties": out["uncertainties"]}


def build(seed: int = 0, beta: float = 0.5, num_frames: int = 5):
    model, ppm = ref_import.build_reference_model(seed=seed, beta=beta, num_frames=num_frames, mixed_precision=True)
    return model.cuda().eval(), ppm


def synthetic_clip(frames: int, H: int, W: int, seed: int = 3, pinned: bool = True) -> torch.Tensor:
    v = torch.rand(frames, 2, 3, H, W, generator=torch.Generator().manual_seed(seed)) * 255.0
    return v.pin_memory() if pinned and torch.cuda.is_available() else v


def measure(frames: int = 40, H: int = 720, W: int = 1280, kernel_size: int = 20, iters: int = 20,
            reference: bool = True, warmup: int = 1, split: bool = True, model_and_ppm=None,
            reference_warmup: int = 0) -> Dict:
    """BASELINE configs[3]: rand(40,2,3,720,1280)*255, ks 20, iters 20, bf16 autocast, one GPU."""
    model, ppm = model_and_ppm if model_and_ppm is not None else build()
    video = synthetic_clip(frames, H, W)
    res: Dict[str, Optional[Dict]] = {"config": {"frames": frames, "H": H, "W": W, "kernel_size": kernel_size,
                                                 "iters": iters, "dtype": "bf16 autocast", "weights": "random init seed 0, "
                                                 "beta=0.5, temporal_fc~N(0,0.02)"}}
    arms = [("installed", True)] + ([("reference", False)] if reference else [])
    for name, inst in arms:
        for _ in range(warmup if inst else reference_warmup):
            run_clip(model, ppm, video, kernel_size, iters, installed=inst, split=False)
        plain = run_clip(model, ppm, video, kernel_size, iters, installed=inst, split=False)
        r = run_clip(model, ppm, video, kernel_size, iters, installed=inst, split=split) if split else plain
        r["frames_per_s"], r["ms_per_clip"] = plain["frames_per_s"], plain["ms_per_clip"]    # headline without the hooks
        r["frames_per_s_wall"], r["wall_s_per_clip"] = plain["frames_per_s_wall"], plain["wall_s_per_clip"]
        res[name] = r
    if reference:
        a, b = res["installed"], res["reference"]
        res["disparity_mad_px"] = float((a["disparity"].float() - b["disparity"].float()).abs().mean())
        res["speedup"] = a["frames_per_s"] / b["frames_per_s"]
    for name, _ in arms:
        res[name].pop("disparity"), res[name].pop("uncertainties")
    return res
