"""Import the UNMODIFIED reference (cocowy1/PPMStereo) on CPU (TEST INFRASTRUCTURE ONLY).

Used by tests/golden/make_golden.py in the build container to generate golden
vectors; /root/reference does not exist on the GPU box, so nothing that runs
there may call ``load_reference``.  No reference source is copied: the modules
are imported from where they lie, with stand-ins for third-party packages that
are absent here:

  timm        only ``trunc_normal_`` / ``DropPath`` (models/core/convnext.py:6) and a
              bare import (models/core/extractor.py:14)
  unfoldNd    imported (models/core/ppmstereo.py:15) but unused unless use_convex_3d
  flash_attn  CUDA-only; replaced by a CPU function with FlashAttention-2 rounding
              (bf16 operands, fp32 softmax, P rounded to bf16, bf16 output)
  torch.load  of the hard-coded ConvNeXt checkpoint (models/core/convnext.py:221-222)
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("PPM_REFERENCE_ROOT", "/root/reference")

FLASH_CALLS = []          # filled by the flash_attn stand-in when recording is on
RECORD = {"on": False}


def _flash_attn_cpu(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, **kw):
    """(b, L, heads, d) bf16 in -> (b, Lq, heads, d) bf16 out; see module docstring."""
    assert not causal and dropout_p == 0.0
    b, lq, nh, d = q.shape
    scale = softmax_scale if softmax_scale is not None else d ** -0.5
    qf, kf, vf = (t.to(torch.bfloat16).float() for t in (q, k, v))
    out = torch.empty(b, lq, nh, d)
    for bb in range(b):
        for hh in range(nh):
            for s in range(0, lq, 4096):
                sc = (qf[bb, s:s + 4096, hh] @ kf[bb, :, hh].t()) * scale
                m = sc.amax(-1, keepdim=True)
                p = torch.exp(sc - m)
                l = p.sum(-1, keepdim=True)
                p = p.to(torch.bfloat16).float()
                out[bb, s:s + 4096, hh] = (p @ vf[bb, :, hh]) / l
    out = out.to(torch.bfloat16)
    if RECORD["on"]:
        FLASH_CALLS.append({"q": q.detach().clone(), "k": k.detach().clone(),
                            "v": v.detach().clone(), "out": out.clone(),
                            "scale": float(scale)})
    return out


def _install_stubs():
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")

        class DropPath(torch.nn.Module):
            def __init__(self, drop_prob=0.0):
                super().__init__()
                self.drop_prob = drop_prob

            def forward(self, x):
                return x

        layers.trunc_normal_ = torch.nn.init.trunc_normal_
        layers.DropPath = DropPath
        timm.models, models.layers = models, layers
        sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers})
    if "unfoldNd" not in sys.modules:
        sys.modules["unfoldNd"] = types.ModuleType("unfoldNd")
    fa = types.ModuleType("flash_attn")
    fa.flash_attn_func = _flash_attn_cpu
    fa.flash_attn_qkvpacked_func = None
    sys.modules["flash_attn"] = fa


def load_reference():
    """Return the reference's ``models.core`` modules as a namespace."""
    if not os.path.isdir(REFERENCE_ROOT):
        raise RuntimeError(f"reference tree not present at {REFERENCE_ROOT}")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    ns = types.SimpleNamespace()
    ns.corr = importlib.import_module("models.core.corr")
    return ns


def build_reference_model(seed: int = 0, beta: float = 0.5, num_frames: int = 5):
    """PPMStereo with the wrapper's arguments (models/ppm_stereo_model.py:27-33), seeded
    random weights, ``aggregator.beta`` set non-zero (it is zero-initialised at
    models/core/ppmtereo_update.py:647 and would hide the memory read) and
    ``temporal_fc`` ~ N(0, 0.02) (zero-initialised at :600-601)."""
    load_reference()
    convnext = importlib.import_module("models.core.convnext")
    real_load = torch.load

    def fake_load(path, *a, **k):
        if "convnext" in str(path):
            torch.manual_seed(1234)
            return {"model": convnext.convnextv2_tiny().state_dict()}
        return real_load(path, *a, **k)

    ppm = importlib.import_module("models.core.ppmstereo")
    ppm.flash_attn_func = _flash_attn_cpu
    torch.manual_seed(seed)
    torch.load = fake_load
    try:
        model = ppm.PPMStereo(mixed_precision=False, num_frames=num_frames,
                              attention_type="self_stereo_temporal_update_time_update_space",
                              use_3d_update_block=True, different_update_blocks=True)
    finally:
        torch.load = real_load
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith("aggregator.beta"):
                p.fill_(beta)
            if "temporal_fc" in name:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    model.eval()
    return model, ppm
