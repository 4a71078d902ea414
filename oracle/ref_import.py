"""Import the UNMODIFIED reference (cocowy1/PPMStereo) (TEST INFRASTRUCTURE ONLY).

Used (a) by tests/golden/make_golden.py in the build container to generate golden vectors from
/root/reference on CPU, and (b) on the GPU box -- where /root/reference does not exist -- from the byte-identical
staged copy under oracle/_ref/ (oracle/stage_ref.py; git-ignored, shipped with the snapshot), by
tests/test_install_gpu.py and bench.py's reference / full-model arms.  Root resolution: $PPM_REFERENCE_ROOT, else
/root/reference when present, else oracle/_ref.  The modules are imported as they lie, with stand-ins for
third-party packages that are absent:

  timm        only ``trunc_normal_`` / ``DropPath`` (models/core/convnext.py:6) and a
              bare import (models/core/extractor.py:14)
  unfoldNd    imported (models/core/ppmstereo.py:15) but unused unless use_convex_3d
  flash_attn  CUDA-only; on a machine WITHOUT a GPU replaced by a CPU function with FlashAttention-2 rounding
              (bf16 operands, fp32 softmax, P rounded to bf16, bf16 output).  With a GPU the installed
              flash_attn is left alone: the reference then runs exactly its own stack.
  torch.load  of the hard-coded ConvNeXt checkpoint (models/core/convnext.py:221-222)
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))


def _resolve_root() -> str:
    env = os.environ.get("PPM_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/models/core"):
        return "/root/reference"
    return os.path.join(_HERE, "_ref")


REFERENCE_ROOT = _resolve_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "core", "ppmstereo.py"))


def _real_flash_attn_usable() -> bool:
    if not torch.cuda.is_available():
        return False
    try:
        import flash_attn  # noqa: F401
        return hasattr(flash_attn, "flash_attn_func")
    except Exception:
        return False


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
    if not _real_flash_attn_usable():
        fa = types.ModuleType("flash_attn")
        fa.flash_attn_func = _flash_attn_cpu
        fa.flash_attn_qkvpacked_func = None
        sys.modules["flash_attn"] = fa


def load_reference():
    """Return the reference's ``models.core`` modules as a namespace."""
    if not available():
        raise RuntimeError(f"reference tree not present at {REFERENCE_ROOT} (stage it with oracle/stage_ref.py)")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    ns = types.SimpleNamespace()
    ns.corr = importlib.import_module("models.core.corr")
    return ns


def build_reference_model(seed: int = 0, beta: float = 0.5, num_frames: int = 5, mixed_precision: bool = False):
    """PPMStereo with the wrapper's arguments (models/ppm_stereo_model.py:27-33; the wrapper sets
    mixed_precision=True and moves the model to "cuda" -- pass mixed_precision=True and call .cuda() for that), seeded
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
    if not _real_flash_attn_usable():
        ppm.flash_attn_func = _flash_attn_cpu
    torch.manual_seed(seed)
    torch.load = fake_load
    try:
        model = ppm.PPMStereo(mixed_precision=mixed_precision, num_frames=num_frames,
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


class MemoryReadRecorder:
    """Records, from an UNMODIFIED reference model while it runs, the inputs and outputs of every PPM memory
    read (models/core/ppmstereo.py:501-555) without touching the reference: forward hooks on ``to_qk`` and the
    update blocks, wrappers around ``get_motion_and_value`` / ``get_uncertainty`` / ``compute_qk_similarity``
    and a recording ``torch.argsort``.  Use as a context manager around ``model(...)``; afterwards
    ``reads(level)`` yields one dict per iteration with the tensors in the layouts PPMMemory takes, on the
    device and in the dtypes the reference produced them (bf16 under autocast on a GPU)."""

    LEVELS = ("l16", "l08", "l04")

    def __init__(self, model, T: int):
        self.model, self.T = model, T
        self.rec = {"qk": {}, "mv": {}, "unc": {}, "upd": {}, "sim": [], "argsort": []}
        self._undo = []

    def __enter__(self):
        m, rec, T = self.model, self.rec, self.T
        blocks = {"l16": m.update_block16, "l08": m.update_block08, "l04": m.update_block04}
        atts = {"l16": m.att[0], "l08": m.att[1], "l04": m.att[2]}
        for lvl, att in atts.items():
            h = att.to_qk.register_forward_hook(
                lambda mod, i, o, lvl=lvl: rec["qk"].setdefault(lvl, []).append(o.detach().clone()))
            self._undo.append(h.remove)
        for lvl, blk in blocks.items():
            gmv, gun = blk.get_motion_and_value, blk.get_uncertainty

            def get_motion_and_value(*a, _f=gmv, _l=lvl, **k):
                r = _f(*a, **k)
                rec["mv"].setdefault(_l, []).append((r[0].detach().clone(), r[2].detach().clone()))
                return r

            def get_uncertainty(*a, _f=gun, _l=lvl, **k):
                r = _f(*a, **k)
                rec["unc"].setdefault(_l, []).append(r.detach().clone())
                return r

            blk.get_motion_and_value, blk.get_uncertainty = get_motion_and_value, get_uncertainty
            self._undo.append(lambda blk=blk: (blk.__dict__.pop("get_motion_and_value", None),
                                               blk.__dict__.pop("get_uncertainty", None)))
            h = blk.register_forward_pre_hook(
                lambda mod, args, lvl=lvl: rec["upd"].setdefault(lvl, []).append(args[3].detach().clone()))
            self._undo.append(h.remove)
        cqs = m.compute_qk_similarity

        def compute_qk_similarity(*a, **k):
            r = cqs(*a, **k)
            rec["sim"].append(r.detach().clone())
            return r

        m.compute_qk_similarity = compute_qk_similarity
        self._undo.append(lambda: m.__dict__.pop("compute_qk_similarity", None))
        real_argsort = torch.argsort

        def argsort(x, *a, **k):
            r = real_argsort(x, *a, **k)
            if x.dim() == 4 and x.shape[1] == 1 and x.shape[2] == x.shape[3] == T:
                rec["argsort"].append((x.detach().clone(), r.clone()))
            return r

        torch.argsort = argsort
        self._undo.append(lambda: setattr(torch, "argsort", real_argsort))
        return self

    def __exit__(self, *exc):
        for u in reversed(self._undo):
            u()
        self._undo.clear()
        return False

    def reads(self, iters: int, top_k: int = 5):
        """Yield (level, iteration, dict) for every memory read of one ``forward(iters=iters)`` call."""
        T, rec = self.T, self.rec
        n_it = {"l16": iters // 2, "l08": iters // 2, "l04": iters}
        a_off = 0
        for li, lvl in enumerate(self.LEVELS):
            qk = rec["qk"][lvl][0]
            _, _, h, w = qk.shape
            b = qk.shape[0] // T
            q = qk[:, :128].reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4)
            k = qk[:, 128:].reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4)
            sim = rec["sim"][li]
            strive = torch.ones_like(sim)
            for it in range(n_it[lvl]):
                score, idx = rec["argsort"][a_off + it]
                mask = torch.zeros_like(score, dtype=torch.bool).scatter_(3, idx[..., :top_k], True)
                mf, val = rec["mv"][lvl][it]
                yield lvl, it, {
                    "query": q, "key": k, "sim": sim, "strive_before": strive.clone(),
                    "value": val.reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4),
                    "motion_features": mf.reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4),
                    "uncertainty": rec["unc"][lvl][it], "frame_score": score, "mask": mask,
                    "motion_features_global": rec["upd"][lvl][it],
                }
                strive = strive.clone()
                strive[mask] += 1
            a_off += n_it[lvl]
