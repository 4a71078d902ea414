"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python tests/golden/make_golden.py

Reads /root/reference through oracle/ref_import.py (stand-ins for timm, unfoldNd,
flash_attn and the ConvNeXt checkpoint; see that file).  The fixtures pin the CPU
oracle (tests/test_oracle_golden.py) and are compared with the CUDA path on the GPU
box (tests/test_*_gpu.py), where /root/reference does not exist.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_import  # noqa: E402


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)


def corr_cases():
    ref = ref_import.load_reference()
    cases = {
        # name: (B, D, H, W, flow range)
        "corr_d256_w82": (2, 256, 3, 82, (-30.0, 6.0)),     # odd widths 82->41->20->10->5
        "corr_d128_w64": (1, 128, 4, 64, (-40.0, 0.0)),
        "corr_d256_w160": (1, 256, 2, 160, (-60.0, 10.0)),   # W2 > 128: two MMA n-tiles
    }
    for i, (name, (B, D, H, W, (lo, hi))) in enumerate(cases.items()):
        g = torch.Generator().manual_seed(100 + i)
        f1 = torch.randn(B, D, H, W, generator=g)
        f2 = torch.randn(B, D, H, W, generator=g)
        flow = torch.zeros(B, 2, H, W)
        flow[:, 0] = torch.rand(B, H, W, generator=g) * (hi - lo) + lo
        blk = ref.corr.CorrBlock1D(f1, f2, num_levels=4, radius=4)
        out = blk(flow)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            fmap1=f1.numpy(), fmap2=f2.numpy(), flow=flow.numpy(),
            **{f"level{l}": p.reshape(p.shape[0], p.shape[-1]).numpy()
               for l, p in enumerate(blk.corr_pyramid)},
            lookup=out.numpy())
        print(name, [tuple(p.shape) for p in blk.corr_pyramid], tuple(out.shape))


def model_cases():
    T, H, W, iters = 6, 64, 256, 4
    model, ppm = ref_import.build_reference_model(seed=0, beta=0.5)
    rec = {"qk": {}, "mv": {}, "unc": {}, "upd": {}, "sim": [], "argsort": []}

    # --- hooks -----------------------------------------------------------
    blocks = {"l16": model.update_block16, "l08": model.update_block08, "l04": model.update_block04}
    atts = {"l16": model.att[0], "l08": model.att[1], "l04": model.att[2]}
    for lvl, att in atts.items():
        att.to_qk.register_forward_hook(
            lambda m, i, o, lvl=lvl: rec["qk"].setdefault(lvl, []).append(o.detach().clone()))
    for lvl, blk in blocks.items():
        def wrap(blk=blk, lvl=lvl):
            gmv, gun = blk.get_motion_and_value, blk.get_uncertainty

            def get_motion_and_value(*a, **k):
                r = gmv(*a, **k)
                rec["mv"].setdefault(lvl, []).append((r[0].detach().clone(), r[2].detach().clone()))
                return r

            def get_uncertainty(*a, **k):
                r = gun(*a, **k)
                rec["unc"].setdefault(lvl, []).append(r.detach().clone())
                return r

            blk.get_motion_and_value, blk.get_uncertainty = get_motion_and_value, get_uncertainty
            blk.register_forward_pre_hook(
                lambda m, args: rec["upd"].setdefault(lvl, []).append(args[3].detach().clone()))
        wrap()
    cqs = model.compute_qk_similarity

    def compute_qk_similarity(*a, **k):
        r = cqs(*a, **k)
        rec["sim"].append(r.detach().clone())
        return r

    model.compute_qk_similarity = compute_qk_similarity
    real_argsort = torch.argsort

    def argsort(x, *a, **k):
        r = real_argsort(x, *a, **k)
        if x.dim() == 4 and x.shape[1] == 1 and x.shape[2] == x.shape[3] == T:
            rec["argsort"].append((x.detach().clone(), r.clone()))
        return r

    g = torch.Generator().manual_seed(7)
    video = torch.rand(T, 2, 3, H, W, generator=g) * 255.0
    ref_import.FLASH_CALLS.clear()
    ref_import.RECORD["on"] = True
    torch.argsort = argsort
    try:
        with torch.no_grad():
            disp, unc = model(video[None, :, 0], video[None, :, 1], iters=iters, test_mode=True)
    finally:
        torch.argsort = real_argsort
        ref_import.RECORD["on"] = False
    np.savez_compressed(os.path.join(HERE, "model_T6_64x256_it4.npz"),
                        disparity=disp.numpy(), uncertainty=unc.numpy(),
                        meta=np.array([T, H, W, iters, 7]))
    print("disparity", tuple(disp.shape), float(disp.abs().mean()))

    # --- assemble per-level PPM-read fixtures ------------------------------
    n_it = {"l16": iters // 2, "l08": iters // 2, "l04": iters}
    order = ["l16", "l08", "l04"]
    a_off = f_off = 0
    for li, lvl in enumerate(order):
        qk = rec["qk"][lvl][0]                       # (T, 256, h, w)
        _, _, h, w = qk.shape
        q = qk[:, :128].reshape(1, T, 128, h, w).permute(0, 2, 1, 3, 4).contiguous()
        k = qk[:, 128:].reshape(1, T, 128, h, w).permute(0, 2, 1, 3, 4).contiguous()
        sim = rec["sim"][li]
        strive = torch.ones_like(sim)
        for it in range(n_it[lvl]):
            score, idx = rec["argsort"][a_off + it]
            mask = torch.zeros_like(score, dtype=torch.bool).scatter_(3, idx[..., :5], True)
            if lvl in ("l16", "l08") and it == 1:
                mf, val = rec["mv"][lvl][it]
                calls = ref_import.FLASH_CALLS[f_off + it * T: f_off + (it + 1) * T]
                np.savez_compressed(
                    os.path.join(HERE, f"ppm_read_{lvl}.npz"),
                    query=q.numpy(), key=k.numpy(),
                    value=val.reshape(1, T, 128, h, w).permute(0, 2, 1, 3, 4).contiguous().numpy(),
                    motion_features=mf.reshape(1, T, 128, h, w).permute(0, 2, 1, 3, 4).contiguous().numpy(),
                    uncertainty=rec["unc"][lvl][it].numpy(),
                    sim=sim.numpy(), strive_before=strive.numpy(),
                    frame_score=score.numpy(), argsort=idx.numpy(), mask=mask.numpy(),
                    hidden_bf16=np.stack([bf16_bits(c["out"][0, :, 0]) for c in calls]),
                    kmod_bf16=np.stack([bf16_bits(c["k"][0, :, 0]) for c in calls]),
                    scale=np.float64(calls[0]["scale"]), beta=np.float32(0.5),
                    motion_features_global=rec["upd"][lvl][it].numpy())
                print(lvl, "hw", h * w, "picked", [torch.nonzero(mask[0, 0, i]).flatten().tolist() for i in range(T)])
            strive = strive.clone()
            strive[mask] += 1
        a_off += n_it[lvl]
        f_off += n_it[lvl] * T


def upsample_cases():
    """PPMStereo.convex_upsample (models/core/ppmstereo.py:185-197) on fp32 and bf16 masks; the method uses no
    module state, so it is called unbound."""
    import importlib
    ref_import.load_reference()
    ppm = importlib.import_module("models.core.ppmstereo")
    g = torch.Generator().manual_seed(21)
    N, H, W = 2, 7, 13
    flow = torch.randn(N, 2, H, W, generator=g) * 6.0
    mask = torch.randn(N, 9 * 16, H, W, generator=g) * 2.0
    out32 = ppm.PPMStereo.convex_upsample(None, flow, mask, 4)
    # A bf16 mask reaches the op only under the reference's torch.cuda.amp.autocast (ppmstereo.py:617), where
    # torch.softmax up-casts its input to fp32 (autocast's fp32 op list).  That context is a no-op on CPU, so the
    # up-cast autocast would perform is applied to the argument here.
    mask16 = mask.bfloat16()
    out16 = ppm.PPMStereo.convex_upsample(None, flow, mask16.float(), 4)
    np.savez_compressed(os.path.join(HERE, "convex_upsample.npz"), flow=flow.numpy(), mask=mask.numpy(),
                        out_fp32_mask=out32.float().numpy(), mask_bf16=bf16_bits(mask16),
                        out_bf16_mask=out16.float().numpy())
    print("convex_upsample", tuple(out32.shape), out16.dtype)


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    if "--upsample-only" in sys.argv:
        upsample_cases()
        sys.exit(0)
    corr_cases()
    model_cases()
    upsample_cases()
