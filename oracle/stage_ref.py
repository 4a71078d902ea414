"""Stage the UNMODIFIED reference model sources next to the oracle (TEST INFRASTRUCTURE ONLY).

    python oracle/stage_ref.py            # in the build container, where /root/reference exists

Copies the files ``PPMStereo`` imports (models/core/*.py + utils) byte for byte into ``oracle/_ref/models/``.
``oracle/_ref/`` is git-ignored (no reference source ever enters the history) but NOT gpurun-ignored, so the
staged copy travels to the GPU box like a built ``.so`` does.  There it serves three purposes, all of them
checker / baseline roles, none on the product path:
  * tests/test_install_gpu.py runs the reference model as the reference runs it on a GPU (its own einsum /
    grid_sample / argsort / real flash_attn, bf16 autocast) next to the same model after
    ``ppmstereo_b200.install()`` and compares disparities, masks and per-iteration memory reads;
  * bench.py --impl reference / cpu_baseline time the reference's own CorrBlock1D and memory block on the
    host cores (kind "reference");
  * bench.py --full-model times the whole model, reference vs installed.
``__graft_entry__.build()`` calls ``stage()`` whenever /root/reference is present.  A manifest with the sha256 of
every staged file is written beside them so a stale or edited copy is detectable.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = "/root/reference"
DST_ROOT = os.path.join(HERE, "_ref")
FILES = [
    "models/core/ppmstereo.py",
    "models/core/ppmtereo_update.py",
    "models/core/extractor.py",
    "models/core/corr.py",
    "models/core/convnext.py",
    "models/core/attention.py",
    "models/core/utils/utils.py",
]


def _sha(path: str) -> str:
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def stage(src_root: str = SRC_ROOT, dst_root: str = DST_ROOT) -> str:
    if not os.path.isdir(src_root):
        raise RuntimeError(f"{src_root} not present: staging happens in the build container only")
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(src_root, rel), os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not (os.path.exists(dst) and _sha(dst) == _sha(src)):
            shutil.copyfile(src, dst)
        manifest[rel] = _sha(dst)
    with open(os.path.join(dst_root, "MANIFEST.json"), "w") as f:
        json.dump({"source": src_root, "sha256": manifest}, f, indent=1)
    return dst_root


def staged_ok(dst_root: str = DST_ROOT) -> bool:
    """True when every file of the manifest is present and unmodified."""
    mf = os.path.join(dst_root, "MANIFEST.json")
    if not os.path.exists(mf):
        return False
    want = json.load(open(mf))["sha256"]
    return all(os.path.exists(os.path.join(dst_root, r)) and _sha(os.path.join(dst_root, r)) == h
               for r, h in want.items()) and set(want) == set(FILES)


if __name__ == "__main__":
    print(stage())
