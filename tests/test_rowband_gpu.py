"""GPU test of the optional row-band split: needs >= 2 GPUs (skipped on a 1-GPU box)."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
def test_row_band_read_equals_full_read_on_two_gpus():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(HERE, "mgpu_rowband_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ROWBAND_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_band_on_one_gpu_matches_full():
    """Single-GPU form of the same property: two bands computed one after the other tile the full result."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import ppmstereo_b200 as P
    g = torch.Generator().manual_seed(9)
    b, c, T, h, w = 1, 128, 6, 19, 33                         # hw = 627
    rn = lambda *s: torch.randn(*s, generator=g).cuda()
    q, k, v, mf = rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w)
    unc = torch.rand(b * T, 1, h, w, generator=g).cuda()
    sim = (torch.rand(b, 1, T, T, generator=g) * 2 - 1).cuda()
    beta = torch.tensor([0.5], device="cuda")
    mem = P.PPMMemory(q, k, None, top_k=5)
    mem.pick(sim, torch.ones_like(sim), unc)
    mem.pack_kv(v)
    want = mem.aggregate(mf, beta)
    got = torch.full_like(want, float("nan"))
    cut = 7 * w + 5                                            # a band boundary inside a query tile
    mem.aggregate(mf, beta, out=got, band=(0, cut))
    assert torch.isnan(got.reshape(b * T, c, -1)[:, :, cut:]).all()      # pixels outside the band untouched
    mem.aggregate(mf, beta, out=got, band=(cut, h * w - cut))
    assert torch.equal(got, want)
