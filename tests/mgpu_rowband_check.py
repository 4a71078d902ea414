"""Multi-GPU check of the optional row-band split (run with torchrun on >= 2 GPUs; launched by
tests/test_rowband_gpu.py).  Every rank computes the full-frame PPM read AND its own row band through
RowBandMemory (value / uncertainty bands all-gathered over NCCL); the band must equal the matching rows
of the full result bit for bit, and the picked frame sets must agree across ranks."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ppmstereo_b200 as P  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    b, c, T, h, w = 1, 128, 7, 21, 40                         # hw = 840: partial tiles, uneven bands
    g = torch.Generator().manual_seed(5)                      # same inputs on every rank
    rn = lambda *s: torch.randn(*s, generator=g)
    q, k, v, mf = rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w), rn(b, c, T, h, w)
    unc = torch.rand(b * T, 1, h, w, generator=g)
    sim = torch.rand(b, 1, T, T, generator=g) * 2 - 1
    q, k, v, mf, unc, sim = (t.to(dev) for t in (q, k, v, mf, unc, sim))
    beta = torch.tensor([0.5], device=dev)
    ok = True
    full = P.PPMMemory(q, k, None, top_k=5)
    band = P.RowBandMemory(q, k, None, top_k=5)
    s_full, s_band = torch.ones_like(sim), torch.ones_like(sim)
    r0, r1 = band.r0, band.r1
    for it in range(3):                                       # strive_time evolves: picks change
        vv = v + 0.1 * it
        want, mask_f = full.read(vv, mf, sim, s_full, unc, beta)
        got, mask_b = band.read(vv[:, :, :, r0:r1].contiguous(), mf[:, :, :, r0:r1].contiguous(), sim, s_band,
                                unc[:, :, r0:r1].contiguous(), beta)
        ok &= bool(torch.equal(got, want[:, :, r0:r1])) and bool(torch.equal(mask_f, mask_b))
        ok &= bool(torch.equal(s_full, s_band))
        picks = [torch.empty_like(band.mem.picked) for _ in range(world)]
        dist.all_gather(picks, band.mem.picked)
        ok &= all(bool(torch.equal(p, picks[0])) for p in picks)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("ROWBAND_OK" if int(flag) == 1 else "ROWBAND_MISMATCH", f"world={world} bands={P.row_bands(h, world)}")
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
