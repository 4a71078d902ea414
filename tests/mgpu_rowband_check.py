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
    h = 21                                                    # hw = 840: partial tiles, uneven bands
    ok = P.rowband.verify_against_full_read(dev, shape=(1, 7, h, 40), iterations=3)
    flag = torch.tensor([1 if ok else 0], device=dev)
    if rank == 0:
        print("ROWBAND_OK" if int(flag) == 1 else "ROWBAND_MISMATCH", f"world={world} bands={P.row_bands(h, world)}")
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 1 else 1)


if __name__ == "__main__":
    main()
