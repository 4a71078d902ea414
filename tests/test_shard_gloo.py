"""Host-side multi-GPU logic on CPU: partition tables and the world_size-2 gloo reduction."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ppmstereo_b200 import shard


def test_clip_windows_match_reference_cut():
    # models/core/ppmstereo.py:277-281 with kernel_size=20 on a 40-frame clip
    assert shard.clip_windows(40, 20) == [(0, 20), (10, 30), (20, 40), (30, 40)]
    assert shard.clip_windows(5, 20) == [(0, 5)]
    assert shard.clip_windows(41, 20)[-1] == (40, 41)          # the T = 1 window the reference also produces
    with pytest.raises(ValueError):
        shard.clip_windows(0, 20)


@pytest.mark.parametrize("n,world", [(64, 1), (64, 2), (64, 8), (5, 8), (0, 4)])
def test_round_robin_partitions_every_unit_once(n, world):
    owned = [shard.assign_round_robin(n, world, r) for r in range(world)]
    flat = sorted(i for o in owned for i in o)
    assert flat == list(range(n))
    assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1


def test_balanced_assignment_of_unequal_windows():
    hw = [46 * 80, 92 * 160, 184 * 320]
    costs = [shard.window_cost(e - s, hw, [10, 10, 20]) for s, e in shard.clip_windows(40, 20)]
    assert costs[0] == costs[1] == costs[2] and abs(costs[3] / costs[0] - 0.5) < 1e-12
    for world in (1, 2, 3, 4):
        owned = shard.assign_balanced(costs, world)
        assert sorted(i for o in owned for i in o) == [0, 1, 2, 3]
    two = shard.assign_balanced(costs, 2)
    loads = [sum(costs[i] for i in o) for o in two]
    assert max(loads) / sum(loads) <= 2.0 / 3.5 + 1e-9          # best possible split of 1,1,1,0.5


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard.assign_round_robin(7, world, rank)                   # 7 clips of 40 frames
        frames = 40.0 * len(mine)
        ms = 1000.0 * (rank + 1)                                          # rank 1 is the slow one
        tot, mx = shard.reduce_throughput(frames, ms)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            out.put((tot, mx, gathered))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_reduction():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    tot, mx, gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert tot == 280.0 and mx == 2000.0                                  # all 7 clips, time of the slowest rank
    assert sorted(i for g in gathered for i in g) == list(range(7))


def test_reduce_without_process_group_is_local():
    assert shard.reduce_throughput(40.0, 12.5) == (40.0, 12.5)


# ---------------------------------------------------------------- row-band split (optional)
from ppmstereo_b200 import rowband


@pytest.mark.parametrize("h,world", [(184, 8), (46, 4), (5, 2), (3, 4), (0, 2)])
def test_row_bands_cover_rows_once(h, world):
    bands = rowband.row_bands(h, world)
    assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == h
    assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
    sizes = [b - a for a, b in bands]
    assert max(sizes) - min(sizes) <= 1


def _band_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h = 5                                                       # uneven: bands of 3 and 2 rows
        full = torch.arange(2 * 3 * 4 * h * 6, dtype=torch.float32).reshape(2, 3, 4, h, 6)
        r0, r1 = rowband.row_bands(h, world)[rank]
        got = rowband.gather_bands(full[:, :, :, r0:r1].contiguous(), h, 3)
        unc = torch.arange(8 * h * 6, dtype=torch.float32).reshape(8, 1, h, 6)
        got_u = rowband.gather_bands(unc[:, :, r0:r1].contiguous(), h, 2)
        out.put((rank, bool(torch.equal(got, full)), bool(torch.equal(got_u, unc))))
    finally:
        dist.destroy_process_group()


def test_gather_bands_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_band_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True, True), (1, True, True)]
