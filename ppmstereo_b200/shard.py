"""Multi-GPU partitioning of the hot path: one process per GPU, units of work are
independent, so there is NO data-path collective -- the only communication is the reduction of the
timing / frame counters at the end (NCCL on GPUs, gloo in the CPU tests).

Two granularities, both exchange-free:
  * clips   -- ``forward_batch_test`` keeps no state across clips (models/core/ppmstereo.py:238-320);
  * windows -- the temporal windows of one clip are independent too (:277-307 only concatenates them),
               so a single clip can use up to ``len(windows)`` GPUs.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def clip_windows(n_frames: int, kernel_size: int) -> List[Tuple[int, int]]:
    """[start, end) of every temporal window, as forward_batch_test cuts them (ppmstereo.py:277-281):
    stride kernel_size // 2, last windows truncated at the clip end."""
    if n_frames <= 0 or kernel_size <= 1:
        raise ValueError("n_frames must be > 0 and kernel_size > 1")
    stride = kernel_size // 2
    return [(i, min(i + kernel_size, n_frames)) for i in range(0, n_frames, stride)]


def assign_round_robin(n_units: int, world: int, rank: int) -> List[int]:
    """Units (clips or windows) owned by ``rank``: i with i % world == rank."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_units, world))


def assign_balanced(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time assignment of units with unequal cost (windows of 20 and 10 frames differ
    4x in attention FLOPs).  Deterministic: every rank computes the same table without communication."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    load = [0.0] * world
    owned: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        owned[r].append(i)
        load[r] += costs[i]
    for o in owned:
        o.sort()
    return owned


def window_cost(T: int, hw_levels: Sequence[int], iters: Sequence[int], top_k: int = 5, c: int = 128) -> float:
    """Attention FLOPs of one window: sum over levels of iters * 4 * min(K,T) * c * hw^2 * T."""
    return float(sum(it * 4.0 * min(top_k, T) * c * float(hw) ** 2 * T for hw, it in zip(hw_levels, iters)))


def reduce_throughput(units_done: float, elapsed_ms: float, device=None, group=None) -> Tuple[float, float]:
    """(total units over all ranks, max elapsed ms over ranks).  Works on any initialised backend; with no
    process group it returns the local values."""
    if not (dist.is_available() and dist.is_initialized()):
        return float(units_done), float(elapsed_ms)
    dev = device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu")
    tot = torch.tensor([float(units_done)], dtype=torch.float64, device=dev)
    mx = torch.tensor([float(elapsed_ms)], dtype=torch.float64, device=dev)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    return float(tot.item()), float(mx.item())
