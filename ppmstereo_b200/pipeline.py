"""Host <-> device staging around the hot path (window scheduler + I/O).

The reference moves every temporal window to the GPU and back synchronously
(``forward_batch_test``, models/core/ppmstereo.py:287-294: ``.cuda()`` before and ``.cpu()`` after each
window).  ``StagedRunner`` keeps the same per-job semantics -- inputs live in host memory, results are
returned in host memory -- but overlaps the copies with the kernels:

  * job i+1's inputs are uploaded on a copy stream while job i computes;
  * job i's results are downloaded on a second copy stream into pinned buffers while job i+1 computes.

Jobs are independent units (one window x cascade level, or one clip); there is no data-path collective.
CUDA only: there is no CPU path.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch


class StagedRunner:
    def __init__(self, device=None, depth: int = 2):
        if not torch.cuda.is_available():
            raise RuntimeError("StagedRunner needs a CUDA device (ppmstereo_b200 has no CPU path)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.depth = max(1, int(depth))
        self.h2d = torch.cuda.Stream(self.device)
        self.d2h = torch.cuda.Stream(self.device)
        self._host_out: Dict[Tuple, torch.Tensor] = {}

    # -- pinned result buffers, one per (result index, shape, dtype, slot), allocated when a slot is first used and
    #    reused by later run() calls
    def _out_buffer(self, t: torch.Tensor, slot: int, idx: int) -> torch.Tensor:
        key = (idx, tuple(t.shape), t.dtype, slot)
        buf = self._host_out.get(key)
        if buf is None:
            buf = self._host_out[key] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        return buf

    def _upload(self, job: Dict[str, torch.Tensor], after=None):
        with torch.cuda.stream(self.h2d):
            if after is not None:
                self.h2d.wait_event(after)                 # at most two jobs' inputs resident at a time
            dev = {}
            for k, v in job.items():
                if not v.is_pinned():
                    raise RuntimeError(f"input '{k}' must be in pinned host memory (torch.Tensor.pin_memory())")
                dev[k] = v.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.h2d)
        return dev, ev

    def run(self, jobs: Sequence[Dict[str, torch.Tensor]],
            fn: Callable[[Dict[str, torch.Tensor], int], Iterable[torch.Tensor]],
            on_result: Optional[Callable[[int, List[torch.Tensor]], None]] = None) -> List[List[torch.Tensor]]:
        """Run ``fn(device_inputs, job_index) -> device tensors`` over ``jobs`` (dicts of pinned host
        tensors); returns only after every copy has finished.

        Without ``on_result``: returns, per job, the results in pinned host memory -- one buffer set per job, no
        two jobs share one (the sets are kept and reused by the next ``run`` of the same length / shapes).
        With ``on_result(i, host_tensors)``: results stream through a ring of ``depth`` pinned buffer sets; the
        callback is invoked on the host once job i's download has completed and before its buffers are reused
        (copy or consume them there); ``run`` then returns an empty list."""
        compute = torch.cuda.current_stream(self.device)
        results: List[List[torch.Tensor]] = []
        done_copy: List[torch.cuda.Event] = []
        computed: List[torch.cuda.Event] = []
        slots = self.depth if on_result is not None else max(1, len(jobs))
        delivered = 0

        def deliver(upto: int) -> int:
            d = delivered
            while on_result is not None and d < upto:
                done_copy[d].synchronize()
                on_result(d, results[d])
                d += 1
            return d
        nxt = self._upload(jobs[0]) if jobs else None
        for i, _ in enumerate(jobs):
            dev, ready = nxt
            nxt = self._upload(jobs[i + 1], computed[i - 1] if i >= 1 else None) if i + 1 < len(jobs) else None
            compute.wait_event(ready)
            for t in dev.values():
                t.record_stream(compute)                   # allocated on the copy stream, used here
            outs = list(fn(dev, i))
            finished = torch.cuda.Event()
            finished.record(compute)
            computed.append(finished)
            if on_result is not None and i >= self.depth:
                delivered = deliver(i - self.depth + 1)     # job i - depth consumed: its ring slot is free again
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(finished)
                host = []
                for n, t in enumerate(outs):
                    t.record_stream(self.d2h)
                    hb = self._out_buffer(t, i % slots, n)
                    hb.copy_(t, non_blocking=True)
                    host.append(hb)
                ev = torch.cuda.Event()
                ev.record(self.d2h)
            done_copy.append(ev)
            results.append(host)
        for ev in done_copy:
            ev.synchronize()
        if on_result is not None:
            deliver(len(results))
            return []
        return results
