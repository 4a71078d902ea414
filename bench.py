"""bench.py — frames/s of the PPMStereo hot path (correlation pyramid + lookup + PPM memory read)
on synthetic 720p 40-frame clips, one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one 40-frame 720x1280 clip exactly as
PPMStereoModel.forward schedules it (reference models/core/ppmstereo.py:238-320, :601-810):
4 temporal windows [20, 20, 20, 10 frames], per window 3 cascade levels (1/16, 1/8, 1/4 of the
736x1280 padded image) with iters//2, iters//2, iters = 10, 10, 20 update iterations; per level one
correlation-pyramid build + one query pack, per iteration one lookup + one QAM pick + K'/V pack
+ one memory read over all T query frames.  The PyTorch glue between the kernels (encoders,
GRU) is not part of the path and is not run; its outputs are replaced by synthetic tensors of
the shapes and dtypes the autocast model hands over (bf16 features/values, fp32 flow).
At N > 1 every rank processes its own clip (weak scaling, no data-path collective).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frames/s @720p 40-frame clips (hot path: corr pyramid + lookup + PPM memory read)"
CLIP_FRAMES, KERNEL_SIZE, ITERS, TOP_K, D_FEAT, C_QKV = 40, 20, 20, 5, 256, 128
PAD_H, PAD_W = 736, 1280                      # 720 -> 736 rows (InputPadder divis_by=32, utils.py:22-38)
LEVELS = [(16, ITERS // 2), (8, ITERS // 2), (4, ITERS)]   # (downscale, iterations)


def clip_windows():
    from ppmstereo_b200 import shard       # host-side partitioning logic (no CUDA needed to import)
    return [e - s for s, e in shard.clip_windows(CLIP_FRAMES, KERNEL_SIZE)]     # [20, 20, 20, 10]


def attention_flops(T: int, hw: int, k: int = TOP_K) -> float:
    """4*K*c*hw^2 per query frame (QK^T + PV, 2 FLOP/MAC), T query frames per launch."""
    return 4.0 * min(k, T) * C_QKV * float(hw) ** 2 * T


def clip_attention_flops() -> float:
    tot = 0.0
    for T in clip_windows():
        for ds, it in LEVELS:
            tot += it * attention_flops(T, (PAD_H // ds) * (PAD_W // ds))
    return tot


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        load = [x for x in sm if x >= 0.5 * (mx[0] if mx else 1)] or sm
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# our arm: the CUDA path through the public Python API (which calls the C ABI)
# --------------------------------------------------------------------------------------------
class LevelInputs:
    """Synthetic inputs of one (window length T, cascade level) in pinned host memory and in HBM."""
    NAMES = ("fmap1", "fmap2", "query", "key", "value", "mf", "unc", "flow_a", "flow_b", "sim")

    def __init__(self, T: int, ds: int, seed: int, device):
        h, w = PAD_H // ds, PAD_W // ds
        self.T, self.h, self.w, self.ds = T, h, w, ds
        g = torch.Generator(device=device).manual_seed(seed)
        rn = lambda *shape: torch.randn(*shape, generator=g, device=device)
        dev = {
            "fmap1": rn(T, D_FEAT, h, w).bfloat16(), "fmap2": rn(T, D_FEAT, h, w).bfloat16(),
            "query": rn(1, C_QKV, T, h, w).bfloat16(), "key": rn(1, C_QKV, T, h, w).bfloat16(),
            "value": rn(1, C_QKV, T, h, w).bfloat16(), "mf": rn(1, C_QKV, T, h, w).bfloat16(),
            "unc": torch.rand(T, 1, h, w, generator=g, device=device).bfloat16(),
            "sim": torch.rand(1, 1, T, T, generator=g, device=device) * 2 - 1,
        }
        for n in ("flow_a", "flow_b"):
            f = torch.zeros(T, 2, h, w, device=device)
            f[:, 0] = torch.rand(T, h, w, generator=g, device=device) * (-160.0 / ds)
            dev[n] = f
        self.dev = dev
        self.host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in dev.items()}
        self.in_bytes = sum(v.numel() * v.element_size() for v in dev.values())
        self.out_bytes = T * C_QKV * h * w * 4 + T * 36 * h * w * 4


class Counters:
    launches = 0
    attend_events = []          # (start, end) CUDA events of the 1/4-res memory-read launches


def run_level(P, t, iters: int, beta, time_attend: bool):
    """One cascade level of one window on device tensors `t`.  Returns the tensors a caller reads back."""
    blk = P.CorrBlock1D(t["fmap1"], t["fmap2"])                       # K1
    mem = P.PPMMemory(t["query"], t["key"], None, top_k=TOP_K)          # pack Q' once per level
    strive = torch.ones_like(t["sim"])
    Counters.launches += 2
    out = corr = None
    for it in range(iters):
        corr = blk(t["flow_a"] if it % 2 == 0 else t["flow_b"])         # K2
        mem.pick(t["sim"], strive, t["unc"])                            # K3a (2 launches)
        mem.pack_kv(t["value"])                                         # K3b (2 launches)
        if time_attend:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        out = mem.aggregate(t["mf"], beta, out=out)                     # K3c
        if time_attend:
            e1.record()
            Counters.attend_events.append((e0, e1))
        Counters.launches += 6
    return out, corr


def clip_jobs():
    return [(T, ds, iters) for T in clip_windows() for ds, iters in LEVELS]


def run_clip(P, inputs, beta, from_host: bool, time_attend: bool = False, runner=None):
    jobs = clip_jobs()
    if not from_host:
        for T, ds, iters in jobs:
            run_level(P, inputs[(T, ds)].dev, iters, beta, time_attend and ds == 4 and T == KERNEL_SIZE)
        return None
    # end to end: every window/level's inputs come from pinned host memory and its results go back to
    # host memory (ppmstereo_b200.StagedRunner: copies on side streams, overlapped with the kernels)
    return runner.run([inputs[(T, ds)].host for T, ds, _ in jobs],
                      lambda dev, i: run_level(P, dev, jobs[i][2], beta, False))


# --------------------------------------------------------------------------------------------
# CPU arm: the reference's path on the host cores (bounded sample)
#   kind "reference": the reference's OWN code -- CorrBlock1D.__init__ / __call__ (models/core/corr.py:56-94) and
#     PPMStereo.forward_update_block (models/core/ppmstereo.py:426-594), imported unmodified from the staged copy
#     under oracle/_ref (or /root/reference), with the glue modules it calls replaced by constants exactly as this
#     file's GPU arm replaces them by synthetic tensors, and the CUDA-only flash_attn by the CPU stand-in of
#     oracle/ref_import.py (the one third-party dependency on the path);
#   kind "port": oracle/ppm_oracle.py, when no reference tree is available.
# --------------------------------------------------------------------------------------------
SAMPLE_DS, SAMPLE_T, SAMPLE_ITERS = 16, KERNEL_SIZE, 2
SAMPLE_WHAT = ("pyramid build + 2 update iterations of the 1/16-res level of a 20-frame 720p window (h x w = 46 x 80, hw = 3680, "
               "T = 20, K = 5; each iteration = lookup + full PPM read; about 10-25 s)")
SAMPLE_DESC = {
    "reference": SAMPLE_WHAT + ": the reference's own CorrBlock1D + PPMStereo.forward_update_block (unmodified, oracle/_ref) "
                 "with constant glue outputs and the CPU stand-in for CUDA-only flash_attn, fp32, all host threads; frames/s "
                 "extrapolated to the 40-frame clip by attention FLOPs",
    "port": SAMPLE_WHAT + ", oracle/ppm_oracle.py (port of the reference's PyTorch CPU path) on all host threads; frames/s "
            "extrapolated to the 40-frame clip by attention FLOPs",
}


def make_cpu_sample(seed: int = 0):
    h, w, T = PAD_H // SAMPLE_DS, PAD_W // SAMPLE_DS, SAMPLE_T
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    flow = torch.zeros(T, 2, h, w)
    flow[:, 0] = torch.rand(T, h, w, generator=g) * -10.0
    return dict(f1=rn(T, D_FEAT, h, w), f2=rn(T, D_FEAT, h, w), q=rn(1, C_QKV, T, h, w), k=rn(1, C_QKV, T, h, w),
                v=rn(1, C_QKV, T, h, w), mf=rn(1, C_QKV, T, h, w), unc=torch.rand(T, 1, h, w, generator=g),
                sim=torch.rand(1, 1, T, T, generator=g) * 2 - 1, strive=torch.ones(1, 1, T, T), flow=flow,
                up_mask=rn(T, 144, h, w))


def cpu_kind() -> str:
    from oracle import ref_import
    return "reference" if ref_import.available() else "port"


_REF_SHELL = {}


def _reference_shell():
    """The reference's own loop shell + correlation class, bound to constant glue (see the section header)."""
    if _REF_SHELL:
        return _REF_SHELL
    import importlib
    from oracle import ref_import
    ref_import.load_reference()
    ppm = importlib.import_module("models.core.ppmstereo")

    class Shell:                                         # borrows the reference's unbound methods; no parameters needed
        mixed_precision = False
        use_convex_3d = False
        forward_update_block = ppm.PPMStereo.forward_update_block
        compute_qk_similarity = ppm.PPMStereo.compute_qk_similarity
        convex_upsample = ppm.PPMStereo.convex_upsample

    _REF_SHELL.update(ppm=ppm, shell=Shell())
    return _REF_SHELL


class _ConstGlue:
    """Stands where SequenceUpdateBlock3D / Attention_qk stand; returns the sample's constant tensors."""

    def __init__(self, s):
        T = s["q"].shape[2]
        self.s = s
        self.mf = s["mf"][0].permute(1, 0, 2, 3).contiguous()                  # (T, 128, h, w)
        self.v = s["v"][0].permute(1, 0, 2, 3).contiguous()
        self.qk = torch.cat([s["q"][0].permute(1, 0, 2, 3), s["k"][0].permute(1, 0, 2, 3)], dim=1).contiguous()
        self.aggregator = type("Agg", (), {"beta": torch.tensor([0.5])})()
        self.delta = torch.zeros_like(s["flow"])

    def to_qk(self, inp):
        return self.qk

    def get_motion_and_value(self, flow, corr, mhs, inp):
        return self.mf, mhs, self.v

    def get_uncertainty(self, x):
        return self.s["unc"]

    def __call__(self, net, inp, mf, mfg, t=1):
        return net, self.s["up_mask"], self.delta


def cpu_sample_step(s, kind: str) -> float:
    if kind == "reference":
        r = _reference_shell()
        glue = _ConstGlue(s)
        T, _, h, w = s["f1"].shape
        inp = torch.zeros(T, C_QKV, h, w)
        net = torch.zeros(T, C_QKV, h, w)
        from oracle import ref_import
        saved = r["ppm"].flash_attn_func
        r["ppm"].flash_attn_func = ref_import._flash_attn_cpu    # CUDA-only library -> CPU stand-in, for this call only
        import contextlib
        try:
            t0 = time.perf_counter()
            with torch.no_grad(), contextlib.redirect_stdout(sys.stderr):     # stdout carries the one JSON line only
                blk = r["ppm"].CorrBlock1D(s["f1"], s["f2"])
                r["shell"].forward_update_block(image1=None, update_block=glue, corr_fn=blk, flow=s["flow"], net=net,
                                                inp=inp, motion_hidden_state=None, attn_block=glue, predictions=[],
                                                uncertainties=[], iters=SAMPLE_ITERS, interp_scale=4, t=T)
            return time.perf_counter() - t0
        finally:
            r["ppm"].flash_attn_func = saved
    from oracle import ppm_oracle as O          # the CPU baseline is the one place bench.py runs the oracle
    t0 = time.perf_counter()
    lv = O.corr_pyramid(s["f1"], s["f2"], mode="bf16")
    strive = s["strive"].clone()
    for _ in range(SAMPLE_ITERS):
        O.corr_lookup(lv, s["flow"])
        O.ppm_memory_read(s["q"], s["k"], O.temporal_pe(SAMPLE_T, C_QKV), s["v"], s["mf"], s["sim"], strive,
                          s["unc"], 0.5)
    return time.perf_counter() - t0


def cpu_frames_per_s(seconds_per_sample: float) -> float:
    f_sample = SAMPLE_ITERS * attention_flops(SAMPLE_T, (PAD_H // SAMPLE_DS) * (PAD_W // SAMPLE_DS))
    return CLIP_FRAMES / (seconds_per_sample * clip_attention_flops() / f_sample)


def run_reference_arm(args, rank: int):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    s = make_cpu_sample()
    kind = cpu_kind()
    for _ in range(min(args.warmup, 1)):
        cpu_sample_step(s, kind)
    ts = [cpu_sample_step(s, kind) for _ in range(args.steps)]
    sec = sum(ts) / len(ts)
    v = cpu_frames_per_s(sec)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": SAMPLE_DESC[kind]},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(n_gpus: int):
    return {"workload": "hot path of PPMStereoModel.forward on one synthetic 40-frame 720x1280 clip per GPU "
                        "(BASELINE configs[3]/[4]): 4 windows [20,20,20,10], levels 1/16,1/8,1/4 of 736x1280, "
                        "iters 10/10/20, D=256 bf16 features, c=128, top-5 of T frames",
            "clips_per_gpu": 1, "kernel_size": KERNEL_SIZE, "iters": ITERS, "top_k": TOP_K,
            "parallelism": f"clip-sharded x{n_gpus} (no data-path collective)",
            "l2_policy": "inputs larger than L2 (per-window working set > 3 GB vs 126 MB L2), no flush needed"}


# --------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling aid: skip the host-buffer pass")
    ap.add_argument("--no-same-box", action="store_true", help="profiling aid: skip the flash_attn comparison")
    ap.add_argument("--full-model", action="store_true",
                    help="also run the whole model (PPMStereo.forward_batch_test, BASELINE configs[3]) on this GPU, the "
                         "reference's own stack vs ppmstereo_b200.install(), with a phase split: extra key `full_model` "
                         "(adds ~5 min; needs the staged reference under oracle/_ref)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch.distributed as dist
    import ppmstereo_b200 as P          # fails loudly without libppm_sm100.so
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    inputs = {(T, ds): LevelInputs(T, ds, 1000 * rank + 10 * T + ds, dev)
              for T in sorted(set(clip_windows())) for ds, _ in LEVELS}
    beta = torch.tensor([0.5], device=dev)
    h2d = sum(inputs[(T, ds)].in_bytes for T in clip_windows() for ds, _ in LEVELS)
    d2h = sum(inputs[(T, ds)].out_bytes for T in clip_windows() for ds, _ in LEVELS)

    runner = P.StagedRunner(dev)

    def timed(from_host: bool, steps: int, time_attend: bool):
        barrier()
        sampler = ClockSampler(local)
        if rank == 0 and not from_host:
            sampler.start()
        Counters.launches = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            run_clip(P, inputs, beta, from_host, time_attend, runner)
        e1.record()
        barrier()
        # every rank ran `steps` clips of its own; time = max over ranks (device events)
        _, ms = P.shard.reduce_throughput(CLIP_FRAMES * steps, e0.elapsed_time(e1), device=dev)
        clocks = sampler.stop() if rank == 0 and not from_host else None
        return ms / steps, clocks

    for _ in range(args.warmup):
        run_clip(P, inputs, beta, False)
    Counters.attend_events.clear()
    ms_step, clocks = timed(False, args.steps, True)
    launches = Counters.launches // args.steps
    att_ms = [a.elapsed_time(b) for a, b in Counters.attend_events]
    ms_e2e = float("nan")
    if not args.no_e2e:
        run_clip(P, inputs, beta, True, False, runner)   # warm the pinned-copy path once
        ms_e2e, _ = timed(True, args.steps, False)

    # ---- N >= 2 only: the optional splits of ONE clip; the weak-scaling line above stays the headline
    multi = None
    if world > 1:
        multi = {"rowband_bit_exact": bool(P.rowband.verify_against_full_read(dev))}
        # (a) strong scaling by temporal windows: the 4 windows of one clip are independent (ppmstereo.py:277-307),
        #     LPT-assigned to min(N, 4) ranks by attention FLOPs; no exchange; time = max over ranks
        wins = clip_windows()
        hw_l = [(PAD_H // ds) * (PAD_W // ds) for ds, _ in LEVELS]
        costs = [P.shard.window_cost(T, hw_l, [it for _, it in LEVELS]) for T in wins]
        owned = P.shard.assign_balanced(costs, min(world, len(wins)))
        mine = owned[rank] if rank < len(owned) else []

        def one_clip_by_windows():
            for wi in mine:
                for ds, iters in LEVELS:
                    run_level(P, inputs[(wins[wi], ds)].dev, iters, beta, False)

        one_clip_by_windows()
        barrier()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        one_clip_by_windows()
        b_.record()
        barrier()
        _, ms_w = P.shard.reduce_throughput(0, a.elapsed_time(b_), device=dev)
        multi["one_clip_window_split"] = {
            "scaling": "strong", "ranks_used": len(owned), "windows_per_rank": [[wins[i] for i in o] for o in owned],
            "ms_per_clip": ms_w, "frames_per_s": CLIP_FRAMES / (ms_w * 1e-3),
            "speedup_vs_1gpu": ms_step / ms_w,
            "note": "hot path of ONE 40-frame clip, windows [20,20,20,10] LPT-assigned by attention FLOPs; no collective; "
                    "ceiling 3.5x (the largest share is one 20-frame window of 3.5 window-equivalents)"}
        # (b) row-band split of one window's memory read at the 1/4 level: per iteration the value + uncertainty bands
        #     are all-gathered (NCCL), the pick is replicated, every rank aggregates its band's queries
        x = inputs[(KERNEL_SIZE, 4)].dev
        h4 = PAD_H // 4
        r0, r1 = P.row_bands(h4, world)[rank]
        band = P.RowBandMemory(x["query"], x["key"], None, top_k=TOP_K)
        vb = x["value"][:, :, :, r0:r1].contiguous()
        mb = x["mf"][:, :, :, r0:r1].contiguous()
        ub = x["unc"][:, :, r0:r1].contiguous()
        strive = torch.ones_like(x["sim"])

        def timed_ms(fn, reps=3):
            fn()
            barrier()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            e[0].record()
            for _ in range(reps):
                fn()
            e[1].record()
            barrier()
            return P.shard.reduce_throughput(0, e[0].elapsed_time(e[1]) / reps, device=dev)[1]

        ms_band = timed_ms(lambda: band.read(vb, mb, x["sim"], strive, ub, beta))
        ms_gather = timed_ms(lambda: (P.gather_bands(vb, h4, 3), P.gather_bands(ub, h4, 2)))
        full = P.PPMMemory(x["query"], x["key"], None, top_k=TOP_K)
        ms_full = timed_ms(lambda: full.read(x["value"], x["mf"], x["sim"], strive, x["unc"], beta))
        multi["rowband_memory_read"] = {
            "scaling": "strong", "level": "1/4 res, T=20, hw=58880, K=5", "ms_per_iteration_1gpu": ms_full,
            "ms_per_iteration_banded": ms_band, "speedup": ms_full / ms_band, "ideal": float(world),
            "allgather_ms": ms_gather, "allgather_bytes": int(x["value"].numel() * 2 + x["unc"].numel() * 2),
            "collective": "NCCL all_gather_into_tensor of the bf16 value bands (+ uncertainty bands) per iteration; "
                          "it is the only exchange of the split and is not overlapped with the attention",
            "note": "memory read only: the PyTorch glue convolutions have no halo exchange (out of scope), so a whole "
                    "clip is split by windows, (a), not by bands"}
        del band, full, vb, mb, ub

    def time_kernel(fn, reps=10, rounds=5):
        """Kernel duration free of the host launch path: `reps` back-to-back launches captured in one CUDA
        graph, the replay timed with CUDA events on its stream, median of `rounds` replays / reps."""
        fn()
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            fn()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):   # NCCL watchdog thread may poll events
                for _ in range(reps):
                    fn()
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(rounds):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            graph.replay()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) / reps)
        return sorted(ts)[len(ts) // 2]

    other = []
    if rank == 0:
        # K1 / K2 at the 1/4-res level of a 20-frame window, timed alone (HBM-bound kernels; burst copy peak)
        x = inputs[(KERNEL_SIZE, 4)].dev
        T4, h4, w4 = KERNEL_SIZE, PAD_H // 4, PAD_W // 4
        hbm = float(peaks()[0].get("hbm_gbs", 6650.0))
        lv = P.build_pyramid(x["fmap1"], x["fmap2"])
        ms1 = time_kernel(lambda: P.build_pyramid(x["fmap1"], x["fmap2"], levels=lv))
        del lv
        b1 = 2.0 * T4 * D_FEAT * h4 * w4 * 2 + 1.9375 * T4 * h4 * w4 * w4 * 2
        blk = P.CorrBlock1D(x["fmap1"], x["fmap2"])
        ms2 = time_kernel(lambda: blk(x["flow_a"]), reps=20)
        b2 = 228.0 * T4 * h4 * w4
        gk = torch.Generator(device=dev).manual_seed(5)
        up_mask = torch.randn(T4, 144, h4, w4, generator=gk, device=dev).bfloat16()       # autocast dtype of mask_2d
        ms4 = time_kernel(lambda: P.convex_upsample(x["flow_a"], up_mask))
        b4 = T4 * h4 * w4 * (144.0 * 2 + 8 + 128)
        del up_mask
        for name, ms, byts, traffic in (("corr_build_kernel (K1, 1/4-res, T=20, D=256, bf16 in/out)", ms1, b1, 2.63e9),
                                        ("corr_lookup_l0_kernel (K2b, 1/4-res, T=20, 4 levels from one level-0 span, r=4)", ms2, b2, 4.70e8),
                                        ("convex_upsample_bf16x2_kernel (K4, SURVEY 8f glue op, 1/4-res, T=20, bf16 mask; "
                                         "not part of the timed step)", ms4, b4, 4.71e8)):
            other.append({"kernel": name, "bound": "hbm", "achieved": byts / (ms * 1e-3) / 1e9, "peak": hbm,
                          "unit": "GB/s", "frac": byts / (ms * 1e-3) / 1e9 / hbm, "algorithmic_bytes": byts,
                          "traffic": traffic, "avg_launch_ms": ms,
                          "note": "average duration of 10 back-to-back launches replayed from a CUDA graph (no host launch "
                                  "path in the number), CUDA events on the replay stream; traffic = dram bytes of "
                                  "one ncu --set full capture (profiles/)"})
        del blk
        # K5: SKSepConvGRU3D at the same level, tensor-core bound.  The layer container below only
        # states the unit's Conv3d shapes (include/ppm_b200.h, K5) with random-init weights; no reference code involved.
        import torch.nn as nn

        def _c(ci, k):
            return nn.Conv3d(ci, 128, k, padding=tuple(x // 2 for x in k))

        class _GruLayers(nn.Module):
            def __init__(self):
                super().__init__()
                self.convz1 = nn.Sequential(_c(512, (1, 1, 15)), nn.GELU(), _c(128, (1, 1, 5)))
                self.convr1 = nn.Sequential(_c(512, (1, 1, 15)), nn.GELU(), _c(128, (1, 1, 5)))
                self.convq1 = _c(512, (1, 1, 5))
                self.convz2, self.convr2, self.convq2 = _c(512, (1, 5, 1)), _c(512, (1, 5, 1)), _c(512, (1, 5, 1))
                self.convz3, self.convr3, self.convq3 = _c(512, (5, 1, 1)), _c(512, (5, 1, 1)), _c(512, (5, 1, 1))

        torch.manual_seed(11)
        k5 = P.SepConvGRU3D(_GruLayers().to(dev))
        gh = torch.tanh(torch.randn(1, 128, T4, h4, w4, generator=gk, device=dev))
        gx = torch.randn(1, 384, T4, h4, w4, generator=gk, device=dev)
        ms5 = time_kernel(lambda: k5(gh, gx), reps=3)
        f5 = 2.0 * T4 * h4 * w4 * (512 * 128 * 35 + 2 * 128 * 128 * 5 + 6 * 512 * 128 * 5)
        tpk = float(peaks()[0].get("bf16_tflops_sustained", 1400.0))
        other.append({"kernel": "gru_gemm_kernel x 8 + packs (K5 SKSepConvGRU3D, SURVEY 8f rank 2, 1/4-res, T=20; the unit the "
                                "installed model calls once per iteration; not part of the timed hot-path step)",
                      "bound": "tensor", "achieved": f5 / (ms5 * 1e-3) / 1e12, "peak": tpk, "unit": "TFLOP/s",
                      "frac": f5 / (ms5 * 1e-3) / 1e12 / tpk, "flops_per_call": f5, "avg_call_ms": ms5, "traffic": None,
                      "note": "one whole unit (2 pack launches + 8 GEMM launches) replayed from a CUDA graph; FLOPs = 8.85 MFLOP "
                              "per position (SURVEY 8f / DESIGN K5); ncu per-GEMM breakdown in profiles/r02_k5_gru.md"})
        del k5, gh, gx

        def _c2(ci, co, k, groups=1, bias=True):
            return nn.Conv2d(ci, co, k, padding=k // 2, groups=groups, bias=bias)

        class _FrontLayers(nn.Module):                  # PCBlock4_Deep_nopool_res(36, 256, [1, 7]) layer shapes
            def __init__(self):
                super().__init__()
                self.ffn1 = nn.Sequential(_c2(36, 54, 1), nn.GELU(), _c2(54, 36, 1))
                self.conv_list = nn.ModuleList([_c2(36, 36, 1, groups=36), _c2(36, 36, 7, groups=36)])
                self.pw = _c2(36, 36, 1)
                self.ffn2 = nn.Sequential(_c2(36, 54, 1), nn.GELU(), _c2(54, 256, 1))

        class _EncoderLayers(nn.Module):                # BasicMotionEncoder_v2 layer shapes
            def __init__(self):
                super().__init__()
                self.convc1, self.convc2 = _FrontLayers(), _c2(256, 192, 3)
                self.convf1, self.convf2, self.final_conv = _c2(2, 128, 7), _c2(128, 64, 3), _c2(320, 190, 3)

        class _ValueLayers(nn.Module):                  # aggregator.to_v + the uncertainty head
            def __init__(self):
                super().__init__()
                self.aggregator = nn.Module()
                self.aggregator.to_v = _c2(128, 128, 1, bias=False)
                self.uncertainty = nn.Sequential(_c2(256, 128, 3), nn.ReLU(), _c2(128, 1, 1), nn.Sigmoid())

        k7 = P.MotionEncoder(_EncoderLayers().to(dev))
        k8 = P.ValueUncertainty(_ValueLayers().to(dev))
        ef = torch.zeros(T4, 2, h4, w4, device=dev)
        ef[:, 0] = torch.rand(T4, h4, w4, generator=gk, device=dev) * -40.0
        f1e = torch.randn(T4, 256, h4, w4, generator=gk, device=dev).bfloat16()
        f2e = torch.randn(T4, 256, h4, w4, generator=gk, device=dev).bfloat16()
        eblk = P.CorrBlock1D(f1e, f2e)
        emh = torch.randn(T4, 64, h4, w4, generator=gk, device=dev).bfloat16()
        enet = torch.tanh(torch.randn(T4, 128, h4, w4, generator=gk, device=dev))
        with torch.autocast("cuda", dtype=torch.bfloat16):
            ms7 = time_kernel(lambda: k7(ef, P.LookupRequest(eblk, ef), emh, None), reps=3)
            emf, _ = k7(ef, P.LookupRequest(eblk, ef), emh, None)
            ms8 = time_kernel(lambda: k8(emf, enet), reps=3)
        pos = float(T4 * h4 * w4)
        f7 = 2.0 * pos * (9 * (256 * 192 + 128 * 64 + 320 * 190) + 98 * 128)
        f8 = 2.0 * pos * (128 * 128 + 9 * 256 * 128 + 128)
        other.append({"kernel": "K7 ppm_motion_encoder (BasicMotionEncoder_v2.forward incl. the fused lookup: im2col + K6a + K6b + "
                                "4 GEMMs + 1 pack; 1/4-res, T=20; not part of the timed hot-path step)",
                      "bound": "tensor", "achieved": f7 / (ms7 * 1e-3) / 1e12, "peak": tpk, "unit": "TFLOP/s",
                      "frac": f7 / (ms7 * 1e-3) / 1e12 / tpk, "flops_per_call": f7, "avg_call_ms": ms7, "traffic": None,
                      "note": "FLOPs = the four convolutions on the tensor cores only; 3.2 of the call's ~6.2 ms are the CUDA-core "
                              "front end K6a/K6b (DESIGN K6), so frac understates the GEMMs: ncu per launch in "
                              "profiles/r02_k7_motion_encoder.md; torch/cuDNN path of the reference: 13.1 ms"})
        other.append({"kernel": "K8 ppm_value_uncertainty (to_v + uncertainty head: 2 packs + 3 GEMMs; 1/4-res, T=20)",
                      "bound": "tensor", "achieved": f8 / (ms8 * 1e-3) / 1e12, "peak": tpk, "unit": "TFLOP/s",
                      "frac": f8 / (ms8 * 1e-3) / 1e12 / tpk, "flops_per_call": f8, "avg_call_ms": ms8, "traffic": None,
                      "note": "0.76 TFLOP per call, of which the 3x3 convolution 0.70; the two K = 128 GEMMs are epilogue / store bound"})
        del k7, k8, ef, f1e, f2e, eblk, emh, enet, emf
        same_box = None
        try:                                            # the library the reference calls at ppmstereo.py:550, same box
            if args.no_same_box:
                raise ImportError("skipped")
            import flash_attn
            g2 = torch.Generator(device=dev).manual_seed(6)
            hw4_ = h4 * w4
            qf = torch.randn(1, hw4_, 1, C_QKV, generator=g2, device=dev).bfloat16()
            kf = torch.randn(1, TOP_K * hw4_, 1, C_QKV, generator=g2, device=dev).bfloat16()
            vf = torch.randn(1, TOP_K * hw4_, 1, C_QKV, generator=g2, device=dev).bfloat16()
            msf = time_kernel(lambda: flash_attn.flash_attn_func(qf, kf, vf, 0.0, softmax_scale=P.softmax_scale(C_QKV),
                                                                 causal=False), reps=5)
            same_box = {"library": "flash_attn " + flash_attn.__version__ + " flash_attn_func (the reference's call, "
                                   "ppmstereo.py:550), one query frame, Lq=58880 Lk=294400 d=128",
                        "ms_per_query_frame": msf, "tflops": attention_flops(KERNEL_SIZE, hw4_) / KERNEL_SIZE / (msf * 1e-3) / 1e12}
            del qf, kf, vf
        except ImportError:
            pass

    if rank == 0:
        pk, pk_src = peaks()
        hw4 = (PAD_H // 4) * (PAD_W // 4)
        fl = attention_flops(KERNEL_SIZE, hw4)
        avg = sum(att_ms) / max(1, len(att_ms))
        ach = fl / (avg * 1e-3) / 1e12
        peak = float(pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 1400.0)))
        out = {
            "metric": METRIC, "value": world * CLIP_FRAMES / (ms_step * 1e-3), "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "config": workload_config(world),
            "e2e": None if args.no_e2e else {"value": world * CLIP_FRAMES / (ms_e2e * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": "per window and level: inputs copied from pinned host memory, the level's final "
                            "motion_features_global and lookup read back into pinned host memory; copies inside "
                            "the timed region, overlapped with the kernels by ppmstereo_b200.StagedRunner"},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": "ppm_attend3_kernel (1/4-res level, T=20, hw=58880, K=5)", "bound": "tensor",
                         "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "peak_source": "bf16_tflops_sustained, " + pk_src, "traffic": 2.09e10,
                         "flops_per_launch": fl, "avg_launch_ms": avg, "launches_timed": len(att_ms),
                         "share_of_step": sum(att_ms) / args.steps / ms_step},
            "attention_tflops_whole_step": clip_attention_flops() / (ms_step * 1e-3) / 1e12,
            "roofline_other_kernels": other,
            "same_box_reference_attention": same_box,
        }
        if multi is not None:
            out["one_clip_splits"] = multi
            out["rowband_bit_exact"] = multi["rowband_bit_exact"]
        if not args.no_cpu_baseline and world == 1:           # the CPU arm is timed on rank 0 at N = 1 only
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            s = make_cpu_sample()
            kind = cpu_kind()
            cpu_sample_step(s, kind)
            sec = cpu_sample_step(s, kind)
            out["cpu_baseline"] = {"value": cpu_frames_per_s(sec), "unit": "frames/s", "cores": cores,
                                   "kind": kind, "sample": SAMPLE_DESC[kind], "sample_seconds": sec}
        if args.full_model and world == 1:
            # Whole model (BASELINE configs[3]; SURVEY 8d brackets PPMStereoModel.forward): the reference's own network
            # code is the host application here (encoders, GRU glue, loop shells -- unmodified, staged under oracle/_ref),
            # run (a) with its own einsum / grid_sample / flash_attn stack and (b) after ppmstereo_b200.install().  The
            # loader lives in oracle/ because it imports the reference; no arithmetic of the product path comes from it.
            del inputs
            torch.cuda.empty_cache()
            from oracle import full_model, ref_import
            if ref_import.available():
                import contextlib
                with contextlib.redirect_stdout(sys.stderr):        # the reference prints ("iters 20"); stdout = one JSON line
                    fm = full_model.measure(CLIP_FRAMES, 720, 1280, KERNEL_SIZE, ITERS, reference=True)
                out["full_model"] = fm
            else:
                out["full_model"] = {"unavailable": "reference sources not staged under oracle/_ref (python oracle/stage_ref.py)"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
