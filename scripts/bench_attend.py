"""GPU aid: time ppm_memory_attend at the 1/4-res 720p shape (T=20, hw=58880, K=5), with the SM clock and board
power sampled during the run.
  python scripts/bench_attend.py [T] [h] [w] [reps]"""
import ctypes, os, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P
from ppmstereo_b200 import _lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 20
h = int(sys.argv[2]) if len(sys.argv) > 2 else 184
w = int(sys.argv[3]) if len(sys.argv) > 3 else 320
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
variants = [(3, 0)]
lib = _lib.load()
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, generator=g, device=dev).bfloat16()
q, k, v, mf = rn(1, 128, T, h, w), rn(1, 128, T, h, w), rn(1, 128, T, h, w), rn(1, 128, T, h, w)
unc = torch.rand(T, 1, h, w, generator=g, device=dev).bfloat16()
sim = torch.rand(1, 1, T, T, generator=g, device=dev) * 2 - 1
beta = torch.tensor([0.5], device=dev)
mem = P.PPMMemory(q, k, None, top_k=5)
mem.pick(sim, torch.ones_like(sim), unc)
mem.pack_kv(v)
flops = 4.0 * min(5, T) * 128 * float(h * w) ** 2 * T
def smi_sampler(stop, out):
    try:
        pr = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"],
                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    except OSError:
        return
    for ln in pr.stdout:
        out.append(ln)
        if stop.is_set():
            break
    pr.terminate()

ref = None
for ver, emu in variants:
    out = mem.aggregate(mf, beta)
    torch.cuda.synchronize()
    stop, lines = threading.Event(), []
    th = threading.Thread(target=smi_sampler, args=(stop, lines), daemon=True); th.start()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    e[0].record()
    for i in range(reps):
        mem.aggregate(mf, beta, out=out)
        e[i + 1].record()
    torch.cuda.synchronize()
    ms = [e[i].elapsed_time(e[i + 1]) for i in range(reps)]
    stop.set()
    vals = [[float(x) for x in ln.split(',')] for ln in lines if ln.count(',') == 1]
    clk = sorted(v[0] for v in vals)[len(vals) // 2] if vals else float('nan')
    pw = max((v[1] for v in vals), default=float('nan'))
    hid = (out.float() - mf.float().permute(0, 2, 1, 3, 4).reshape(out.shape)) / 0.5
    if ref is None:
        ref = hid
    d = (hid - ref).abs()
    print(f"v{ver} emu{emu}: {min(ms):8.2f} ms best, {sum(ms)/len(ms):8.2f} avg  -> {flops/min(ms)/1e9:7.1f} TFLOP/s best, "
          f"{flops/(sum(ms)/len(ms))/1e9:7.1f} avg | vs first: max {float(d.max()):.3e} mean {float(d.mean()):.3e} "
          f"(|hidden| max {float(ref.abs().max()):.3f}) | sm clock median {clk:.0f} MHz, power max {pw:.0f} W ({len(vals)} samples)", flush=True)
