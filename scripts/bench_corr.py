"""GPU aid: time K1 (pyramid build) and K2 (lookup) at the 720p 1/4-res shape and report achieved
algorithmic GB/s (K1 2*B*D*H*W*e_in + 1.9375*B*H*W1*W2*e_out; K2 228 B/pixel bf16)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P

B, D, H, W = (int(a) for a in sys.argv[1:5]) if len(sys.argv) > 4 else (20, 256, 184, 320)
dev = torch.device("cuda")
torch.zeros(1, device=dev)
if os.environ.get("PPM_L2_FETCH"):
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(int(os.environ["PPM_L2_FETCH"])))   # cudaLimitMaxL2FetchGranularity
    v = ctypes.c_size_t(); rt.cudaDeviceGetLimit(ctypes.byref(v), 5)
    print("L2 fetch granularity set rc", rc, "now", v.value)
g = torch.Generator(device=dev).manual_seed(1)
f1 = torch.randn(B, D, H, W, generator=g, device=dev).bfloat16()
f2 = torch.randn(B, D, H, W, generator=g, device=dev).bfloat16()
flow = torch.zeros(B, 2, H, W, device=dev)
flow[:, 0] = torch.rand(B, H, W, generator=g, device=dev) * -40.0
peak = 6549.4
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass

def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn(); ev[i + 1].record()
    torch.cuda.synchronize()
    t = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    return t[0], t[len(t) // 2]

def graph_time(fn, reps=10, rounds=5):
    """Kernel time free of the host launch path: `reps` launches captured in one CUDA graph, replayed."""
    fn(); torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
    torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(rounds):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / reps)
    return min(ts), sorted(ts)[len(ts) // 2]

blk = P.CorrBlock1D(f1, f2)
k1_bytes = 2 * B * D * H * W * 2 + 1.9375 * B * H * W * W * 2
best, med = timeit(lambda: P.CorrBlock1D(f1, f2))
print(f"K1 build  B={B} D={D} H={H} W={W}: best {best*1e3:.1f} us median {med*1e3:.1f} us -> "
      f"{k1_bytes/med/1e6:.0f} GB/s algorithmic ({k1_bytes/med/1e6/peak:.2f} of {peak:.0f})")
lv = P.build_pyramid(f1, f2)
import time
t0 = time.perf_counter()
for _ in range(20): P.build_pyramid(f1, f2, levels=lv)
host_us = (time.perf_counter() - t0) / 20 * 1e6
torch.cuda.synchronize()
best, med = graph_time(lambda: P.build_pyramid(f1, f2, levels=lv))
print(f"K1 graph replay: best {best*1e3:.1f} us median {med*1e3:.1f} us -> {k1_bytes/med/1e6:.0f} GB/s algorithmic "
      f"({k1_bytes/med/1e6/peak:.2f} of {peak:.0f}); host launch path {host_us:.0f} us per call")
k2_bytes = B * H * W * 228.0
best, med = timeit(lambda: P.lookup(blk.corr_pyramid, flow, W2=W), reps=20)
print(f"K2  per-level lookup {B*H*W} px: best {best*1e3:.1f} us median {med*1e3:.1f} us -> "
      f"{k2_bytes/med/1e6:.0f} GB/s algorithmic ({k2_bytes/med/1e6/peak:.2f} of {peak:.0f})")
best, med = graph_time(lambda: blk(flow), reps=20)
print(f"K2b graph replay: best {best*1e3:.1f} us median {med*1e3:.1f} us -> {k2_bytes/med/1e6:.0f} GB/s algorithmic "
      f"({k2_bytes/med/1e6/peak:.2f} of {peak:.0f})")
best, med = timeit(lambda: blk(flow), reps=20)
print(f"K2b lookup from level 0 {B*H*W} px: best {best*1e3:.1f} us median {med*1e3:.1f} us -> "
      f"{k2_bytes/med/1e6:.0f} GB/s algorithmic ({k2_bytes/med/1e6/peak:.2f} of {peak:.0f})")
