"""GPU aid: installed model on a small clip (BASELINE configs[0] size), eager iterations vs CUDA-graph replay.
  python scripts/small_clip_probe.py [T H W iters]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P
from oracle import ref_import

T, H, W, iters = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (5, 320, 512, 10)
model, ppm = ref_import.build_reference_model(seed=0, beta=0.5, mixed_precision=True)
model = model.cuda().eval()
video = torch.rand(T, 2, 3, H, W, generator=torch.Generator().manual_seed(1)) * 255.0
l, r = video[None, :, 0].cuda(), video[None, :, 1].cuda()


def run(n=5):
    with torch.no_grad():
        model(l, r, iters=iters, test_mode=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            d, _ = model(l, r, iters=iters, test_mode=True)
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, d


ms_ref, d_ref = run()
out = {}
for mode in ("never", "auto", "always", "never", "auto", "always"):        # alternated: shows run-to-run variance
    P.install(ppm, graph=mode)
    ms, d = run()
    print(f"  {mode:7s} {ms:7.1f} ms/forward", flush=True)
    if mode not in out or ms < out[mode][0]:
        out[mode] = (ms, d)
P.uninstall(ppm)
print(f"T={T} {H}x{W} iters={iters}: reference {ms_ref:.1f} ms/forward; installed eager {out['never'][0]:.1f}, graph auto "
      f"{out['auto'][0]:.1f}, graph always {out['always'][0]:.1f} ms/forward ({T / out['auto'][0] * 1e3:.1f} frames/s); "
      f"graph == eager: {bool(torch.equal(out['never'][1], out['always'][1]))}; vs reference mean |d| "
      f"{float((out['auto'][1].float() - d_ref.float()).abs().mean()):.4f} px")
