"""GPU aid: torch.profiler kernel table of ONE installed 20-frame 720p window (iters 20), attention excluded from the
listing so that the glue shows.  python scripts/exp/profile_window.py [frames H W iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import profile, ProfilerActivity
import ppmstereo_b200 as P
from oracle import full_model

frames, H, W, iters = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (20, 720, 1280, 20)
model, ppm = full_model.build()
video = full_model.synthetic_clip(frames, H, W)
P.install(ppm)
batch = {"stereo_video": video}
with torch.no_grad():
    model.forward_batch_test(batch, kernel_size=frames, iters=iters)          # warm
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model.forward_batch_test(batch, kernel_size=frames, iters=iters)
        torch.cuda.synchronize()
P.uninstall(ppm)
rows = [(e.key, e.device_time_total / 1e3, e.count) for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print(f"total CUDA kernel time {tot:.1f} ms over {sum(r[2] for r in rows)} launches")
for k, ms, n in rows[:45]:
    print(f"{ms:9.2f} ms {100 * ms / tot:6.2f} % {n:6d}  {k[:110]}")
