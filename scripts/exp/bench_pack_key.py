"""GPU aid: ppm_pack_key at the headline shape (1/4 res, T = 20, K = 5): ms per call and HBM fraction."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import ppmstereo_b200 as P
b, T, h, w = 1, 20, 184, 320
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(b, 128, T, h, w, generator=g, device="cuda").bfloat16()
k = torch.randn(b, 128, T, h, w, generator=g, device="cuda").bfloat16()
v = torch.randn(b * T, 128, h, w, generator=g, device="cuda").bfloat16()
mem = P.PPMMemory(q, k, None, top_k=5)
mem.pick(torch.rand(b, 1, T, T, generator=g, device="cuda"), torch.ones(b, 1, T, T, device="cuda"),
         torch.rand(b * T, 1, h, w, generator=g, device="cuda"))
val = v.reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4)
for _ in range(3): mem.pack_kv(val)
torch.cuda.synchronize()
a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20): mem.pack_kv(val)
e.record(); torch.cuda.synchronize()
ms = a.elapsed_time(e) / 20
hw = h * w
wr = T * 5 * hw * 128 * 2
rd = T * hw * 128 * 2
print(f"ppm_pack_key: {ms:.3f} ms per call; {wr / 1e9:.2f} GB written + {rd / 1e9:.2f} GB read (each key frame once) = "
      f"{(wr + rd) / ms / 1e6:.0f} GB/s")
