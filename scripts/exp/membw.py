"""GPU aid: write-only / read-only / copy HBM bandwidth with torch kernels (context for K1's store side)."""
import torch
dev = torch.device("cuda")
n = 1 << 30                                   # 2 GB of bf16
x = torch.empty(n, dtype=torch.bfloat16, device=dev)
y = torch.empty(n, dtype=torch.bfloat16, device=dev)
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
ms = t(lambda: x.zero_());            print(f"memset 2 GB: {ms:.3f} ms -> {2*n/ms/1e6:.0f} GB/s written")
ms = t(lambda: x.fill_(1.5));         print(f"fill   2 GB: {ms:.3f} ms -> {2*n/ms/1e6:.0f} GB/s written")
ms = t(lambda: y.copy_(x));           print(f"copy   2 GB: {ms:.3f} ms -> {4*n/ms/1e6:.0f} GB/s read+written")
ms = t(lambda: x.sum());              print(f"sum    2 GB: {ms:.3f} ms -> {2*n/ms/1e6:.0f} GB/s read")
