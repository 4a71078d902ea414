"""GPU aid: time K4 (convex upsampling) at the 720p 1/4-res shape against (a) its HBM roofline and (b) the
reference's own op sequence in PyTorch on the same GPU (restated here: view/softmax/unfold/mul/sum/permute)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import ppmstereo_b200 as P

N, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (20, 184, 320)
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
flow = torch.randn(N, 2, H, W, generator=g, device=dev) * 8
peak = 6549.4
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def torch_sequence(flow, mask, rate=4):          # the op sequence of models/core/ppmstereo.py:185-197
    n, _, h, w = flow.shape
    m = torch.softmax(mask.view(n, 1, 9, rate, rate, h, w).contiguous(), dim=2)
    up = F.unfold(rate * flow, [3, 3], padding=1).view(n, 2, 9, 1, 1, h, w)
    up = torch.sum(m * up, dim=2).permute(0, 1, 4, 2, 5, 3).contiguous()
    return up.reshape(n, 2, rate * h, rate * w)


def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn(); ev[i + 1].record()
    torch.cuda.synchronize()
    return sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))[reps // 2]


for dt in (torch.bfloat16, torch.float32):
    mask = (torch.randn(N, 144, H, W, generator=g, device=dev) * 2).to(dt)
    byts = N * H * W * (144.0 * mask.element_size() + 8 + 128)
    ms = timeit(lambda: P.convex_upsample(flow, mask))
    ms_t = timeit(lambda: torch_sequence(flow, mask))
    d = (P.convex_upsample(flow, mask) - torch_sequence(flow, mask).float()).abs().max()
    print(f"K4 convex_upsample N={N} {H}x{W} mask {str(dt)[6:]}: {ms*1e3:.1f} us -> {byts/ms/1e6:.0f} GB/s algorithmic "
          f"({byts/ms/1e6/peak:.2f} of {peak:.0f}); PyTorch op sequence on the same GPU {ms_t*1e3:.1f} us "
          f"({ms_t/ms:.1f}x); max |diff| {float(d):.2e}")
