"""GPU aid: time K5 (ppm_sepconv_gru3d) against the reference module on cuDNN at a given shape.
  python scripts/bench_gru.py [T H W]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P
from oracle import ref_import

T, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (20, 184, 320)
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
gru = upd.SKSepConvGRU3D(hidden_dim=128, input_dim=384).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(1)
h = torch.tanh(torch.randn(1, 128, T, H, W, generator=g, device="cuda")).bfloat16()
x = torch.randn(1, 384, T, H, W, generator=g, device="cuda")
k = P.SepConvGRU3D(gru)
flop = 2.0 * T * H * W * (512 * 128 * (15 + 15 + 5) + 2 * 128 * 128 * 5 + 2 * 3 * 512 * 128 * 5)


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


ms_k, got = timeit(lambda: k(h, x))
with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    ms_r, want = timeit(lambda: gru(h, x))
d = (got.float() - want.float()).abs()
print(f"shape T={T} {H}x{W}: K5 {ms_k:.2f} ms = {flop / ms_k / 1e9:.0f} TFLOP/s | cuDNN module {ms_r:.2f} ms = {flop / ms_r / 1e9:.0f} TFLOP/s "
      f"| speedup {ms_r / ms_k:.2f}x | max diff {float(d.max()):.4f} mean {float(d.mean()):.2e} identical {float((got == want).float().mean()) * 100:.2f} %")
