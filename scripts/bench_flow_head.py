"""GPU aid: K5b (ppm_flow_head3d) against the reference FlowHead3D on cuDNN.  python scripts/bench_flow_head.py [T H W]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P
from oracle import ref_import
T, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (20, 184, 320)
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
head = upd.FlowHead3D(128, hidden_dim=256).cuda().eval()
x = torch.tanh(torch.randn(1, 128, T, H, W, generator=torch.Generator(device="cuda").manual_seed(1), device="cuda"))
k = P.FlowHead3D(head)

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): out = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out

ms_k, got = timeit(lambda: k(x))
with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    ms_r, want = timeit(lambda: head(x))
same = float((got == want).float().mean()) * 100
print(f"shape T={T} {H}x{W}: K5b {ms_k:.3f} ms | cuDNN module {ms_r:.3f} ms | {ms_r / ms_k:.2f}x | identical {same:.2f} % | "
      f"max diff {float((got.float() - want.float()).abs().max()):.4f}")
