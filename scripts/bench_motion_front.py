"""GPU aid: K6 (fused lookup + PCBlock front) vs the reference's lookup + gelu(convc1(.)) on torch/cuDNN, 720p 1/4 res."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import ppmstereo_b200 as P
from oracle import ref_import
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
pc = upd.PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]).cuda().eval()
B, D, H, W = 20, 256, 184, 320
g = torch.Generator(device="cuda").manual_seed(1)
f1 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
f2 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
flow = torch.zeros(B, 2, H, W, device="cuda"); flow[:, 0] = torch.rand(B, H, W, generator=g, device="cuda") * -40
blk = P.CorrBlock1D(f1, f2)
front = P.MotionCorrFront(pc)

def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): r = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r

ms_k, got = t(lambda: front(P.LookupRequest(blk, flow)))
with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    ms_r, want = t(lambda: F.gelu(pc(blk(flow))))
print(f"K6 fused lookup + front: {ms_k:.3f} ms | K2b lookup + cuDNN PCBlock + gelu: {ms_r:.3f} ms | {ms_r / ms_k:.2f}x | identical "
      f"{float((got == want).float().mean()) * 100:.3f} %")
