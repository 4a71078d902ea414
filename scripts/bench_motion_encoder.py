"""GPU aid: K7 (BasicMotionEncoder_v2.forward as one call) vs the reference module on torch/cuDNN and vs the K6-only
replacement (fused front, the three 3x3 convolutions and the cat on cuDNN), 720p 1/4 res x 20 frames by default.
  python scripts/bench_motion_encoder.py [B H W] [--kernel-only]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ppmstereo_b200 as P
from ppmstereo_b200 import motion
from oracle import ref_import
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
enc = upd.BasicMotionEncoder_v2(36).cuda().eval()
args = [a for a in sys.argv[1:] if not a.startswith("--")]
B, H, W = (int(x) for x in args[:3]) if len(args) >= 3 else (20, 184, 320)
D = 256
g = torch.Generator(device="cuda").manual_seed(1)
f1 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
f2 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
flow = torch.zeros(B, 2, H, W, device="cuda"); flow[:, 0] = torch.rand(B, H, W, generator=g, device="cuda") * -40
mhs = torch.randn(B, 64, H, W, generator=g, device="cuda").bfloat16()
blk = P.CorrBlock1D(f1, f2)
impl = P.MotionEncoder(enc)

def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): r = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r

with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    ms_k, got = t(lambda: impl(flow, P.LookupRequest(blk, flow), mhs, None))
    if "--kernel-only" in sys.argv:
        print(f"K7 encoder: {ms_k:.3f} ms")
        sys.exit(0)
    ms_r, want = t(lambda: enc(flow, blk(flow), mhs, None))
    front = impl.front
    def k6_only():
        cor = front(P.LookupRequest(blk, flow))
        cor = torch.relu(enc.convc2(cor))
        flo = torch.relu(enc.convf2(torch.relu(enc.convf1(flow))))
        out = torch.relu(enc.final_conv(torch.cat([cor, flo, mhs], dim=1)))
        o, h = torch.split(out, [126, 64], dim=1)
        return torch.cat([o, flow], dim=1), h
    ms_6, _ = t(k6_only)
same_f = float((got[0] == want[0]).float().mean()) * 100
same_h = float((got[1] == want[1]).float().mean()) * 100
P_ = B * H * W
flop = 2.0 * P_ * 9 * (256 * 192 + 128 * 64 + 320 * 190)
print(f"K7 encoder: {ms_k:.3f} ms | K6 front + cuDNN convs + cat: {ms_6:.3f} ms | reference (K2b lookup + torch/cuDNN): {ms_r:.3f} ms | "
      f"{ms_r / ms_k:.2f}x | identical: features {same_f:.2f} %, hidden {same_h:.2f} % | 3x3 convs {flop / 1e12:.2f} TFLOP")
