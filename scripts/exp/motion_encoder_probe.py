"""GPU aid: where does BasicMotionEncoder_v2 (+ to_v, uncertainty, mask heads) spend its time at 720p 1/4 res x 20 frames?"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn.functional as F
from oracle import ref_import
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
blk = upd.SequenceUpdateBlock3D(hidden_dim=128, cor_planes=36, mask_size=4).cuda().eval()
enc = blk.encoder
B, h, w = 20, 184, 320
g = torch.Generator(device="cuda").manual_seed(0)
corr = torch.randn(B, 36, h, w, generator=g, device="cuda")
flow = torch.randn(B, 2, h, w, generator=g, device="cuda")
inp = torch.randn(B, 128, h, w, generator=g, device="cuda").bfloat16()
mhs = torch.randn(B, 64, h, w, generator=g, device="cuda").bfloat16()
net = torch.randn(B, 128, h, w, generator=g, device="cuda")

def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): r = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r

with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    ms, _ = t(lambda: enc(flow, corr, mhs, inp)); print(f"encoder total            {ms:7.3f} ms")
    ms, c1 = t(lambda: F.gelu(enc.convc1(corr))); print(f"  convc1 PCBlock + gelu  {ms:7.3f} ms  -> {tuple(c1.shape)} {c1.dtype}")
    pc = enc.convc1
    ms, x1 = t(lambda: F.gelu(corr + pc.ffn1(corr))); print(f"    ffn1 + res + gelu    {ms:7.3f} ms {x1.dtype}")
    x = x1
    for i, conv in enumerate(pc.conv_list):
        ms, x = t(lambda: F.gelu(x + conv(x))); print(f"    depthwise k={conv.kernel_size[0]} + gelu  {ms:7.3f} ms")
    ms, x2 = t(lambda: F.gelu(x + pc.pw(x))); print(f"    pw + gelu            {ms:7.3f} ms")
    ms, x3 = t(lambda: pc.ffn2(x2)); print(f"    ffn2                 {ms:7.3f} ms")
    ms, c2 = t(lambda: F.relu(enc.convc2(c1))); print(f"  convc2 3x3 256->192    {ms:7.3f} ms")
    ms, f1 = t(lambda: F.relu(enc.convf1(flow))); print(f"  convf1 7x7 2->128      {ms:7.3f} ms")
    ms, f2 = t(lambda: F.relu(enc.convf2(f1))); print(f"  convf2 3x3 128->64     {ms:7.3f} ms")
    cat = torch.cat([c2, f2, mhs], 1)
    ms, _ = t(lambda: torch.cat([c2, f2, mhs], 1)); print(f"  cat 320                {ms:7.3f} ms")
    ms, fo = t(lambda: F.relu(enc.final_conv(cat))); print(f"  final 3x3 320->190     {ms:7.3f} ms")
    mf = torch.cat([fo[:, :126], flow], 1)
    ms, v = t(lambda: blk.aggregator.to_v(mf)); print(f"to_v 1x1                 {ms:7.3f} ms  {v.dtype}")
    ms, _ = t(lambda: blk.get_uncertainty(torch.cat([net, v], 1))); print(f"uncertainty head         {ms:7.3f} ms")
    ms, _ = t(lambda: 0.25 * blk.mask_2d(net)); print(f"mask head                {ms:7.3f} ms")
