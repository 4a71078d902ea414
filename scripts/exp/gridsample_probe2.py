"""GPU aid: discriminate cuDNN's lerp form with sample positions near 0 (fine mantissa bits in the weights)."""
import torch, torch.nn.functional as F
dev = "cuda"
N, W = 400000, 320
g = torch.Generator(device=dev).manual_seed(1)
img = torch.randn(N, 1, 1, W, device=dev, generator=g).bfloat16().float()
x = torch.rand(N, device=dev, generator=g) * 5 - 1.5
xg = 2 * x / (W - 1) - 1
grid = torch.stack([xg, torch.zeros_like(xg)], -1).view(N, 1, 1, 2)
ref = F.grid_sample(img, grid, align_corners=True).view(N)
rows = torch.arange(N, device=dev)
def gather(i):
    ok = (i >= 0) & (i < W)
    return torch.where(ok, img.view(N, W)[rows, i.clamp(0, W - 1)], torch.zeros((), device=dev))
ix = ((xg + 1) / 2) * (W - 1)
i0 = torch.floor(ix)
nw = (i0 + 1) - ix
ne = 1 - nw
ne_b = ix - i0
v0, v1 = gather(i0.long()), gather(i0.long() + 1)
fma = lambda a, b, c: torch.addcmul(c, a, b)
forms = {
    "A fma(v0,nw,v1*(1-nw))": fma(v0, nw, v1 * ne),
    "A' fma(v0,nw,v1*(ix-i0))": fma(v0, nw, v1 * ne_b),
    "B fma(v0,nw,fma(-nw,v1,v1))": fma(v0, nw, fma(-nw, v1, v1)),
    "C fma(nw,v0-v1,v1)": fma(nw, v0 - v1, v1),
    "D v1+nw*(v0-v1)": v1 + nw * (v0 - v1),
    "E fma(ne,v1-v0,v0)": fma(ne, v1 - v0, v0),
    "F fma(v1,ne_b,v0*nw)": fma(v1, ne_b, v0 * nw),
    "G fma(v1,ne_b,fma(-ne_b,v0,v0))": fma(v1, ne_b, fma(-ne_b, v0, v0)),
    "H fma(ne_b,v1-v0,v0)": fma(ne_b, v1 - v0, v0),
    "I exact64": (v0.double() * nw.double() + v1.double() * (1 - nw.double())).float(),
    "J exact64 w/ ix64": None,
}
x64 = ((xg.double() + 1) / 2) * (W - 1)
i64 = torch.floor(x64)
forms["J exact64 w/ ix64"] = (gather(i64.long()).double() * (i64 + 1 - x64) + gather(i64.long() + 1).double() * (x64 - i64)).float()
for name, o in forms.items():
    for lo, hi in ((-1.5, 0.0), (0.0, 1.0), (1.0, 3.5)):
        m = (x >= lo) & (x < hi)
        print(f"{name:34s} x in [{lo},{hi}): identical {float((o == ref)[m].float().mean()):.5f}", end=" | ")
    print()
print("---- left border: recover cuDNN's weight of element 0 for ix in (-1, 0)")
imp = torch.zeros(N, 1, 1, W, device=dev); imp[..., 0] = 1.0
w0 = F.grid_sample(imp, grid, align_corners=True).view(N)
m = (ix > -1) & (ix < 0)
u = (xg + 1) / 2
cands = {
    "1 - (0 - ix)": 1 - (0 - ix), "ix + 1": ix + 1,
    "fma(u, W-1, 1)": torch.addcmul(torch.ones_like(u), u, torch.full_like(u, W - 1)),
    "fma(xg+1, (W-1)/2, 1)": torch.addcmul(torch.ones_like(u), xg + 1, torch.full_like(u, (W - 1) / 2)),
    "fma(xg, h, h+1)": torch.addcmul(torch.full_like(u, (W - 1) / 2 + 1), xg, torch.full_like(u, (W - 1) / 2)),
    "float(ix64 + 1)": (ix.double() + 1).float(),
    "float(x64 + 1)": (x64 + 1).float(),
    "frac via ix - trunc: ix-(-1)": ix - torch.floor(ix),
}
for nm, c in cands.items():
    print(f"{nm:30s} identical {float((c == w0)[m].float().mean()):.5f}  max diff {float((c - w0)[m].abs().max()):.2e}")
d = (w0 - (ix + 1))[m]
print("sign of (cudnn - (ix+1)): +", int((d > 0).sum()), " -", int((d < 0).sum()), " 0", int((d == 0).sum()))
