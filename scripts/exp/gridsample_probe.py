"""GPU aid: which fp32 op sequence reproduces torch's grid_sample(align_corners=True) on a height-1 image bit for bit?"""
import torch, torch.nn.functional as F
torch.manual_seed(0)
dev = "cuda"
N, W = 200000, 320
img = torch.randn(N, 1, 1, W, device=dev).bfloat16().float()
x = (torch.rand(N, device=dev) * (W + 20) - 10)            # sample position in pixels
for variant in ("recip", "div"):
    xg = (2 * x * (1.0 / (W - 1)) - 1) if variant == "recip" else (2 * x / (W - 1) - 1)
    # what torch itself does for `2 * x / (W - 1) - 1`
    xg_t = 2 * x / (W - 1) - 1
    print(variant, "xgrid equals torch expr:", float((xg == xg_t).float().mean()))
xg = 2 * x / (W - 1) - 1
grid = torch.stack([xg, torch.zeros_like(xg)], -1).view(N, 1, 1, 2)
ref = F.grid_sample(img, grid, align_corners=True).view(N)
rows = torch.arange(N, device=dev)
def gather(i):
    ok = (i >= 0) & (i < W)
    return torch.where(ok, img.view(N, W)[rows, i.clamp(0, W - 1)], torch.zeros((), device=dev))
cands = {
    "((g+1)/2)*(W-1)": ((xg + 1) / 2) * (W - 1),
    "((g+1)*0.5f)*(W-1)": ((xg + 1) * 0.5) * (W - 1),
    "(g+1)*((W-1)/2)": (xg + 1) * ((W - 1) / 2),
    "fma(g,(W-1)/2,(W-1)/2)": torch.addcmul(torch.full_like(xg, (W - 1) / 2), xg, torch.full_like(xg, (W - 1) / 2)),
}
for name, ix in cands.items():
    i0 = torch.floor(ix)
    nw, ne = (i0 + 1) - ix, ix - i0
    v0, v1 = gather(i0.long()), gather(i0.long() + 1)
    outs = {
        "fma(v1,ne,v0*nw)": torch.addcmul(v0 * nw, v1, ne),
        "v0*nw+v1*ne": v0 * nw + v1 * ne,
        "fma64": (v1.double() * ne.double() + (v0 * nw).double()).float(),
        "exact64": (v0.double() * nw.double() + v1.double() * ne.double()).float(),
    }
    for on, o in outs.items():
        print(f"{name:26s} {on:18s} identical {float((o == ref).float().mean()):.5f}  max {float((o - ref).abs().max()):.2e}")

print("---- more lerp forms (cudnn on), and the ATen kernel (cudnn off)")
ix = ((xg + 1) / 2) * (W - 1)
i0 = torch.floor(ix)
nw, ne = (i0 + 1) - ix, ix - i0
v0, v1 = gather(i0.long()), gather(i0.long() + 1)
one_m = 1 - ne
forms = {
    "v0+ne*(v1-v0)": v0 + ne * (v1 - v0),
    "fma(ne,v1-v0,v0)": torch.addcmul(v0, ne, v1 - v0),
    "fma(v0,nw,v1*ne)": torch.addcmul(v1 * ne, v0, nw),
    "(1-ne)*v0+ne*v1": one_m * v0 + ne * v1,
    "fma(v1,ne,(1-ne)*v0)": torch.addcmul(one_m * v0, v1, ne),
    "fma(v0,1-ne,v1*ne)": torch.addcmul(v1 * ne, v0, one_m),
    "v0*nw+v1*ne": v0 * nw + v1 * ne,
}
with torch.backends.cudnn.flags(enabled=False):
    ref_aten = F.grid_sample(img, grid, align_corners=True).view(N)
print("cudnn vs aten identical:", float((ref == ref_aten).float().mean()))
for on, o in forms.items():
    print(f"{on:24s} vs cudnn {float((o == ref).float().mean()):.5f}   vs aten {float((o == ref_aten).float().mean()):.5f}")
print("aten: fma(v1,ne,v0*nw)", float((torch.addcmul(v0 * nw, v1, ne) == ref_aten).float().mean()))
# does the position differ?  look at samples where floor differs: compare weights recovered from an impulse image
imp = torch.zeros(N, 1, 1, W, device=dev)
j = i0.long().clamp(0, W - 2)
imp.view(N, W)[rows, j + 1] = 1.0
w_ne = F.grid_sample(imp, grid, align_corners=True).view(N)          # = weight of element i0+1 as cudnn computes it
inside = (i0 >= 0) & (i0 < W - 1)
print("cudnn ne weight == ix - i0:", float((w_ne == ne)[inside].float().mean()), " max diff", float((w_ne - ne)[inside].abs().max()))
imp.zero_(); imp.view(N, W)[rows, j] = 1.0
w_nw = F.grid_sample(imp, grid, align_corners=True).view(N)
print("cudnn nw weight == i0+1-ix:", float((w_nw == nw)[inside].float().mean()), " == 1-ne:", float((w_nw == one_m)[inside].float().mean()))
print("---- cudnn candidates")
ne2 = 1 - nw
for nm, nee in (("ne=ix-i0", ne), ("ne=1-nw", ne2)):
    print(nm, "weight match", float((w_ne == nee)[inside].float().mean()))
    for on, o in {"fma(v0,nw,v1*ne)": torch.addcmul(v1 * nee, v0, nw), "fma(v1,ne,v0*nw)": torch.addcmul(v0 * nw, v1, nee),
                  "v0*nw+v1*ne": v0 * nw + v1 * nee}.items():
        print(f"   {on:20s} vs cudnn {float((o == ref).float().mean()):.5f} max {float((o - ref).abs().max()):.2e}")
# maybe cudnn's coordinate differs slightly: recover ix from weights: ix_c = i0 + 1 - w_nw (exact when inside)
for nm, cand in {"(g+1)*(W-1)/2 [mul then /2]": ((xg + 1) * (W - 1)) / 2, "fma(g,h,h)": torch.addcmul(torch.full_like(xg, (W - 1) / 2), xg, torch.full_like(xg, (W - 1) / 2)),
                 "0.5*((g+1)*(W-1))": 0.5 * ((xg + 1) * (W - 1)), "((g+1)/2)*(W-1)": ix}.items():
    i0c = torch.floor(cand)
    print(f"{nm:30s} floor equal {float((i0c == i0).float().mean()):.6f}  nw equal {float(((i0c + 1 - cand) == w_nw)[inside].float().mean()):.6f}  ne equal {float(((cand - i0c) == w_ne)[inside].float().mean()):.6f}")
print("---- where does fma(v0,nw,v1*(1-nw)) differ from cudnn?")
o = torch.addcmul(v1 * ne2, v0, nw)
bad = o != ref
print("mismatches", int(bad.sum()), "of", N, " border(i0<0 or i0>=W-1):", int((bad & ~inside).sum()))
idx = torch.nonzero(bad & inside).flatten()[:8]
for i in idx.tolist():
    print(f"ix {float(ix[i]):.6f} nw {float(nw[i]):.8f} ne {float(ne2[i]):.8f} v0 {float(v0[i]):.6f} v1 {float(v1[i]):.6f} mine {float(o[i]):.9f} cudnn {float(ref[i]):.9f} "
          f"alt fma(v1,ne,v0*nw) {float(torch.addcmul(v0 * nw, v1, ne2)[i]):.9f} plain {float((v0 * nw + v1 * ne2)[i]):.9f} exact {float(v0[i].double() * nw[i].double() + v1[i].double() * ne2[i].double()):.12f}")
idx = torch.nonzero(bad).flatten()[:10]
for i in idx.tolist():
    print(f"ix {float(ix[i]):.6f} i0 {int(i0[i])} nw {float(nw[i]):.8f} ne {float(ne2[i]):.8f} v0 {float(v0[i]):.6f} v1 {float(v1[i]):.6f} mine {float(o[i]):.9f} cudnn {float(ref[i]):.9f}")
nb = ~inside
print("border samples", int(nb.sum()), "left", int((i0 == -1).sum()), "right", int((i0 == W - 1).sum()), "bad left", int((bad & (i0 == -1)).sum()), "bad right", int((bad & (i0 == W - 1)).sum()),
      "bad beyond", int((bad & ((i0 < -1) | (i0 > W - 1))).sum()))
