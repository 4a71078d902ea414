"""GPU aid: fp32 (TF32) feature maps through the 2-CTA cluster path of K1 (W2 = 320), every row against torch."""
import os, sys, torch
sys.path.insert(0, os.getcwd())
import ppmstereo_b200 as P
g = torch.Generator(device="cuda").manual_seed(1)
for shape in ((1, 256, 3, 320, 320), (2, 128, 40, 320, 288)):
    B, D, H, W1, W2 = shape
    f1 = torch.randn(B, D, H, W1, generator=g, device="cuda"); f2 = torch.randn(B, D, H, W2, generator=g, device="cuda")
    for od in (torch.float32, torch.bfloat16):
        lv = P.build_pyramid(f1, f2, 4, out_dtype=od)
        want = torch.einsum("bdhi,bdhj->bhij", f1.double(), f2.double()).reshape(B * H * W1, W2) / D ** 0.5
        got = lv[0].reshape(B * H * W1, W2).double()
        rel = [float((got[:, a:b] - want[:, a:b]).norm() / want[:, a:b].norm()) for a, b in ((0, W2 // 2), (W2 // 2, W2))]
        print(shape, od, "rel err per N half", rel, "finite", bool(torch.isfinite(got).all()))
