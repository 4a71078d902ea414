import sys, os, torch
sys.path.insert(0, os.getcwd())
import ppmstereo_b200 as P
g = torch.Generator(device="cuda").manual_seed(3)
f1 = torch.randn(20, 256, 184, 320, generator=g, device="cuda").bfloat16()
f2 = torch.randn(20, 256, 184, 320, generator=g, device="cuda").bfloat16()
for dt in (torch.float32, torch.bfloat16):
    bad_runs = 0
    for rep in range(5):
        lv = P.build_pyramid(f1, f2, 4, out_dtype=dt)
        torch.cuda.synchronize()
        L0 = lv[0].reshape(20 * 184, 320, 320)
        nonfinite = (~torch.isfinite(L0.float())).sum().item()
        big = (L0.float().abs() > 100).sum().item()
        rows = []
        if nonfinite or big:
            bad = ((~torch.isfinite(L0.float())) | (L0.float().abs() > 100)).nonzero()
            rows = sorted(set((int(r) , int(m) // 128, int(n) // 160) for r, m, n in bad[:200000:997].tolist()))[:12]
            bad_runs += 1
        print(dt, "rep", rep, "nonfinite", nonfinite, "big", big, rows)
    # reference check on a few rows
    r = 1234; b, h = divmod(r, 184)
    want = (f1[b, :, h].float().t() @ f2[b, :, h].float()) / 16.0
    print("row check", float((L0[r].float() - want).abs().max()))
