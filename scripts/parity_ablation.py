"""GPU aid: where does the installed-vs-reference disparity difference come from, and how large is the reference's own
rounding-noise floor?  (configs[0] shape by default.)
  python scripts/parity_ablation.py [T H W iters]
Arms (all against the unmodified reference, bf16 autocast, same weights / clip):
  floor_kv_flip   reference with the K/V sequence of every flash_attn call reversed (mathematically identical
                  softmax attention, different summation order inside FA2) = the reference's own rounding noise
  floor_fp32_sm   reference with flash_attn replaced by an fp32 softmax attention rounded to bf16 at the end
  corr_only       reference loop + our CorrBlock1D
  memory_only     our loop shell + memory read, reference CorrBlock1D and reference convex_upsample
  upsample_only   reference everything, convex_upsample replaced by K4
  installed       everything"""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import ref_import
import ppmstereo_b200 as P

a = [int(x) for x in sys.argv[1:5]] if len(sys.argv) >= 5 else [5, 320, 512, 10]
T, H, W, iters = a
model, ppm = ref_import.build_reference_model(seed=0, beta=0.5, mixed_precision=True)
model = model.cuda().eval()
inst = importlib.import_module("ppmstereo_b200.install")
video = torch.rand(T, 2, 3, H, W, generator=torch.Generator().manual_seed(11)) * 255.0
l, r = video[None, :, 0].cuda(), video[None, :, 1].cuda()


def run():
    with torch.no_grad():
        d, u = model(l, r, iters=iters, test_mode=True)
    torch.cuda.synchronize()
    return d.float()


ref = run()
out = {"shape": a, "disparity_mean_abs": float(ref.abs().mean())}
real_fa = ppm.flash_attn_func

ppm.flash_attn_func = lambda q, k, v, **kw: real_fa(q, k.flip(1).contiguous(), v.flip(1).contiguous(), **kw)
out["floor_kv_flip"] = float((run() - ref).abs().mean())


def fa_fp32(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False):
    qf, kf, vf = q[:, :, 0].float(), k[:, :, 0].float(), v[:, :, 0].float()
    o = torch.empty_like(qf)
    for s in range(0, qf.shape[1], 8192):
        p = torch.softmax(qf[:, s:s + 8192] @ kf.transpose(1, 2) * softmax_scale, dim=-1)
        o[:, s:s + 8192] = p @ vf
    return o.to(torch.bfloat16)[:, :, None]


torch.backends.cuda.matmul.allow_tf32 = False
ppm.flash_attn_func = fa_fp32
out["floor_fp32_sm"] = float((run() - ref).abs().mean())
ppm.flash_attn_func = real_fa

ref_corr = ppm.CorrBlock1D
ppm.CorrBlock1D = P.CorrBlock1D
out["corr_only"] = float((run() - ref).abs().mean())
ppm.CorrBlock1D = ref_corr

ref_up = ppm.PPMStereo.convex_upsample
ppm.PPMStereo.convex_upsample = lambda self, flow, mask, rate=4: P.convex_upsample(flow, mask, rate)
out["upsample_only"] = float((run() - ref).abs().mean())
ppm.PPMStereo.convex_upsample = ref_up

P.install(ppm)
ppm.CorrBlock1D = ref_corr
k4 = inst.convex_upsample
inst.convex_upsample = lambda flow, mask, rate=4: ref_up(None, flow, mask, rate)
out["memory_only"] = float((run() - ref).abs().mean())
inst.convex_upsample = k4
P.uninstall(ppm)

P.install(ppm)
out["installed"] = float((run() - ref).abs().mean())
P.uninstall(ppm)
print(json.dumps(out, indent=1))
