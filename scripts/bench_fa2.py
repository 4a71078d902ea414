"""Same-box comparison: the reference's attention call (flash_attn.flash_attn_func, models/core/ppmstereo.py:550)
on the installed flash_attn, next to ppm_memory_attend, at one shape.  Usage: bench_fa2.py [T h w]."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import flash_attn
import ppmstereo_b200 as P

T, h, w = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (20, 184, 320)
hw, c, K = h * w, 128, 5
g = torch.Generator(device="cuda").manual_seed(2)
q = torch.randn(1, hw, 1, c, device="cuda", generator=g).bfloat16()
k = torch.randn(1, K * hw, 1, c, device="cuda", generator=g).bfloat16()
v = torch.randn(1, K * hw, 1, c, device="cuda", generator=g).bfloat16()
scale = P.softmax_scale(c)
flops = 4.0 * K * c * hw * hw


def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms = timeit(lambda: flash_attn.flash_attn_func(q, k, v, 0.0, softmax_scale=scale, causal=False), 5)
print(f"flash_attn {flash_attn.__version__} flash_attn_func  Lq={hw} Lk={K*hw} d=128: {ms:.2f} ms/query frame -> {flops/ms/1e9:.0f} TFLOP/s")

qq = torch.randn(1, c, T, h, w, device="cuda", generator=g)
kk = torch.randn(1, c, T, h, w, device="cuda", generator=g)
vv = torch.randn(1, c, T, h, w, device="cuda", generator=g)
mf = torch.randn(1, c, T, h, w, device="cuda", generator=g)
mem = P.PPMMemory(qq, kk, None, top_k=K)
mem.pick(torch.rand(1, 1, T, T, device="cuda"), torch.ones(1, 1, T, T, device="cuda"), torch.rand(T, 1, h, w, device="cuda"))
mem.pack_kv(vv)
beta = torch.tensor([0.5], device="cuda")
out = torch.empty(T, c, h, w, device="cuda")
ms2 = timeit(lambda: mem.aggregate(mf, beta, scale, out=out), 3)
print(f"ppm_memory_attend (T={T} query frames per launch): {ms2/T:.2f} ms/query frame -> {flops*T/ms2/1e9:.0f} TFLOP/s  "
      f"({ms*T/ms2:.2f}x flash_attn)")

# numerics at the same shape: query frame 0 of the launch above against flash_attn on the same packed operands
hw_pad = (hw + 7) // 8 * 8
q0 = mem.q_packed.reshape(1, T, hw, c)[0, 0].reshape(1, hw, 1, c)
k0 = mem.k_packed.reshape(1, T, K, hw, c)[0, 0].reshape(1, K * hw, 1, c)
v0 = mem.v_packed.reshape(1, T, c, hw_pad)[0, mem.picked.long()[0, 0], :, :hw].permute(0, 2, 1).reshape(1, K * hw, 1, c).contiguous()
ref = flash_attn.flash_attn_func(q0, k0, v0, 0.0, softmax_scale=scale, causal=False).float().reshape(hw, c).t()
mine = (out[0].reshape(c, hw) - mf[0, :, 0].reshape(c, hw)) / 0.5
d = (mine - ref).abs()
print(f"hidden vs flash_attn on identical bf16 operands: max |diff| {float(d.max()):.3e}, mean {float(d.mean()):.3e}, "
      f"max |hidden| {float(ref.abs().max()):.3e}, bit-identical after bf16 rounding {float((mine.bfloat16() == ref.bfloat16()).float().mean()):.4f}")
