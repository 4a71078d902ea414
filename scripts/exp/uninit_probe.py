"""GPU aid: does any installed kernel read memory it (or its producers) never wrote?  The caching allocator is
poisoned before every forward (a large block filled with a byte pattern, then freed, so that torch.empty hands the
pattern back): 0xFF bytes are NaN in fp32 / bf16 and propagate into the result if they are ever read; two different
finite patterns must give bit-identical results.  Arms: reference, install(gru=False), install()."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import ppmstereo_b200 as P
from oracle import ref_import

model, ppm = ref_import.build_reference_model(seed=0, beta=0.5, mixed_precision=True)
model = model.cuda().eval()
T, H, W, iters = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (6, 64, 256, 6)
video = torch.rand(T, 2, 3, H, W, generator=torch.Generator().manual_seed(17)) * 255.0
left, right = video[None, :, 0].cuda(), video[None, :, 1].cuda()


def poison(byte, gb=6):
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    blocks = [torch.full((1 << 28,), byte, dtype=torch.uint8, device="cuda") for _ in range(gb * 4)]
    small = [torch.full((n,), byte, dtype=torch.uint8, device="cuda") for n in (1 << 12, 1 << 16, 1 << 20, 1 << 24) for _ in range(64)]
    torch.cuda.synchronize()
    del blocks, small


def fwd():
    with torch.no_grad():
        d, u = model(left, right, iters=iters, test_mode=True)
    torch.cuda.synchronize()
    return d.float(), u.float()


for name, setup in (("reference", lambda: P.uninstall(ppm)), ("install(gru=False)", lambda: P.install(ppm, gru=False)),
                    ("install()", lambda: P.install(ppm)), ("install(graph=always)", lambda: P.install(ppm, graph="always"))):
    setup()
    fwd()                                                   # warm (lazy workspaces allocated here keep their blocks)
    res = {}
    for byte in (0xFF, 0x5A, 0x00):
        poison(byte)
        res[byte] = fwd()
    nan = bool(torch.isnan(res[0xFF][0]).any() or torch.isnan(res[0xFF][1]).any())
    same_5a = bool(torch.equal(res[0x5A][0], res[0x00][0]) and torch.equal(res[0x5A][1], res[0x00][1]))
    same_ff = bool(torch.equal(res[0xFF][0], res[0x00][0]))
    mad = float((res[0x5A][0] - res[0x00][0]).abs().mean())
    print(f"{name:24s} NaN with 0xFF poison: {nan} | 0x5A == 0x00: {same_5a} (mean |d| {mad:.2e}) | 0xFF == 0x00: {same_ff}", flush=True)
    P.uninstall(ppm)
