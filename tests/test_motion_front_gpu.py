"""K6: lookup fused into the motion encoder's front end -- gelu(PCBlock(corr)) -- against the
reference's block run by torch / cuDNN under bf16 autocast on the same GPU (staged sources), against the CPU oracle's
bf16-mode restatement, and fused-lookup vs materialised-lookup (must be bit-identical: same arithmetic, the taps just
stay in registers)."""
import importlib

import pytest
import torch
import torch.nn.functional as F

from oracle import ppm_oracle as O
from oracle import ref_import

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if not ref_import.available():
        pytest.skip("reference sources not staged")
    import ppmstereo_b200 as P
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(5)
    pc = upd.PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]).cuda().eval()
    return P, pc


@pytest.mark.parametrize("shape", [(1, 5, 7), (2, 16, 32), (3, 23, 40), (2, 46, 80), (20, 92, 160)])
def test_vs_reference_block_under_autocast(env, shape):
    P, pc = env
    B, H, W = shape
    corr = torch.randn(B, 36, H, W, generator=torch.Generator(device="cuda").manual_seed(B + H), device="cuda") * 2
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        want = F.gelu(pc(corr))
    assert want.dtype == torch.bfloat16
    got = P.MotionCorrFront(pc)(corr)
    assert got.shape == want.shape == (B, 256, H, W) and got.dtype == torch.bfloat16
    d = (got.float() - want.float()).abs()
    same = float((got == want).float().mean())
    scale = float(want.float().abs().max())
    print(f"K6 {shape}: {same * 100:.3f} % bit-identical to torch/cuDNN, max |diff| {float(d.max()):.4f} of {scale:.2f}, mean {float(d.mean()):.2e}")
    assert float(d.max()) <= 2.0 ** -6 * max(scale, 1.0)
    assert float(d.mean()) <= 1e-4 * max(scale, 1.0)
    assert same >= 0.98


def test_vs_cpu_oracle_bf16_mode(env):
    P, pc = env
    corr = torch.randn(2, 36, 9, 13, generator=torch.Generator(device="cuda").manual_seed(1), device="cuda") * 2
    params = {k: v.detach().float().cpu() for k, v in pc.state_dict().items()}
    want = O.motion_corr_front(corr.cpu(), params, mode="bf16")
    got = P.MotionCorrFront(pc)(corr).float().cpu()
    d = (got - want).abs()
    assert float(d.max()) <= 2.0 ** -6 * max(1.0, float(want.abs().max()))
    assert float((d == 0).float().mean()) >= 0.98


@pytest.mark.parametrize("shape", [(2, 256, 6, 64), (3, 256, 23, 40), (1, 64, 16, 328)])
def test_fused_lookup_equals_materialised_lookup(env, shape):
    """LookupRequest (level-0 pyramid + flow, taps in registers) vs the (B,36,h,w) tensor CorrBlock1D.__call__ writes."""
    P, pc = env
    B, D, H, W = shape
    g = torch.Generator(device="cuda").manual_seed(W)
    f1 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    flow = torch.zeros(B, 2, H, W, device="cuda")
    flow[:, 0] = torch.rand(B, H, W, generator=g, device="cuda") * -(W * 0.6) + 8.0
    blk = P.CorrBlock1D(f1, f2)
    assert blk._l0_only
    front = P.MotionCorrFront(pc)
    a = front(P.LookupRequest(blk, flow))
    b_ = front(blk(flow))
    assert torch.equal(a, b_)


def test_no_write_outside_output(env):
    import ctypes  # noqa: F401
    from ppmstereo_b200 import _lib
    P, pc = env
    lib = _lib.load()
    B, H, W = 2, 19, 37                                     # partial 32 x 16 tiles in both directions
    corr = torch.randn(B, 36, H, W, generator=torch.Generator(device="cuda").manual_seed(2), device="cuda")
    front = P.MotionCorrFront(pc)
    want = front(corr)
    wa, wb = front._params()
    n = B * 256 * H * W
    buf = torch.full((n + 2048,), -3.0, dtype=torch.bfloat16, device="cuda")
    ws = torch.full((B * 36 * H * W + 2048,), float("nan"), device="cuda")
    _lib.check(lib.ppm_motion_corr_front(None, None, corr.data_ptr(), wa.data_ptr(), wb.data_ptr(), B, H, W, 0,
                                         ws.data_ptr() + 1024 * 4, buf.data_ptr() + 1024 * 2, _lib.stream_ptr()), "k6")
    torch.cuda.synchronize()
    assert bool((buf[:1024] == -3.0).all()) and bool((buf[1024 + n:] == -3.0).all())
    assert bool(torch.isnan(ws[:1024]).all()) and bool(torch.isnan(ws[1024 + B * 36 * H * W:]).all())
    assert torch.equal(buf[1024:1024 + n].view(B, 256, H, W), want)
