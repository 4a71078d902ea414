"""K8: value projection + uncertainty head (ppm_value_uncertainty) against the reference's modules run by torch / cuDNN
under bf16 autocast on the same GPU, against the CPU oracle's bf16-mode restatement, and inside the installed loop shell.
Bars as for the other shifted-GEMM convolutions (K5b / K7): value is ONE K = 128 contraction, so it is compared
strictly (>= 99 % bit-identical, one bf16 ulp); the uncertainty map sits behind a 2304-term convolution whose one-ulp
flips pass through the sigmoid: >= 97 % identical, |diff| <= one bf16 ulp of 1 (2^-8)."""
import ctypes
import importlib

import pytest
import torch

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
    torch.manual_seed(21)
    blk = upd.SequenceUpdateBlock3D(hidden_dim=128, cor_planes=36, mask_size=4).cuda().eval()
    return P, blk, upd


def _inputs(B, H, W, seed, net_dtype=torch.float32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mf = torch.randn(B, 128, H, W, generator=g, device="cuda")
    net = torch.tanh(torch.randn(B, 128, H, W, generator=g, device="cuda")).to(net_dtype)
    return mf, net


def _reference(blk, mf, net):
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        value = blk.aggregator.to_v(mf)
        unc = blk.get_uncertainty(torch.cat([net, value], dim=1))
    return value, unc


@pytest.mark.parametrize("shape", [(1, 5, 7), (2, 16, 32), (3, 23, 40), (2, 46, 80), (20, 92, 160)])
@pytest.mark.parametrize("net_dtype", [torch.float32, torch.bfloat16])
def test_vs_reference_modules_under_autocast(env, shape, net_dtype):
    P, blk, _ = env
    mf, net = _inputs(*shape, seed=sum(shape), net_dtype=net_dtype)
    want_v, want_u = _reference(blk, mf, net)
    assert want_v.dtype == torch.bfloat16 and want_u.dtype == torch.bfloat16
    got_v, got_u = P.ValueUncertainty(blk)(mf, net)
    assert got_v.shape == want_v.shape and got_u.shape == want_u.shape == (shape[0], 1, shape[1], shape[2])
    assert got_v.dtype == torch.bfloat16 and got_u.dtype == torch.bfloat16 and got_v.is_contiguous()
    same_v = float((got_v == want_v).float().mean())
    same_u = float((got_u == want_u).float().mean())
    dv = float((got_v.float() - want_v.float()).abs().max())
    du = float((got_u.float() - want_u.float()).abs().max())
    print(f"K8 {shape} net {net_dtype}: value {same_v * 100:.2f} % bit-identical (max |d| {dv:.4f}), uncertainty {same_u * 100:.2f} % (max |d| {du:.4f})")
    assert same_v >= 0.99 and dv <= 2.0 ** -6 * max(1.0, float(want_v.float().abs().max()))
    assert same_u >= 0.97 and du <= 2.0 ** -8


def test_vs_cpu_oracle_bf16_mode(env):
    P, blk, _ = env
    mf, net = _inputs(2, 9, 13, seed=1)
    want_v, want_u = O.value_uncertainty(mf.cpu(), net.cpu(), blk.aggregator.to_v.weight.detach().float().cpu(),
                                         {k: v.detach().float().cpu() for k, v in blk.uncertainty.state_dict().items()}, mode="bf16")
    got_v, got_u = P.ValueUncertainty(blk)(mf, net)
    assert float((got_v.float().cpu() == want_v).float().mean()) >= 0.99
    assert float((got_u.float().cpu() - want_u).abs().max()) <= 2.0 ** -8
    assert float((got_u.float().cpu() == want_u).float().mean()) >= 0.97


def test_value_is_the_memory_reads_operand_in_place(env):
    """The NCHW bf16 value tensor is what PPMMemory.pack_kv uses without a copy (frame-major, hw % 8 == 0)."""
    P, blk, _ = env
    b, T, H, W = 1, 6, 8, 16
    mf, net = _inputs(b * T, H, W, seed=5)
    value, _ = P.ValueUncertainty(blk)(mf, net)
    g = torch.Generator(device="cuda").manual_seed(2)
    q = torch.randn(b, 128, T, H, W, generator=g, device="cuda")
    k = torch.randn(b, 128, T, H, W, generator=g, device="cuda")
    mem = P.PPMMemory(q, k, None, top_k=5)
    mem.pick(torch.rand(b, 1, T, T, generator=g, device="cuda"), torch.ones(b, 1, T, T, device="cuda"),
             torch.rand(b * T, 1, H, W, generator=g, device="cuda"))
    mem.pack_kv(value.reshape(b, T, 128, H, W).permute(0, 2, 1, 3, 4))
    assert mem.v_packed.data_ptr() == value.data_ptr()


def test_no_write_outside_outputs_and_argument_checks(env):
    P, blk, _ = env
    from ppmstereo_b200 import _lib
    B, H, W = 2, 13, 21
    mf, net = _inputs(B, H, W, seed=4)
    impl = P.ValueUncertainty(blk)
    want_v, want_u = impl(mf, net)
    n = ctypes.c_size_t()
    _lib.check(impl.lib.ppm_value_uncertainty_workspace(B, H, W, ctypes.byref(n)), "ws")
    G = 4096
    ws = torch.full((n.value + 2 * G,), 0x5A, dtype=torch.uint8, device="cuda")
    val = torch.full((B * 128 * H * W + 2 * G,), -77.0, device="cuda").bfloat16()
    unc = torch.full((B * H * W + 2 * G,), -77.0, device="cuda").bfloat16()
    args = [mf.data_ptr(), 0, net.data_ptr(), 0, impl.wv.data_ptr(), impl.bv.data_ptr(), impl.wu1.data_ptr(),
            impl.bu1.data_ptr(), impl.wu2.data_ptr(), impl.bu2.data_ptr(), B, H, W]
    _lib.check(impl.lib.ppm_value_uncertainty(*args, ws.data_ptr() + G, n.value, val.data_ptr() + 2 * G, unc.data_ptr() + 2 * G,
                                              _lib.stream_ptr()), "ppm_value_uncertainty")
    torch.cuda.synchronize()
    for t in (ws, val, unc):
        fill = 0x5A if t.dtype == torch.uint8 else -77.0
        assert bool((t[:G] == fill).all()) and bool((t[-G:] == fill).all())
    assert torch.equal(val[G:-G].view(B, 128, H, W), want_v) and torch.equal(unc[G:-G].view(B, 1, H, W), want_u)
    rc = impl.lib.ppm_value_uncertainty(*args, ws.data_ptr() + G, 16, val.data_ptr(), unc.data_ptr(), None)
    assert rc != 0 and "workspace" in impl.lib.ppm_last_error().decode()
    args[0] = None
    rc = impl.lib.ppm_value_uncertainty(*args, ws.data_ptr() + G, n.value, val.data_ptr(), unc.data_ptr(), None)
    assert rc != 0 and "null" in impl.lib.ppm_last_error().decode()


def test_installed_loop_uses_it_and_falls_back(env):
    """value_and_uncertainty(): K8 when the encoder class is installed and autocast is on; the block's own methods
    otherwise (fp32: bit-identical to them)."""
    P, blk, upd = env
    from ppmstereo_b200 import motion
    B, H, W = 2, 16, 32
    g = torch.Generator(device="cuda").manual_seed(8)
    flow = torch.randn(B, 2, H, W, generator=g, device="cuda") * 3
    corr = torch.randn(B, 36, H, W, generator=g, device="cuda")
    mhs = torch.randn(B, 64, H, W, generator=g, device="cuda")
    inp = torch.randn(B, 128, H, W, generator=g, device="cuda")
    net = torch.tanh(torch.randn(B, 128, H, W, generator=g, device="cuda"))
    with torch.no_grad():
        mf0, h0, v0 = blk.get_motion_and_value(flow, corr, mhs, inp)
        u0 = blk.get_uncertainty(torch.cat([net, v0], dim=1))
        got = motion.value_and_uncertainty(blk, flow, corr, mhs, inp, net)          # not installed, fp32: the block's methods
    for a, b_ in zip(got, (mf0, h0, v0, u0)):
        assert torch.equal(a, b_)
    ppm = importlib.import_module("models.core.ppmstereo")
    P.install(ppm)
    try:
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            mf1, h1, v1, u1 = motion.value_and_uncertainty(blk, flow, corr, mhs.bfloat16(), inp, net)
            want_v = blk.aggregator.to_v(mf1)
            want_u = blk.get_uncertainty(torch.cat([net, want_v], dim=1))
    finally:
        P.uninstall(ppm)
    assert v1.dtype == torch.bfloat16 and u1.dtype == torch.bfloat16
    assert float((v1 == want_v).float().mean()) >= 0.99 and float((u1.float() - want_u.float()).abs().max()) <= 2.0 ** -8
