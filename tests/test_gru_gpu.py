"""K5 SKSepConvGRU3D: the sm_100a kernel against (i) the reference's own module run by torch /
cuDNN on the same GPU under bf16 autocast (staged reference sources, oracle/stage_ref.py) and (ii) the CPU oracle's
bf16-mode restatement (oracle/ppm_oracle.py::sk_sep_conv_gru3d, pinned on the reference module in
tests/test_oracle_golden.py).

Bar: the unit is three chained gated stages whose every intermediate is a bf16 tensor.  Two correct evaluations that
differ only in fp32 accumulation order flip a bf16 rounding of a conv output here and there (one ulp = 2^-8 relative);
a flip in stage 1 propagates through stages 2-3 with gains < 1 (sigmoid / tanh gates, |h| <= 1).  So: max |diff| <= 4
bf16 ulps of the unit range (2^-6), mean |diff| <= 2e-4 -- and cuDNN-vs-oracle is reported next to ours for scale."""
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
    torch.manual_seed(0)
    gru = upd.SKSepConvGRU3D(hidden_dim=128, input_dim=384).cuda().eval()
    return P, gru


def make(b, T, H, W, seed, hdtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    h = torch.tanh(torch.randn(b, 128, T, H, W, generator=g, device="cuda")).to(hdtype)
    x = torch.randn(b, 384, T, H, W, generator=g, device="cuda")
    return h, x


@pytest.mark.parametrize("hdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 4, 6, 9), (1, 5, 20, 32), (2, 3, 23, 40), (1, 6, 16, 64), (1, 20, 46, 80)])
def test_vs_reference_module_under_autocast(env, shape, hdtype):
    """hdtype fp32 is what the model hands over (type promotion keeps the state in fp32); bf16 the other promotion."""
    P, gru = env
    h, x = make(*shape, seed=sum(shape), hdtype=hdtype)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        want = gru(h, x)
    assert want.dtype == hdtype
    got = P.SepConvGRU3D(gru)(h, x)
    torch.cuda.synchronize()
    assert got.shape == want.shape and got.dtype == hdtype
    d = (got.float() - want.float()).abs()
    same = float(((got.float() - want.float()).abs() <= 1e-6).float().mean()) if hdtype == torch.float32 else float((got == want).float().mean())
    print(f"GRU {shape} h {hdtype}: {same * 100:.2f} % bit-identical to torch/cuDNN, max |diff| {float(d.max()):.4f}, mean {float(d.mean()):.2e}")
    assert float(d.max()) <= 2.0 ** -6
    assert float(d.mean()) <= 2e-4
    # (an fp32 state shows EVERY bf16 flip of a conv output as a different fp32 value; a bf16 state rounds most away)
    assert same >= (0.97 if hdtype == torch.bfloat16 else 0.93)


@pytest.mark.parametrize("hdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 4, 6, 9), (2, 3, 5, 17)])
def test_vs_cpu_oracle_bf16_mode(env, shape, hdtype):
    P, gru = env
    h, x = make(*shape, seed=7 + sum(shape), hdtype=hdtype)
    params = {k: v.detach().float().cpu() for k, v in gru.state_dict().items()}
    want = O.sk_sep_conv_gru3d(h.float().cpu(), x.cpu(), params, mode="bf16",
                               hidden="fp32" if hdtype == torch.float32 else "bf16")
    got = P.SepConvGRU3D(gru)(h, x).float().cpu()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        cudnn = gru(h, x).float().cpu()
    d = (got - want).abs()
    print(f"GRU {shape} vs oracle: max {float(d.max()):.4f} mean {float(d.mean()):.2e}; cuDNN vs oracle: max "
          f"{float((cudnn - want).abs().max()):.4f} mean {float((cudnn - want).abs().mean()):.2e}")
    assert float(d.max()) <= 2.0 ** -6
    assert float(d.mean()) <= 2e-4


def test_strided_frame_major_inputs_and_fp32_hidden(env):
    """The model hands over views: h / x are (b, C, T, H, W) views of frame-major (b*T, C, H, W) tensors."""
    P, gru = env
    b, T, H, W = 1, 5, 12, 20
    h, x = make(b, T, H, W, seed=3)
    hv = h.permute(0, 2, 1, 3, 4).contiguous().permute(0, 2, 1, 3, 4)             # frame-major storage
    xv = x.permute(0, 2, 1, 3, 4).contiguous().permute(0, 2, 1, 3, 4)
    assert not hv.is_contiguous()
    k = P.SepConvGRU3D(gru)
    assert torch.equal(k(hv, xv), k(h, x))
    assert torch.equal(k(h, x.bfloat16()), k(h, x))                                # the cast the conv would do anyway


def test_weights_are_repacked_when_parameters_change(env):
    P, gru = env
    h, x = make(1, 3, 6, 16, seed=5)
    k = P.SepConvGRU3D(gru)
    a = k(h, x)
    with torch.no_grad():
        gru.convq3.bias.add_(0.5)
    try:
        b_ = k(h, x)
        assert not torch.equal(a, b_)
    finally:
        with torch.no_grad():
            gru.convq3.bias.sub_(0.5)
    assert torch.equal(k(h, x), a)


# ------------------------------------------------------------------------------------------------ FlowHead3D (K5b)
@pytest.fixture(scope="module")
def flow_head(env):
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(3)
    return upd.FlowHead3D(128, hidden_dim=256).cuda().eval()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 4, 6, 9), (2, 3, 23, 40), (1, 1, 16, 64), (1, 20, 46, 80)])
def test_flow_head_vs_reference_module_under_autocast(env, flow_head, shape, dtype):
    P, _ = env
    b, T, H, W = shape
    x = torch.tanh(torch.randn(b, 128, T, H, W, generator=torch.Generator(device="cuda").manual_seed(sum(shape)),
                               device="cuda")).to(dtype)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        want = flow_head(x)
    assert want.dtype == torch.bfloat16 and want.shape == (b, 2, T, H, W)
    got = P.FlowHead3D(flow_head)(x)
    assert got.shape == want.shape and got.dtype == torch.bfloat16
    d = (got.float() - want.float()).abs()
    same = float((got == want).float().mean())
    scale = float(want.float().abs().max())
    print(f"FlowHead3D {shape} {dtype}: {same * 100:.2f} % bit-identical to torch/cuDNN, max |diff| {float(d.max()):.4f} of {scale:.3f}")
    assert float(d.max()) <= 2.0 ** -6 * max(scale, 1.0)
    assert same >= 0.93          # (differences are single bf16 ulps of small values: see the max |diff| bound)


def test_flow_head_vs_cpu_oracle_and_state_reuse(env, flow_head):
    """(i) the CPU oracle's bf16-mode restatement; (ii) fed the tensor K5 just returned, the head reads K5's channels-last
    state instead of re-packing -- same result as with a fresh copy of that tensor."""
    P, gru = env
    h, x = make(1, 4, 10, 24, seed=11, hdtype=torch.float32)
    k = P.SepConvGRU3D(gru)
    head = P.FlowHead3D(flow_head)
    net = k(h, x)
    from ppmstereo_b200 import gru as G
    assert G._state_channels_last(net) != 0 and G._state_channels_last(net.clone()) == 0
    a = head(net)                     # channels-last state of K5
    b_ = head(net.clone())            # packed from the (b, c, T, H, W) tensor
    assert torch.equal(a, b_)
    params = {n: v.detach().float().cpu() for n, v in flow_head.state_dict().items()}
    want = O.flow_head3d(net.float().cpu(), params, mode="bf16")
    d = (a.float().cpu() - want).abs()
    assert float(d.max()) <= 2.0 ** -6 * max(float(want.abs().max()), 1.0)
    assert float(d.mean()) <= 2e-4


@pytest.mark.parametrize("shape", [(1, 3, 6, 9), (2, 2, 5, 17), (1, 4, 7, 70)])
def test_no_write_outside_output_and_workspace(env, flow_head, shape):
    """Guard bands instead of compute-sanitizer (closed on this pool): K5 / K5b called through the C ABI with the output
    and the workspace placed INSIDE larger sentinel-filled buffers, at shapes with partial tiles in x and y (masked
    epilogue stores, TMA boxes hanging over the tensor edge).  Nothing outside the declared regions may change, every
    output element must be written."""
    import ctypes
    from ppmstereo_b200 import _lib
    P, gru = env
    lib = _lib.load()
    b, T, H, W = shape
    h, x = make(b, T, H, W, seed=19, hdtype=torch.float32)
    k = P.SepConvGRU3D(gru)
    want = k(h, x)                                             # also packs the weights
    wptr, bptr = k._params()
    n = ctypes.c_size_t()
    _lib.check(lib.ppm_sepconv_gru3d_workspace(b, T, H, W, ctypes.byref(n)), "ws")
    guard = 4096
    ws = torch.full((n.value + 2 * guard,), 0x5A, dtype=torch.uint8, device="cuda")
    n_out = b * 128 * T * H * W
    outbuf = torch.full((n_out + 2 * 1024,), float("nan"), device="cuda")
    base = (ws.data_ptr() + guard + 255) // 256 * 256
    lo = base - ws.data_ptr()
    out_ptr = outbuf.data_ptr() + 1024 * 4
    _lib.check(lib.ppm_sepconv_gru3d(h.data_ptr(), _lib.dtype_code(h), h.stride(0), h.stride(1), h.stride(2), x.data_ptr(),
                                     _lib.dtype_code(x), x.stride(0), x.stride(1), x.stride(2), wptr, bptr, b, T, H, W,
                                     base, n.value, out_ptr, _lib.PPM_F32, _lib.stream_ptr()), "ppm_sepconv_gru3d")
    torch.cuda.synchronize()
    assert bool((ws[:lo] == 0x5A).all()) and bool((ws[lo + n.value:] == 0x5A).all())
    assert bool(torch.isnan(outbuf[:1024]).all()) and bool(torch.isnan(outbuf[1024 + n_out:]).all())
    got = outbuf[1024:1024 + n_out].view(b, 128, T, H, W)
    assert bool(torch.isfinite(got).all()) and torch.equal(got, want)
    # flow head, same scheme
    head = P.FlowHead3D(flow_head)
    want_f = head(want.clone())
    _lib.check(lib.ppm_flow_head3d_workspace(b, T, H, W, ctypes.byref(n)), "ws")
    ws = torch.full((n.value + 2 * guard,), 0x5A, dtype=torch.uint8, device="cuda")
    base = (ws.data_ptr() + guard + 255) // 256 * 256
    lo = base - ws.data_ptr()
    n_out = b * 2 * T * H * W
    outbuf = torch.full((n_out + 2 * 1024,), -7.0, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.ppm_flow_head3d(want.data_ptr(), _lib.dtype_code(want), want.stride(0), want.stride(1), want.stride(2),
                                   None, head.w1.data_ptr(), head.b1.data_ptr(), head.w2.data_ptr(), head.b2.data_ptr(),
                                   b, T, H, W, base, n.value, outbuf.data_ptr() + 1024 * 2, _lib.PPM_BF16,
                                   _lib.stream_ptr()), "ppm_flow_head3d")
    torch.cuda.synchronize()
    assert bool((ws[:lo] == 0x5A).all()) and bool((ws[lo + n.value:] == 0x5A).all())
    assert bool((outbuf[:1024] == -7.0).all()) and bool((outbuf[1024 + n_out:] == -7.0).all())
    assert torch.equal(outbuf[1024:1024 + n_out].view(b, 2, T, H, W), want_f)


# ------------------------------------------------------------------------------------------------ whole update block (K5c)
@pytest.mark.parametrize("shape", [(1, 4, 6, 9), (2, 3, 23, 40), (1, 6, 46, 80)])
def test_update_block_vs_reference_forward_under_autocast(env, shape):
    """SequenceUpdateBlock3D.forward (no update attention, 2-D mask) as ONE call -- GRU + flow head + mask head, the
    384-channel cat never built, frame-major in and out -- against the reference block on torch/cuDNN."""
    P, _ = env
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(9)
    blk = upd.SequenceUpdateBlock3D(hidden_dim=128, cor_planes=36, mask_size=4).cuda().eval()
    b, T, H, W = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    rn = lambda c: torch.randn(b * T, c, H, W, generator=g, device="cuda")
    net, inp, mf, mfg = torch.tanh(rn(128)), torch.relu(rn(128)).bfloat16(), rn(128), rn(128)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        want_net, want_mask, want_df = blk(net, inp, mf, mfg, t=T)
    k = P.UpdateBlock3D(blk)
    assert k.supported()
    got_net, got_mask, got_df = k(net, inp, mf, mfg, T)
    # the heads' own arithmetic: the reference heads applied to OUR GRU result (the 3 % one-ulp differences of `net` would
    # otherwise smear over every output of the 3 x 3 convolutions behind it)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        net5 = got_net.reshape(b, T, 128, H, W).permute(0, 2, 1, 3, 4)
        same_in_mask = 0.25 * blk.mask_2d(got_net)
        same_in_df = blk.flow_head(net5).permute(0, 2, 1, 3, 4).reshape(b * T, 2, H, W)
    checks = (("net", got_net, want_net, 0.93), ("mask (same input)", got_mask, same_in_mask, 0.97),
              ("delta_flow (same input)", got_df, same_in_df, 0.93), ("mask", got_mask, want_mask, 0.0),
              ("delta_flow", got_df, want_df, 0.0))
    for name, got, want, min_same in checks:
        assert got.shape == want.shape and got.dtype == want.dtype, (name, got.shape, want.shape, got.dtype, want.dtype)
        d = (got.float() - want.float()).abs()
        scale = max(1.0, float(want.float().abs().max()))
        same = float((d <= 1e-6).float().mean())
        print(f"update block {shape} {name}: {same * 100:.2f} % identical, max |diff| {float(d.max()):.4f} of {scale:.2f}")
        assert float(d.max()) <= 2.0 ** -6 * scale, name
        assert float(d.mean()) <= 2e-4 * scale, name
        assert same >= min_same, name
