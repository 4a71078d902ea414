"""GPU parity tests for K1 (correlation pyramid build) and K2 (lookup), through the C ABI.

Checked against (i) golden vectors produced by the unmodified reference (tests/golden), (ii) the CPU
oracle on seeded inputs, (iii) size-independent properties at BASELINE.json's full shapes.
Tolerances (north_star): volumes <= 1e-3 relative with the operand precision stated per test
(bf16 operands / TF32 operands vs the fp32 oracle); lookup <= 1e-4 abs vs grid_sample.
"""
import os

import numpy as np
import pytest
import torch

from oracle import ppm_oracle as O

pytestmark = pytest.mark.gpu

CASES = ["corr_d256_w82", "corr_d128_w64", "corr_d256_w160"]


@pytest.fixture(scope="module")
def P():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ppmstereo_b200 as P
    return P


def load(golden_dir, name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(golden_dir, name + ".npz")).items()}


def rel(a, b):
    return float((a - b).norm() / b.norm())


def flat(levels):
    return [l.reshape(l.shape[0], l.shape[-1]).float().cpu() for l in levels]


@pytest.mark.parametrize("name", CASES)
def test_build_tf32_vs_reference_golden(P, golden_dir, name):
    """fp32 feature maps -> kind::tf32 MMA, fp32 output; vs the reference's fp32 CPU pyramid."""
    g = load(golden_dir, name)
    blk = P.CorrBlock1D(g["fmap1"].cuda(), g["fmap2"].cuda())
    assert len(blk.corr_pyramid) == 5
    got = flat(blk.corr_pyramid)
    want_tf32 = O.corr_pyramid(g["fmap1"], g["fmap2"], mode="tf32")
    for l in range(5):
        ref = g[f"level{l}"]
        assert tuple(blk.corr_pyramid[l].shape) == (ref.shape[0], 1, 1, ref.shape[1])
        assert rel(got[l], ref) <= 1e-3, (l, rel(got[l], ref))          # TF32 operands vs fp32
        assert rel(got[l], want_tf32[l]) <= 2e-5, (l, rel(got[l], want_tf32[l]))
    assert tuple(blk.coords.shape) == tuple(g["flow"].shape)


@pytest.mark.parametrize("name", CASES)
def test_build_bf16_operands(P, golden_dir, name):
    """bf16 feature maps (what autocast hands over on the GPU) -> kind::f16 MMA."""
    g = load(golden_dir, name)
    f1, f2 = g["fmap1"].bfloat16(), g["fmap2"].bfloat16()
    # fp32 output: exact products, fp32 accumulation
    got = flat(P.build_pyramid(f1.cuda(), f2.cuda(), 4, out_dtype=torch.float32))
    want = O.corr_pyramid(f1.float(), f2.float(), mode="fp32")
    for l in range(5):
        assert rel(got[l], want[l]) <= 1e-5, (l, rel(got[l], want[l]))
        assert rel(got[l], g[f"level{l}"]) <= 5e-3        # bf16 operands vs the fp32 reference
    # bf16 output with the reference's cascaded rounding
    got = flat(P.build_pyramid(f1.cuda(), f2.cuda(), 4))
    want = O.corr_pyramid(f1.float(), f2.float(), mode="bf16")
    for l in range(5):
        d = (got[l] - want[l]).abs()
        # one bf16 ulp of the element, or of a typical element of the level: pooled values can
        # cancel, so a 1-ulp difference of an input pair may be several ulps of their small mean
        ulp = (want[l].abs() + want[l].pow(2).mean().sqrt()) * 2.0 ** -7
        assert bool((d <= ulp).all()), (l, float((d / ulp).max()))
        assert float((d == 0).float().mean()) >= 0.995, (l, float((d == 0).float().mean()))


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("vol", ["fp32", "bf16"])
def test_lookup_vs_reference_golden(P, golden_dir, name, vol):
    g = load(golden_dir, name)
    levels = [g[f"level{l}"] for l in range(5)]
    if vol == "bf16":
        levels = [l.bfloat16() for l in levels]
    dev = [l.reshape(l.shape[0], 1, 1, l.shape[1]).cuda() for l in levels]
    got = P.lookup(dev, g["flow"].cuda(), W2=levels[0].shape[1]).cpu()
    assert got.shape == g["lookup"].shape and got.dtype == torch.float32
    if vol == "fp32":
        # the fixture was made by the CPU reference, whose `/ (W - 1)` is a true division; the kernels follow the GPU
        # reference, where ATen multiplies by the reciprocal (see lk_sample): <= 6e-5 apart, inside the 1e-4 bar.  The
        # bit-level pin is test_lookup_bit_identical_to_torch_grid_sample_on_gpu.
        assert float((got - g["lookup"]).abs().max()) <= 1e-4
    else:
        want = O.corr_lookup([l.float() for l in levels], g["flow"])
        assert float((got - want).abs().max()) <= 1e-4


def test_block_end_to_end_G5(P):
    """BASELINE config 2 shape: (5, D, 80, 128), D in {128, 256}, flow in [-40, 0], r=4, 4 levels."""
    for D, seed in ((128, 1), (256, 2)):
        g = torch.Generator().manual_seed(seed)
        f1 = torch.randn(5, D, 80, 128, generator=g).bfloat16()
        f2 = torch.randn(5, D, 80, 128, generator=g).bfloat16()
        flow = torch.zeros(5, 2, 80, 128)
        flow[:, 0] = torch.rand(5, 80, 128, generator=g) * -40.0
        blk = P.CorrBlock1D(f1.cuda(), f2.cuda())
        out = blk(flow.cuda()).cpu()
        want_lv = O.corr_pyramid(f1.float(), f2.float(), mode="bf16")
        want = O.corr_lookup(want_lv, flow)
        got_lv = flat(blk.corr_pyramid)
        for l in range(5):
            assert rel(got_lv[l], want_lv[l]) <= 1e-3
        # lookup of the kernel's own volume, exact interpolation check
        want_own = O.corr_lookup(got_lv, flow)
        assert float((out - want_own).abs().max()) <= 1e-4
        assert rel(out, want) <= 1e-3
        c = P.CorrBlock1D.corr(f1.cuda(), f2.cuda())
        assert tuple(c.shape) == (5, 80, 128, 1, 128)
        assert torch.equal(c.reshape(-1, 128).float().cpu(), got_lv[0])


def test_full_size_properties_G720(P):
    """(20, 256, 184, 320) bf16: pooling identity on every level + sampled rows vs the oracle."""
    g = torch.Generator(device="cuda").manual_seed(3)
    f1 = torch.randn(20, 256, 184, 320, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(20, 256, 184, 320, generator=g, device="cuda").bfloat16()
    lv = P.build_pyramid(f1, f2, 4, out_dtype=torch.float32)
    for l in range(4):
        a = lv[l].reshape(-1, lv[l].shape[-1])
        w = a.shape[1] // 2
        pooled = (a[:, 0:2 * w:2] + a[:, 1:2 * w:2]) * 0.5
        assert torch.equal(pooled, lv[l + 1].reshape(-1, w)), l            # bit-exact in fp32
    rows = torch.randint(0, 20 * 184, (48,), generator=torch.Generator().manual_seed(4))
    L0 = lv[0].reshape(20 * 184, 320, 320)
    for r in rows.tolist():
        b, h = divmod(r, 184)
        want = (f1[b, :, h].float().t().cpu() @ f2[b, :, h].float().cpu()) / 16.0
        assert rel(L0[r].cpu(), want) <= 1e-5
    # lookup at full size: out-of-range taps give exact zeros, in-range centre tap is a lerp of L0
    flow = torch.zeros(20, 2, 184, 320, device="cuda")
    flow[:, 0] = -1000.0
    blk_out = P.lookup(lv, flow, W2=320)
    assert float(blk_out.abs().max()) == 0.0
    flow[:, 0] = 0.0
    out = P.lookup(lv, flow, W2=320)
    diag = L0[:, torch.arange(320), torch.arange(320)].reshape(20, 184, 320)
    # level 0, k = 0 -> corr[i, i], up to the reference's own grid_sample normalise/un-normalise
    # round trip: a few ulp(320) = 3e-5 of position times |dL/dx| ~ 1..5
    assert float((out[:, 4] - diag).abs().max()) <= 5e-4


def test_errors(P):
    f = torch.zeros(1, 16, 2, 8, device="cuda")
    with pytest.raises(RuntimeError, match="too small"):
        P.CorrBlock1D(f, f)                                                 # reference: W/16 >= 1
    with pytest.raises(RuntimeError, match="dtype"):
        P.CorrBlock1D(f.half(), f.half())


@pytest.mark.parametrize("shape", [(2, 128, 3, 64, 64), (1, 256, 2, 40, 88), (3, 256, 5, 320, 320), (1, 128, 1, 8, 16), (1, 128, 2, 24, 40)])
def test_lookup_from_level0_equals_generic_lookup(P, shape):
    """K2b (all four levels rebuilt from one level-0 span) must equal the per-level lookup bit for bit on a
    pyramid written by the build kernel -- including taps left / right of the row, the first and last rows of
    the volume (guarded loads) and far out-of-range flows."""
    B, D, H, W1, W2 = shape
    g = torch.Generator().manual_seed(W1 * 7 + W2)
    f1 = torch.randn(B, D, H, W1, generator=g).bfloat16().cuda()
    f2 = torch.randn(B, D, H, W2, generator=g).bfloat16().cuda()
    lv = P.build_pyramid(f1, f2, 4)
    for lo, hi in ((-W2 - 24.0, 24.0), (-3.0, 3.0), (-1e6, 1e6)):
        flow = torch.zeros(B, 2, H, W1)
        flow[:, 0] = torch.rand(B, H, W1, generator=g) * (hi - lo) + lo
        flow[0, 0, 0, 0] = -5.5                                  # first pixel of the volume, taps left of it
        flow[-1, 0, -1, -1] = 6.25 + (W2 - W1)                   # last pixel, taps right of the row end
        want = P.lookup(lv, flow.cuda(), W2=W2)
        got = P.lookup_l0(lv[0], flow.cuda(), W2=W2)
        assert torch.equal(got, want), (shape, lo, hi, float((got - want).abs().max()))
    blk = P.CorrBlock1D(f1, f2)
    assert torch.equal(blk(flow.cuda()), P.lookup(blk.corr_pyramid, flow.cuda(), W2=W2))


@pytest.mark.parametrize("shape", [(1, 256, 1, 320, 320), (2, 128, 75, 320, 320), (1, 256, 3, 288, 320), (1, 64, 2, 130, 272),
                                   (2, 192, 2, 320, 320)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_build_cluster_path_all_rows(P, shape, out_dtype):
    """Rows with exactly two N tiles (256 < W2 <= 512) and several M tiles are built by 2-CTA clusters: CTA rank =
    N tile, F1 tiles multicast to both CTAs.  Every output row of every level is checked (1 row = 1 cluster, 150 rows = more rows than
    co-resident clusters, partial M tiles, W1 != W2, a D that is not a multiple of the 64-channel stage)."""
    B, D, H, W1, W2 = shape
    g = torch.Generator(device="cuda").manual_seed(D + H + W1)
    f1 = torch.randn(B, D, H, W1, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(B, D, H, W2, generator=g, device="cuda").bfloat16()
    lv = P.build_pyramid(f1, f2, 4, out_dtype=out_dtype)
    want = torch.einsum("bdhi,bdhj->bhij", f1.float(), f2.float()).reshape(B * H * W1, W2) / float(D) ** 0.5
    for l in range(5):
        got = lv[l].reshape(B * H * W1, W2 >> l).float()
        assert bool(torch.isfinite(got).all()), l
        tol = 1e-5 if out_dtype == torch.float32 else 5e-3          # fp32 out: accumulation order only; bf16 out: cascaded bf16 rounding
        assert rel(got, want) <= tol, (l, rel(got, want))
        # per-tile check: no (M tile, N half) may be off (a wrong descriptor shows up in one CTA rank only)
        for n_half in range(2):
            c0, c1 = (n_half * (W2 >> l)) // 2, ((n_half + 1) * (W2 >> l)) // 2
            assert rel(got[:, c0:c1], want[:, c0:c1]) <= tol, (l, n_half)
        want = (want[:, 0:2 * (want.shape[1] // 2):2] + want[:, 1:2 * (want.shape[1] // 2):2]) * 0.5


@pytest.mark.parametrize("shape", [(1, 256, 3, 320, 320), (2, 128, 40, 320, 288)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_build_cluster_path_tf32_inputs(P, shape, out_dtype):
    """fp32 feature maps (kind::tf32, four 32-pixel boxes per F1 stage, two multicast by each CTA of the cluster):
    level 0 of both N halves vs the fp64 contraction.  TF32 operands vs fp32: <= 1e-3 relative (measured 7.7e-4);
    a bf16 pyramid adds its own rounding (measured 1.8e-3 .. 2.5e-3)."""
    B, D, H, W1, W2 = shape
    g = torch.Generator(device="cuda").manual_seed(1)
    f1 = torch.randn(B, D, H, W1, generator=g, device="cuda")
    f2 = torch.randn(B, D, H, W2, generator=g, device="cuda")
    lv = P.build_pyramid(f1, f2, 4, out_dtype=out_dtype)
    want = torch.einsum("bdhi,bdhj->bhij", f1.double(), f2.double()).reshape(B * H * W1, W2) / float(D) ** 0.5
    got = lv[0].reshape(B * H * W1, W2).double()
    tol = 1e-3 if out_dtype == torch.float32 else 5e-3
    for a, b in ((0, W2 // 2), (W2 // 2, W2)):
        assert rel(got[:, a:b], want[:, a:b]) <= tol, (a, b, rel(got[:, a:b], want[:, a:b]))
    for l in range(5):
        assert bool(torch.isfinite(lv[l].float()).all()), l


def test_block_at_headline_shape_sampled_pixels_vs_oracle(P):
    """(20, 256, 184, 320) bf16 through CorrBlock1D.__init__ / __call__ -- the route the model takes (K1, then the
    level-0-span lookup K2b) -- against O.corr_pyramid(mode="bf16") + O.corr_lookup on 96 sampled image rows."""
    g = torch.Generator(device="cuda").manual_seed(31)
    B, D, H, W = 20, 256, 184, 320
    f1 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    flow = torch.zeros(B, 2, H, W, device="cuda")
    flow[:, 0] = torch.rand(B, H, W, generator=g, device="cuda") * -60.0 + 10.0      # incl. taps left of the row
    blk = P.CorrBlock1D(f1, f2)
    out = blk(flow)
    assert out.shape == (B, 36, H, W) and out.dtype == torch.float32
    rows = torch.randperm(B * H, generator=torch.Generator().manual_seed(32))[:96].tolist() + [0, B * H - 1]
    for r in rows:
        b, h = divmod(r, H)
        lv = O.corr_pyramid(f1[b:b + 1, :, h:h + 1].float().cpu(), f2[b:b + 1, :, h:h + 1].float().cpu(), mode="bf16")
        want = O.corr_lookup(lv, flow[b:b + 1, :, h:h + 1].cpu())
        got = out[b:b + 1, :, h:h + 1].cpu()
        # bf16 volume entries: the tensor-core accumulation order differs from the oracle's matmul, which flips a
        # rare bf16 rounding (one ulp = 2^-8 relative of an O(1) entry); everything else is exact lerp arithmetic
        assert float((got - want).abs().max()) <= 2.0 ** -6, r
        assert float((got - want).abs().mean()) <= 1e-4, r
        assert rel(got, want) <= 1e-3, r


def _torch_lookup_sequence(pyramid, flow, num_levels=4, r=4):
    """CorrBlock1D.__call__ + bilinear_sampler (corr.py:74-94, :10-27) verbatim, run by torch on the GPU."""
    import torch.nn.functional as F
    B, _, H, W = flow.shape
    ys, xs = torch.meshgrid(torch.arange(H, device=flow.device), torch.arange(W, device=flow.device), indexing="ij")
    coords = torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1) + flow
    coords = coords[:, :1].permute(0, 2, 3, 1)
    out_pyramid = []
    for i in range(num_levels):
        corr = pyramid[i]
        dx = torch.linspace(-r, r, 2 * r + 1).view(1, 1, 2 * r + 1, 1).to(coords.device)
        x0 = dx + coords.reshape(B * H * W, 1, 1, 1) / 2 ** i
        y0 = torch.zeros_like(x0)
        grid = torch.cat([x0, y0], dim=-1)
        Wl = corr.shape[-1]
        xgrid, ygrid = grid.split([1, 1], dim=-1)
        xgrid = 2 * xgrid / (Wl - 1) - 1
        img = F.grid_sample(corr.contiguous().float(), torch.cat([xgrid, ygrid], dim=-1).contiguous().float(),
                            align_corners=True)
        out_pyramid.append(img.view(B, H, W, -1))
    return torch.cat(out_pyramid, dim=-1).permute(0, 3, 1, 2).contiguous().float()


@pytest.mark.parametrize("shape", [(2, 5, 80, 128), (1, 3, 46, 80), (1, 2, 23, 328)])
def test_lookup_bit_identical_to_torch_grid_sample_on_gpu(P, shape):
    """The lookup kernels evaluate the reference's fp32 op sequence (x0 = dx + c/2^l; 2*x0*(1/(W-1)) - 1;
    ((g+1)/2)*(W-1); floor; cuDNN's weight / blend form, see lk_sample) so the result equals what torch's
    grid_sample (-> cuDNN spatial-transformer sampler) produces on the same GPU -- both the level-0-span kernel (K2b,
    via CorrBlock1D.__call__) and the generic kernel (P.lookup).  Not identical: taps with -1 < ix < 0 (one-ulp
    weight rounding of cuDNN's that is not identified) and window-edge taps at a rounding boundary."""
    B, H, W1, W2 = shape
    g = torch.Generator(device="cuda").manual_seed(W2)
    f1 = torch.randn(B, 64, H, W1, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(B, 64, H, W2, generator=g, device="cuda").bfloat16()
    flow = torch.zeros(B, 2, H, W1, device="cuda")
    flow[:, 0] = torch.rand(B, H, W1, generator=g, device="cuda") * -(W2 * 0.6) + 8.0     # incl. both row ends
    blk = P.CorrBlock1D(f1, f2)
    pyr = blk.corr_pyramid
    want = _torch_lookup_sequence(pyr, flow)
    for got in (blk(flow), P.lookup(pyr, flow, W2=W2)):
        same = float((got == want).float().mean())
        print(f"lookup vs torch grid_sample: {same * 100:.4f} % bit-identical, max |diff| {float((got - want).abs().max()):.2e}")
        assert same >= 0.975, same
        assert float((got - want).abs().max()) <= 2e-6
    # fp32 pyramid through the generic kernel
    pyr32 = [p.float() for p in pyr]
    want32 = _torch_lookup_sequence(pyr32, flow)
    got32 = P.lookup(pyr32, flow, W2=W2)
    assert float((got32 == want32).float().mean()) >= 0.975


@pytest.mark.parametrize("shape", [(2, 256, 6, 128), (1, 256, 3, 320), (1, 128, 4, 64)])
def test_build_vs_torch_autocast_sequence_on_gpu(P, shape):
    """CorrBlock1D.__init__ (corr.py:56-72, 96-104) verbatim under the reference's bf16 autocast, run by torch
    (cuBLAS bf16 bmm + division + avg_pool2d) on the same GPU: the pyramid must agree to the bit except where the
    two fp32 accumulation orders straddle a bf16 rounding boundary."""
    import torch.nn.functional as F
    B, D, H, W = shape
    g = torch.Generator(device="cuda").manual_seed(D + W)
    f1 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    f2 = torch.randn(B, D, H, W, generator=g, device="cuda").bfloat16()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        corr = torch.einsum("aijk, aijh->ajkh", f1, f2)
        corr = corr.reshape(B, H, W, 1, W).contiguous()
        corr = corr / torch.sqrt(torch.tensor(D).float())
        assert corr.dtype == torch.bfloat16
        corr = corr.reshape(B * H * W, 1, 1, W)
        want = [corr]
        for _ in range(4):
            corr = F.avg_pool2d(corr, [1, 2], stride=[1, 2])
            want.append(corr)
    got = P.CorrBlock1D(f1, f2).corr_pyramid
    for l in range(5):
        assert got[l].shape == want[l].shape and got[l].dtype == want[l].dtype == torch.bfloat16
        same = float((got[l] == want[l]).float().mean())
        d = (got[l].float() - want[l].float()).abs()
        print(f"level {l}: {same * 100:.4f} % bit-identical to torch, max |diff| {float(d.max()):.3e}")
        assert same >= 0.998, (l, same)
        assert float(d.max()) <= 2.0 ** -7 * float(want[l].float().abs().max())
