"""Pin the CPU oracle against fixtures produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import ppm_oracle as O

CORR_CASES = ["corr_d256_w82", "corr_d128_w64", "corr_d256_w160"]


def _load(golden_dir, name):
    return {k: v for k, v in np.load(os.path.join(golden_dir, name + ".npz")).items()}


@pytest.mark.parametrize("name", CORR_CASES)
def test_corr_pyramid_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    lv = O.corr_pyramid(torch.from_numpy(g["fmap1"]), torch.from_numpy(g["fmap2"]))
    assert len(lv) == 5
    for l in range(5):
        ref = torch.from_numpy(g[f"level{l}"])
        assert lv[l].shape == ref.shape
        # fp32 matmul summation order differs from einsum's: 1e-5 relative to the level's scale
        assert (lv[l] - ref).abs().max() <= 1e-5 * ref.abs().max()


@pytest.mark.parametrize("name", CORR_CASES)
def test_corr_lookup_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    levels = [torch.from_numpy(g[f"level{l}"]) for l in range(5)]
    out = O.corr_lookup(levels, torch.from_numpy(g["flow"]))
    ref = torch.from_numpy(g["lookup"])
    assert out.shape == ref.shape
    # north_star bar: 1e-4 abs vs grid_sample (its normalise/denormalise round trip is ~2e-5)
    assert (out - ref).abs().max() <= 1e-4


@pytest.mark.parametrize("name", ["ppm_read_l16", "ppm_read_l08"])
def test_ppm_read_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16 and v.ndim > 0}
    b, c, T, h, w = t["query"].shape
    pe = O.temporal_pe(T, c)
    sim = O.qk_similarity(t["query"], t["key"])
    assert (sim - t["sim"]).abs().max() <= 1e-5
    assert abs(O.ppm_softmax_scale(c) - float(g["scale"])) < 1e-12
    r = O.ppm_memory_read(t["query"], t["key"], pe, t["value"], t["motion_features"], t["sim"],
                          t["strive_before"], t["uncertainty"], float(g["beta"]))
    assert (r["frame_score"] - t["frame_score"]).abs().max() <= 1e-6
    assert torch.equal(r["mask"], t["mask"])                       # bit-exact top-k sets
    top5 = torch.sort(t["argsort"][..., :5], dim=-1).values[:, 0]
    assert torch.equal(r["picked"], top5)
    # modulated keys as fed to flash_attn (bf16 bits), frame-major then pixel
    kmod = torch.from_numpy(g["kmod_bf16"].astype(np.int16)).view(torch.bfloat16).float()
    hw = h * w
    for i in range(T):
        for n, j in enumerate(r["picked"][0, i].tolist()):
            mine = O.bf16_round((t["key"][0, :, j].reshape(c, hw) * r["score_norm"][0, i, n]
                                 + pe[j][:, None]).t())
            ref = kmod[i, n * hw:(n + 1) * hw]
            # one bf16 ulp: the reference forms score_norm with a different op order
            assert (mine - ref).abs().max() <= 2.0 ** -7 * ref.abs().max()
    hid = torch.from_numpy(g["hidden_bf16"].astype(np.int16)).view(torch.bfloat16).float()
    mine = r["hidden"][0].permute(0, 2, 1)                           # T,hw,c
    assert (mine - hid).abs().max() <= 2e-2 * hid.abs().max()
    assert (mine - hid).abs().mean() <= 2e-3 * hid.abs().mean() + 1e-6
    out = r["motion_features_global"]
    assert (out - t["motion_features_global"]).abs().max() <= 2e-2 * hid.abs().max()


def test_temporal_pe_endpoints():
    pe = O.temporal_pe(20, 128)
    assert torch.equal(pe[0, 0::2], torch.zeros(64)) and torch.equal(pe[0, 1::2], torch.ones(64))
    assert abs(float(pe[-1, 0]) - np.sin(1.0)) < 1e-6
    assert torch.isnan(O.temporal_pe(1, 8)).all()


def test_topk_no_op_when_T_le_k():
    s = torch.rand(2, 1, 4, 4)
    assert O.topk_pick(s, 5).all()


@pytest.mark.parametrize("kind", ["fp32", "bf16"])
def test_convex_upsample_matches_reference(golden_dir, kind):
    g = _load(golden_dir, "convex_upsample")
    flow = torch.from_numpy(g["flow"])
    if kind == "fp32":
        mask, ref = torch.from_numpy(g["mask"]), torch.from_numpy(g["out_fp32_mask"])
    else:
        mask = torch.from_numpy(g["mask_bf16"].astype(np.int16)).view(torch.bfloat16)
        ref = torch.from_numpy(g["out_bf16_mask"])
    out = O.convex_upsample(flow, mask, 4)
    assert out.shape == ref.shape
    # bf16 mask = storage only: under the reference's CUDA autocast the softmax runs (and stays) in fp32
    assert (out - ref).abs().max() <= 1e-5 * ref.abs().max()   # fp32 summation order only


def test_gru_oracle_matches_reference_module():
    """SKSepConvGRU3D restated as shifted channel contractions (oracle) == the reference module (fp32, CPU)."""
    import importlib
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present")
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(0)
    gru = upd.SKSepConvGRU3D(hidden_dim=128, input_dim=384).eval()
    g = torch.Generator().manual_seed(1)
    for shape in ((1, 4, 6, 9), (2, 2, 3, 17)):
        b, T, H, W = shape
        h = torch.tanh(torch.randn(b, 128, T, H, W, generator=g))
        x = torch.randn(b, 384, T, H, W, generator=g)
        with torch.no_grad():
            want = gru(h, x)
        got = O.sk_sep_conv_gru3d(h, x, dict(gru.state_dict()))
        assert float((got - want).abs().max()) <= 2e-6
        # the bf16-autocast restatements stay within bf16 rounding of the fp32 result
        for hidden in ("bf16", "fp32"):
            lo = O.sk_sep_conv_gru3d(h, x, dict(gru.state_dict()), mode="bf16", hidden=hidden)
            assert float((lo - want).abs().max()) <= 3e-2


def test_flow_head_oracle_matches_reference_module():
    import importlib
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present")
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(1)
    fh = upd.FlowHead3D(128, hidden_dim=256).eval()
    x = torch.randn(2, 128, 3, 5, 7, generator=torch.Generator().manual_seed(2))
    with torch.no_grad():
        want = fh(x)
    got = O.flow_head3d(x, dict(fh.state_dict()))
    assert got.shape == want.shape == (2, 2, 3, 5, 7)
    assert float((got - want).abs().max()) <= 1e-5


def test_motion_corr_front_oracle_matches_reference_module():
    import importlib
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present")
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(2)
    pc = upd.PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]).eval()
    corr = torch.randn(2, 36, 9, 13, generator=torch.Generator().manual_seed(3)) * 2
    with torch.no_grad():
        want = torch.nn.functional.gelu(pc(corr))
    got = O.motion_corr_front(corr, dict(pc.state_dict()))
    assert got.shape == want.shape == (2, 256, 9, 13)
    assert float((got - want).abs().max()) <= 2e-6
    lo = O.motion_corr_front(corr, dict(pc.state_dict()), mode="bf16")
    assert float((lo - want).abs().max()) <= 3e-2 * float(want.abs().max())


def test_motion_encoder_oracle_matches_reference_module():
    import importlib
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present")
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(4)
    enc = upd.BasicMotionEncoder_v2(36).eval()
    g = torch.Generator().manual_seed(5)
    flow = torch.randn(2, 2, 7, 11, generator=g) * 3
    corr = torch.randn(2, 36, 7, 11, generator=g) * 2
    mhs = torch.randn(2, 64, 7, 11, generator=g)
    with torch.no_grad():
        want_f, want_h = enc(flow, corr, mhs, None)
    got_f, got_h = O.motion_encoder(flow, corr, mhs, dict(enc.state_dict()))
    assert got_f.shape == want_f.shape == (2, 128, 7, 11) and got_h.shape == want_h.shape == (2, 64, 7, 11)
    assert float((got_f - want_f).abs().max()) <= 5e-6 and float((got_h - want_h).abs().max()) <= 5e-6
    assert torch.equal(got_f[:, 126:], flow)
    lo_f, lo_h = O.motion_encoder(flow, corr, mhs, dict(enc.state_dict()), mode="bf16")
    assert float((lo_f - want_f).abs().max()) <= 3e-2 * float(want_f.abs().max())
    assert float((lo_h - want_h).abs().max()) <= 3e-2 * max(1.0, float(want_h.abs().max()))


def test_value_uncertainty_oracle_matches_reference_modules():
    import importlib
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present")
    ref_import.load_reference()
    upd = importlib.import_module("models.core.ppmtereo_update")
    torch.manual_seed(6)
    blk = upd.SequenceUpdateBlock3D(hidden_dim=128, cor_planes=36, mask_size=4).eval()
    g = torch.Generator().manual_seed(7)
    mf = torch.randn(2, 128, 6, 9, generator=g)
    net = torch.tanh(torch.randn(2, 128, 6, 9, generator=g))
    with torch.no_grad():
        want_v = blk.aggregator.to_v(mf)
        want_u = blk.get_uncertainty(torch.cat([net, want_v], dim=1))
    got_v, got_u = O.value_uncertainty(mf, net, blk.aggregator.to_v.weight.detach(), dict(blk.uncertainty.state_dict()))
    assert got_v.shape == want_v.shape == (2, 128, 6, 9) and got_u.shape == want_u.shape == (2, 1, 6, 9)
    assert float((got_v - want_v).abs().max()) <= 2e-6 and float((got_u - want_u).abs().max()) <= 2e-6
    lo_v, lo_u = O.value_uncertainty(mf, net, blk.aggregator.to_v.weight.detach(), dict(blk.uncertainty.state_dict()), mode="bf16")
    assert float((lo_v - want_v).abs().max()) <= 2e-2 * float(want_v.abs().max())
    assert float((lo_u - want_u).abs().max()) <= 1e-2
