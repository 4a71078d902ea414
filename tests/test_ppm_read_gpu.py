"""GPU parity tests for K3 (QAM score, top-k pick, key modulation, softmax aggregation, residual),
through the C ABI, against the reference-generated goldens and the CPU oracle.

Bars (north_star): top-k frame indices bit-exact; aggregation within bf16 rounding of the
FlashAttention-2 arithmetic the reference calls (P and the output are bf16 there), stated per test.
"""
import os

import numpy as np
import pytest
import torch

from oracle import ppm_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ppmstereo_b200 as P
    return P


def bf16_from_bits(a):
    return torch.from_numpy(a.astype(np.int16)).view(torch.bfloat16).float()


def run_read(P, t, top_k=5, scale=None, beta=0.5):
    mem = P.PPMMemory(t["query"].cuda(), t["key"].cuda(), None, top_k=top_k)
    strive = t["strive_before"].clone().cuda()
    out, mask = mem.read(t["value"].cuda(), t["motion_features"].cuda(), t["sim"].cuda(), strive,
                         t["uncertainty"].cuda(), torch.tensor([beta], device="cuda"), scale)
    torch.cuda.synchronize()
    return mem, out.cpu(), mask.cpu(), strive.cpu()


@pytest.mark.parametrize("name", ["ppm_read_l16", "ppm_read_l08"])
def test_read_vs_reference_golden(P, golden_dir, name):
    g = dict(np.load(os.path.join(golden_dir, name + ".npz")))
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16 and v.ndim > 0}
    b, c, T, h, w = t["query"].shape
    hw = h * w
    mem, out, mask, strive = run_read(P, t)
    assert float((mem.frame_score.cpu() - t["frame_score"]).abs().max()) <= 1e-6
    assert torch.equal(mask, t["mask"])                                    # bit-exact top-k sets
    top5 = torch.sort(t["argsort"][..., :5], dim=-1).values[:, 0].int()
    assert torch.equal(mem.picked.cpu(), top5)                             # bit-exact indices, ascending
    assert torch.equal(strive, t["strive_before"] + t["mask"].float())
    # modulated keys exactly as the reference fed them to flash_attn (bf16 bits)
    kmod = bf16_from_bits(g["kmod_bf16"]).reshape(T, 5, hw, c)
    mine = mem.k_packed.float().cpu().reshape(T, 5, hw, c)
    assert float((mine - kmod).abs().max()) <= 2.0 ** -7 * float(kmod.abs().max())
    assert float((mine == kmod).float().mean()) >= 0.99
    # aggregation + residual vs the reference's motion_features_global
    hid = bf16_from_bits(g["hidden_bf16"])                                 # T, hw, c
    mine_hid = (out.reshape(T, c, hw) - t["motion_features"][0].permute(1, 0, 2, 3).reshape(T, c, hw)) / 0.5
    d = (mine_hid.permute(0, 2, 1) - hid).abs()
    assert float(d.max()) <= 2e-2 * float(hid.abs().max()), float(d.max())
    assert float(d.mean()) <= 2e-3 * float(hid.abs().mean()) + 1e-6
    ref = t["motion_features_global"]
    assert float((out - ref).abs().max()) <= 1e-2 * float(hid.abs().max())


def make_case(b, T, h, w, seed, peaky=False):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(b, 128, T, h, w, generator=g)
    k = q + 0.1 * torch.randn(b, 128, T, h, w, generator=g) if peaky else torch.randn(b, 128, T, h, w, generator=g)
    return {
        "query": q, "key": k, "value": torch.randn(b, 128, T, h, w, generator=g),
        "motion_features": torch.randn(b, 128, T, h, w, generator=g),
        "uncertainty": torch.rand(b * T, 1, h, w, generator=g),
        "sim": torch.rand(b, 1, T, T, generator=g) * 2 - 1,
        "strive_before": torch.ones(b, 1, T, T) + torch.randint(0, 3, (b, 1, T, T), generator=g).float(),
    }


def check_vs_oracle(P, t, top_k=5, rows=None, scale=None):
    b, c, T, h, w = t["query"].shape
    want = O.ppm_memory_read(t["query"], t["key"], O.temporal_pe(T, c), t["value"], t["motion_features"],
                             t["sim"], t["strive_before"], t["uncertainty"], 0.5, scale=scale, top_k=top_k,
                             rows=rows)
    mem, out, mask, strive = run_read(P, t, top_k=top_k, scale=scale)
    assert float((mem.frame_score.cpu() - want["frame_score"]).abs().max()) <= 2e-6
    gap = torch.sort(want["frame_score"], dim=-1, descending=True).values
    K = min(top_k, T)
    if K < T:
        assert float((gap[..., K - 1] - gap[..., K]).min()) > 1e-5           # pick is well-posed
    assert torch.equal(mask, want["mask"])
    assert torch.equal(mem.picked.cpu().long()[..., :K], want["picked"])
    assert torch.equal(strive, want["strive"])
    assert float((mem.score_norm.cpu()[..., :K] - want["score_norm"]).abs().max()) <= 5e-6
    got = out.reshape(b, T, c, h * w)
    ref = want["motion_features_global"].reshape(b, T, c, h * w)
    if rows is not None:
        got, ref = got[..., rows], ref[..., rows]
    hmax = float(want["hidden"].abs().max())
    d = (got - ref).abs()
    # bf16 rounding of P (2^-9 relative, independent in kernel and oracle) and of the output
    assert float(d.max()) <= 3e-2 * hmax + 1e-6, (float(d.max()), hmax)
    assert float(d.mean()) <= 3e-3 * hmax + 1e-7
    return out


def test_ragged_tiles_batch2(P):
    """hw = 13*23 = 299: partial query and key tiles; b = 2 couples score_norm across the batch."""
    check_vs_oracle(P, make_case(2, 8, 13, 23, seed=11))


def test_peaky_softmax_rescale_path(P):
    """key ~ query: sharply peaked rows; a large scale forces running-max growth > 2^8 (lazy rescale)."""
    check_vs_oracle(P, make_case(1, 7, 16, 24, seed=12, peaky=True), scale=1.0)


def test_T_below_topk_picks_everything(P):
    t = make_case(1, 3, 8, 16, seed=13)
    out = check_vs_oracle(P, t)
    assert torch.isfinite(out).all()


@pytest.mark.parametrize("top_k", [4, 8, 16])
def test_config3_sweep_sampled_rows(P, top_k):
    """BASELINE config 3: hw = 92*160 = 14720, T = 20, K in {4, 8, 16}; 192 sampled query rows per frame
    against the oracle (full rows would be 0.4-1.8 TFLOP per frame on the CPU)."""
    t = make_case(1, 20, 92, 160, seed=20 + top_k)
    rows = torch.randperm(92 * 160, generator=torch.Generator().manual_seed(5))[:192]
    check_vs_oracle(P, t, top_k=top_k, rows=rows)


def test_qam_ties_and_strive(P):
    """Exact ties resolve to the lower frame index; strive_time accumulates over iterations."""
    b, T, h, w = 1, 6, 8, 16
    t = make_case(b, T, h, w, seed=14)
    t["sim"] = torch.zeros(b, 1, T, T)
    t["uncertainty"] = torch.full((b * T, 1, h, w), 0.25)
    t["strive_before"] = torch.ones(b, 1, T, T)
    mem = P.PPMMemory(t["query"].cuda(), t["key"].cuda(), None, top_k=5)
    strive = t["strive_before"].clone().cuda()
    for it in range(3):
        mask = mem.pick(t["sim"].cuda(), strive, t["uncertainty"].cuda()).cpu()
        assert torch.equal(mem.picked.cpu()[0, 0], torch.arange(5, dtype=torch.int32))
        assert int(mask.sum()) == T * 5
    assert torch.equal(strive.cpu()[0, 0, 0], torch.tensor([4.0, 4, 4, 4, 4, 1]))


def test_linearity_in_value_full_size(P):
    """Size-independent property at the 1/8-res 720p shape: the read is linear in V (up to bf16)."""
    t = make_case(1, 20, 92, 160, seed=30)
    mem = P.PPMMemory(t["query"].cuda(), t["key"].cuda(), None, top_k=5)
    strive = t["strive_before"].clone().cuda()
    mem.pick(t["sim"].cuda(), strive, t["uncertainty"].cuda())
    zero = torch.zeros_like(t["value"]).cuda()
    one = torch.tensor([1.0], device="cuda")
    v1, v2 = t["value"].cuda(), t["motion_features"].cuda()
    a = mem.attend(v1, zero, one)
    b2 = mem.attend(v2, zero, one)
    ab = mem.attend(v1 + v2, zero, one)
    const = mem.attend(torch.full_like(v1, 0.75), zero, one)
    assert float((const - 0.75).abs().max()) <= 0.75 * 2.0 ** -7              # softmax weights sum to 1
    scale = float(ab.abs().max())
    assert float((a + b2 - ab).abs().max()) <= 6e-2 * scale
    assert float((a + b2 - ab).abs().mean()) <= 3e-3 * scale


# ---------------------------------------------------------------- row a6: QAM frame similarity
@pytest.mark.parametrize("name", ["ppm_read_l16", "ppm_read_l08"])
def test_similarity_vs_reference_golden(P, golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    q, k = torch.from_numpy(g["query"]).cuda(), torch.from_numpy(g["key"]).cuda()
    sim = P.PPMMemory(q, k, None).similarity().cpu()
    ref = torch.from_numpy(g["sim"])                       # captured from the reference's compute_qk_similarity
    assert sim.shape == ref.shape and sim.dtype == torch.float32
    assert float((sim - ref).abs().max()) <= 1e-5


@pytest.mark.parametrize("shape", [(1, 6, 46, 80), (2, 5, 9, 13), (1, 3, 4, 4)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_similarity_vs_oracle_adaptive_windows(P, shape, dtype):
    """h, w not divisible by 4: AdaptiveMaxPool2d windows of uneven size (46 -> 11 rows at the 1/16 level)."""
    b, T, h, w = shape
    g = torch.Generator().manual_seed(h * 100 + w)
    q = torch.randn(b, 128, T, h, w, generator=g).to(dtype)
    k = torch.randn(b, 128, T, h, w, generator=g).to(dtype)
    want = O.qk_similarity(q, k)                    # dtype rule: bf16 inputs -> bf16-rounded pooled means
    got = P.PPMMemory(q.cuda(), k.cuda(), None).similarity().cpu()
    # (bf16: a pooled mean at a rounding boundary may flip by one bf16 ulp between the two fp32 summation orders)
    assert float((got - want).abs().max()) <= (1e-5 if dtype == torch.float32 else 2e-4)
    again = P.PPMMemory(q.cuda(), k.cuda(), None).similarity().cpu()
    assert torch.equal(got, again)                        # deterministic: it feeds the top-k pick


@pytest.mark.parametrize("shape,peaky", [((1, 8, 24, 40), False), ((1, 20, 46, 80), False), ((2, 7, 13, 23), True)])
def test_aggregation_vs_installed_flash_attn(P, shape, peaky):
    """The reference's attention IS flash_attn.flash_attn_func (models/core/ppmstereo.py:550).  The image carries
    flash_attn 2.8.3 with sm_100 code, so on the GPU box the real library is the checker: same bf16 q / K' / V
    (the (b, K*hw, 1, c) layout of :547-548), same scale, compare `hidden`.  Skipped where flash_attn is absent."""
    fa = pytest.importorskip("flash_attn")
    b, T, h, w = shape
    hw, c, K = h * w, 128, 5
    t = make_case(b, T, h, w, seed=77 + T, peaky=peaky)
    mem, out, mask, strive = run_read(P, t, beta=1.0)
    scale = P.softmax_scale(c)
    mf = t["motion_features"].permute(0, 2, 1, 3, 4).reshape(b, T, c, hw)
    mine = out.reshape(b, T, c, hw) - mf                                   # beta = 1 -> hidden (fp32 of bf16 + fp32 add)
    q_p = mem.q_packed.reshape(b, T, hw, c)
    k_p = mem.k_packed.reshape(b, T, K, hw, c)
    hw_pad = (hw + 7) // 8 * 8
    v_p = mem.v_packed.reshape(b, T, c, hw_pad)[..., :hw]
    picked = mem.picked.long()
    worst, hmax, tot, n = 0.0, 0.0, 0.0, 0
    for bb in range(b):
        for i in range(T):
            q = q_p[bb, i].reshape(1, hw, 1, c)
            k = k_p[bb, i].reshape(1, K * hw, 1, c)
            v = v_p[bb, picked[bb, i]].permute(0, 2, 1).reshape(1, K * hw, 1, c).contiguous()
            ref = fa.flash_attn_func(q, k, v, 0.0, softmax_scale=scale, causal=False)   # (1, hw, 1, c) bf16
            ref = ref.float().reshape(hw, c).t().cpu()
            d = (mine[bb, i] - ref).abs()
            worst, hmax = max(worst, float(d.max())), max(hmax, float(ref.abs().max()))
            tot, n = tot + float(d.sum()), n + d.numel()
    # both sides round P and the output to bf16 (2^-9 relative each); the residual add is fp32
    assert worst <= 2e-2 * hmax + 1e-6, (worst, hmax)
    assert tot / n <= 2e-3 * hmax, (tot / n, hmax)


def test_oracle_attention_vs_installed_flash_attn():
    """Pins the oracle's restatement of flash_attn (oracle.attention_bf16) against the real library where it can
    run: the GPU box.  Same bf16 operands, soft and peaky score distributions."""
    fa = pytest.importorskip("flash_attn")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    g = torch.Generator().manual_seed(5)
    for peaky in (False, True):
        q = torch.randn(300, 128, generator=g)
        k = torch.cat([q, q, q, q, q])[torch.randperm(1500, generator=g)] + 0.1 * torch.randn(1500, 128, generator=g) \
            if peaky else torch.randn(1500, 128, generator=g)
        v = torch.randn(1500, 128, generator=g)
        scale = O.ppm_softmax_scale(128) * (4.0 if peaky else 1.0)
        want = O.attention_bf16(q, k, v, scale)
        got = fa.flash_attn_func(q.bfloat16().cuda()[None, :, None], k.bfloat16().cuda()[None, :, None],
                                 v.bfloat16().cuda()[None, :, None], 0.0, softmax_scale=scale, causal=False)
        got = got.float().cpu().reshape(300, 128)
        d = (got - want).abs()
        assert float(d.max()) <= 2.0 ** -7 * float(want.abs().max()), (peaky, float(d.max()))
        # P is rounded to bf16 against a different running maximum on each side, so about half of the outputs of a
        # diffuse softmax land on the neighbouring bf16 value: bound the mean at a quarter ulp of the largest output
        assert float(d.mean()) <= 2.0 ** -10 * float(want.abs().max()), (peaky, float(d.mean()))


def test_beta_zero_returns_motion_features_exactly(P):
    """beta is initialised to 0 in the reference (ppmtereo_update.py:647): the read must then be the identity
    on motion_features, bit for bit (mf + 0 * hidden)."""
    t = make_case(1, 6, 12, 20, seed=31)
    mem, out, mask, strive = run_read(P, t, beta=0.0)
    assert torch.equal(out.reshape(1, 6, 128, 12, 20), t["motion_features"].permute(0, 2, 1, 3, 4))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_frame_major_views_as_the_model_passes_them(P, dtype):
    """forward_update_block hands over (b*T, c, h, w) tensors viewed as (b, c, T, h, w) (permute, no copy) and,
    under autocast, bf16 value / motion features: the strided path must equal the contiguous one."""
    b, T, h, w = 2, 6, 9, 14
    t = make_case(b, T, h, w, seed=32)
    cont = {k: t[k].to(dtype).cuda() for k in ("query", "key", "value", "motion_features")}
    view = {k: v.permute(0, 2, 1, 3, 4).reshape(b * T, 128, h, w).clone()
            .reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4) for k, v in cont.items()}
    assert not view["query"].is_contiguous()
    outs = []
    for src in (cont, view):
        mem = P.PPMMemory(src["query"], src["key"], None, top_k=5)
        strive = t["strive_before"].clone().cuda()
        out, _ = mem.read(src["value"], src["motion_features"], t["sim"].cuda(), strive,
                          t["uncertainty"].to(dtype).cuda(), torch.tensor([0.5], device="cuda"))
        outs.append(out)
    assert torch.equal(outs[0], outs[1])
    if dtype == torch.float32:
        want = O.ppm_memory_read(t["query"], t["key"], O.temporal_pe(T, 128), t["value"], t["motion_features"],
                                 t["sim"], t["strive_before"], t["uncertainty"], 0.5)
        d = (outs[0].cpu() - want["motion_features_global"]).abs()
        assert float(d.max()) <= 3e-2 * float(want["hidden"].abs().max())


@pytest.mark.parametrize("shape", [(1, 6, 1, 1), (1, 6, 2, 3), (2, 7, 5, 5), (1, 2, 3, 43)])
def test_tiny_and_odd_pixel_counts(P, shape):
    """hw below one tile, below one 32-key chunk, not a multiple of 8 (the packed V pitch) and hw = 129 (one
    key past a tile): every masking path of the aggregation."""
    b, T, h, w = shape
    check_vs_oracle(P, make_case(b, T, h, w, seed=40 + h * w))


@pytest.mark.parametrize("shape", [(1, 7, 8, 16), (2, 6, 12, 10), (1, 6, 5, 5)])
def test_frame_major_bf16_value_is_used_in_place(P, shape):
    """SURVEY 8f rank 1 (`to_v` -> V pack): a contiguous (b*T, 128, h, w) bf16 conv output, handed over as the
    reference's rearranged (b, c, T, h, w) view, already has the packed V layout when hw % 8 == 0 -- no pack kernel runs,
    the attention kernel reads the tensor in place, and the result is bit-identical to the packed path.  hw % 8 != 0
    (5 x 5) must still go through the pack."""
    b, T, h, w = shape
    t = make_case(b, T, h, w, seed=5 + T)
    conv_out = torch.randn(b * T, 128, h, w, generator=torch.Generator().manual_seed(9)).bfloat16().cuda()   # to_v output
    view = conv_out.reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4)          # "(b t) c h w -> b c t h w"
    beta = torch.tensor([0.5], device="cuda")
    outs = []
    for value in (view, view.contiguous()):                                   # in place / channel-major copy -> pack kernel
        mem = P.PPMMemory(t["query"].cuda(), t["key"].cuda(), None, top_k=5)
        mem.pick(t["sim"].cuda(), t["strive_before"].clone().cuda(), t["uncertainty"].cuda())
        mem.pack_kv(value)
        aliased = mem.v_packed.data_ptr() == conv_out.data_ptr()
        outs.append((mem.aggregate(t["motion_features"].cuda(), beta), aliased))
    assert outs[0][1] == ((h * w) % 8 == 0) and not outs[1][1]
    assert torch.equal(outs[0][0], outs[1][0])


@pytest.mark.parametrize("shape", [(1, 6, 8, 16), (2, 7, 5, 13), (1, 3, 9, 9), (1, 20, 23, 40)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_modulated_key_pack_equals_torch_op_sequence(P, shape, dtype):
    """ppm_pack_key (one staged slab per SOURCE frame, written once per pick) against the reference's literal ops run by
    torch on the GPU: ``(key_j * s + pe_j).to(bf16)`` (ppmstereo.py:541-548) -- bit-identical, incl. ragged 64-pixel
    tiles, batch 2, frames nobody picked, a frame picked by every query, and zero rows for unused slots (T < top_k)."""
    b, T, h, w = shape
    hw, K = h * w, 5
    g = torch.Generator(device="cuda").manual_seed(T * h + w)
    q = torch.randn(b, 128, T, h, w, generator=g, device="cuda").to(dtype)
    k = torch.randn(b, 128, T, h, w, generator=g, device="cuda").to(dtype)
    mem = P.PPMMemory(q, k, None, top_k=K)
    sim = torch.rand(b, 1, T, T, generator=g, device="cuda")
    sim[:, :, :, 0] += 5.0                                                     # frame 0 is in every pick
    mem.pick(sim, torch.ones(b, 1, T, T, device="cuda"), torch.rand(b * T, 1, h, w, generator=g, device="cuda"))
    mem.pack_kv(torch.randn(b, 128, T, h, w, generator=g, device="cuda"))
    torch.cuda.synchronize()
    got = mem.k_packed.view(b, T, K, hw, 128)
    pe = P.temporal_pe_table(T, 128, "cuda")
    picked, score = mem.picked.view(b, T, K).long(), mem.score_norm.view(b, T, K)
    kk = k.float().reshape(b, 128, T, hw)
    for bb in range(b):
        for i in range(T):
            for n in range(K):
                j = int(picked[bb, i, n])
                if j < 0:
                    assert n >= T and not bool(got[bb, i, n].any())
                    continue
                want = (kk[bb, :, j].t() * score[bb, i, n] + pe[j]).to(torch.bfloat16)     # (hw, 128)
                assert torch.equal(got[bb, i, n], want), (bb, i, n, j)
    assert int((picked == 0).sum()) == b * T                                   # frame 0: picked by every query frame


def test_value_operand_at_many_addresses_with_resident_query_and_key(P):
    """Regression: the library keeps the TMA descriptors of recent operands in a 16-slot round-robin cache.  A value
    operand at a NEW address (every to_v output inside a fresh CUDA-graph pool, every allocator miss) with the query
    and key operands still cached used to overwrite the slot the query's (or key's) descriptor had just been taken
    from -- that one attention call then ran with the value's descriptor in the query's place.  40 in-place value
    tensors at 40 addresses, same contents: every result must equal the first bit for bit."""
    b, T, h, w = 1, 6, 8, 16
    t = make_case(b, T, h, w, seed=77)
    beta = torch.tensor([0.5], device="cuda")
    mem = P.PPMMemory(t["query"].cuda(), t["key"].cuda(), None, top_k=5)
    mem.pick(t["sim"].cuda(), t["strive_before"].clone().cuda(), t["uncertainty"].cuda())
    mf = t["motion_features"].cuda()
    src = torch.randn(b * T, 128, h, w, generator=torch.Generator().manual_seed(3)).bfloat16().cuda()
    values = [src.clone() for _ in range(40)]                                   # all alive: 40 distinct addresses
    assert len({v.data_ptr() for v in values}) == 40
    outs = []
    for v in values:
        mem.pack_kv(v.reshape(b, T, 128, h, w).permute(0, 2, 1, 3, 4))
        assert mem.v_packed.data_ptr() == v.data_ptr()                          # used in place: the address is the operand's
        outs.append(mem.aggregate(mf, beta))
    torch.cuda.synchronize()
    bad = [i for i, o in enumerate(outs) if not torch.equal(o, outs[0])]
    assert not bad, f"attention calls {bad} differ from call 0"


# ------------------------------------------------------------------------------------------------------------
# The reference's GPU dtype flow (bf16 autocast, ppmstereo.py:617), pinned against the reference's LITERAL torch op
# sequence executed by torch on the GPU under autocast -- needs no reference code, only torch.  What is pinned:
#   * `uncertainty_score...mean(-1)` of a bf16 map returns bf16 (mean is not an autocast-fp32 op)  -> conf rounded
#   * `q_.mean(dim=1)` of the bf16 pooled tensor returns bf16; F.cosine_similarity up-casts          -> pooled rounded
# ------------------------------------------------------------------------------------------------------------
def _torch_qam_sequence(sim, strive, unc, b, T, top_k=5):
    """ppmstereo.py:505-513 verbatim (names shortened), under the reference's autocast."""
    with torch.autocast("cuda", dtype=torch.bfloat16):
        penality_score = torch.exp(-(strive) / (strive.sum(3, keepdim=True) + T))
        frame_confidence = unc.reshape(b, 1, 1, T, -1).contiguous().mean(-1)
        frame_score = penality_score * sim + frame_confidence
        frame_idx = torch.argsort(frame_score, dim=-1, descending=True)
        mask = torch.zeros_like(frame_score, dtype=torch.bool).scatter_(3, frame_idx[:, :, :, :top_k], True)
    return frame_score, frame_confidence, mask


@pytest.mark.parametrize("T", [6, 10, 20])
def test_qam_pick_vs_torch_autocast_sequence_bf16(P, T):
    """>= 100 seeded cases per T, half of them with frame confidences packed closer than one bf16 ulp (the case
    an unrounded fp32 mean gets wrong): identical masks, frame_score <= 1e-6."""
    b, h, w = 1, 23, 40
    n_cases = n_close = n_checked_rows = 0
    for seed in range(104):
        g = torch.Generator(device="cuda").manual_seed(1000 * T + seed)
        close = seed % 2 == 1
        if close:      # per-frame means differ by < 2^-9 (bf16 spacing in [0.25, 0.5) is 2^-9)
            base = 0.3 + 0.0015 * torch.rand(T, 1, 1, 1, generator=g, device="cuda")
            unc = (base + 0.05 * (torch.rand(T, 1, h, w, generator=g, device="cuda") - 0.5)).bfloat16()
            sim = (torch.rand(b, 1, T, T, generator=g, device="cuda") * 2 - 1) * 1e-3
        else:
            unc = torch.rand(T, 1, h, w, generator=g, device="cuda").bfloat16()
            sim = torch.rand(b, 1, T, T, generator=g, device="cuda") * 2 - 1
        strive = torch.ones(b, 1, T, T, device="cuda") + torch.randint(0, 4, (b, 1, T, T), generator=g, device="cuda").float()
        score, conf, mask = _torch_qam_sequence(sim, strive, unc, b, T)
        assert conf.dtype == torch.bfloat16 and score.dtype == torch.float32         # the premise of the rule
        q = torch.zeros(b, 128, T, h, w, device="cuda", dtype=torch.bfloat16)
        mem = P.PPMMemory(q, q, None, top_k=5)
        s2 = strive.clone()
        got = mem.pick(sim, s2, unc)
        assert float((mem.frame_score - score).abs().max()) <= 1e-6, (T, seed)
        srt = torch.sort(score, dim=-1, descending=True).values
        well_posed = (srt[..., 4] != srt[..., 5])[0, 0]        # exact ties are undefined in the reference (unstable argsort)
        assert torch.equal(got[0, 0][well_posed], mask[0, 0][well_posed]), (T, seed)
        assert torch.equal(s2[0, 0][well_posed], (strive + mask.float())[0, 0][well_posed])
        n_cases += 1
        n_close += int(close)
        n_checked_rows += int(well_posed.sum())
    assert n_cases >= 100 and n_close >= 50 and n_checked_rows >= 0.9 * n_cases * T


def _torch_similarity_sequence(q, k, t):
    """PPMStereo.compute_qk_similarity (ppmstereo.py:397-423) verbatim, under the reference's autocast."""
    import torch.nn as nn
    import torch.nn.functional as F
    from einops import rearrange
    with torch.autocast("cuda", dtype=torch.bfloat16):
        q = rearrange(q, "b c t h w -> (b t) c h w")
        k = rearrange(k, "b c t h w -> (b t) c h w")
        bt, channels, height, width = q.shape
        b = bt // t
        global_max_pool = nn.AdaptiveMaxPool2d((height // 4, width // 4))
        q_, k_ = global_max_pool(q), global_max_pool(k)
        q_, k_ = q_.mean(dim=1).reshape(b, t, -1), k_.mean(dim=1).reshape(b, t, -1)
        assert q_.dtype == torch.bfloat16
        sim = F.cosine_similarity(q_.unsqueeze(1), k_.unsqueeze(2), dim=-1).unsqueeze(1)
    return sim


@pytest.mark.parametrize("shape", [(1, 6, 46, 80), (1, 20, 23, 40), (2, 10, 12, 16), (1, 7, 9, 13)])
def test_similarity_vs_torch_autocast_sequence_bf16(P, shape):
    b, T, h, w = shape
    g = torch.Generator(device="cuda").manual_seed(b * 1000 + T * 10 + h)
    q = torch.randn(b, 128, T, h, w, generator=g, device="cuda").bfloat16()
    k = (0.5 * q.float() + torch.randn(b, 128, T, h, w, generator=g, device="cuda")).bfloat16()
    want = _torch_similarity_sequence(q, k, T)
    assert want.dtype == torch.float32
    got = P.PPMMemory(q, k, None, top_k=5).similarity()
    # pooled means are bf16 on both sides; a differing fp32 summation order can flip one of them by one bf16 ulp,
    # which moves a cosine over >= 12 pooled positions by < 1e-4
    assert float((got - want).abs().max()) <= 1e-4
    assert float((got - want).abs().mean()) <= 1e-5
    # and the rule matters: an unrounded fp32 pooled mean is further away than the tolerance allows at this size
    from oracle import ppm_oracle as O
    unrounded = O.qk_similarity(q.float().cpu(), k.float().cpu())
    rounded = O.qk_similarity(q.cpu(), k.cpu())
    assert float((rounded - want.cpu()).abs().max()) <= 1e-4
    print("similarity: |kernel - torch| max %.2e, |fp32-mean oracle - torch| max %.2e" %
          (float((got - want).abs().max()), float((unrounded - want.cpu()).abs().max())))


# ------------------------------------------------------------------------------------------------------------
# The shape the roofline and frames/s are quoted on: 720p 1/4 res, hw = 58 880, T = 20, K = 5
# ------------------------------------------------------------------------------------------------------------
def test_headline_shape_sampled_rows_vs_oracle(P):
    """256 sampled query rows per frame against O.ppm_memory_read(rows=...) on the full (b, 128, 20, 184, 320)
    operands in the autocast dtypes (bf16 q / k / value / uncertainty)."""
    b, T, h, w = 1, 20, 184, 320
    g = torch.Generator(device="cuda").manual_seed(2024)
    rn = lambda *s: torch.randn(*s, generator=g, device="cuda")
    t = {"query": rn(b, 128, T, h, w).bfloat16(), "key": rn(b, 128, T, h, w).bfloat16(),
         "value": rn(b, 128, T, h, w).bfloat16(), "motion_features": rn(b, 128, T, h, w),
         "uncertainty": torch.rand(b * T, 1, h, w, generator=g, device="cuda").bfloat16(),
         "sim": torch.rand(b, 1, T, T, generator=g, device="cuda") * 2 - 1,
         "strive_before": torch.ones(b, 1, T, T, device="cuda")}
    rows = torch.randperm(h * w, generator=torch.Generator().manual_seed(5))[:256].sort().values
    rows = torch.cat([torch.tensor([0, 127, 128, h * w - 129, h * w - 128, h * w - 1]), rows]).unique()
    mem = P.PPMMemory(t["query"], t["key"], None, top_k=5)
    strive = t["strive_before"].clone()
    out, mask = mem.read(t["value"], t["motion_features"], t["sim"], strive, t["uncertainty"],
                         torch.tensor([0.5], device="cuda"))
    torch.cuda.synchronize()
    cpu = {k: v.cpu() for k, v in t.items()}
    want = O.ppm_memory_read(cpu["query"], cpu["key"], O.temporal_pe(T, 128), cpu["value"], cpu["motion_features"],
                             cpu["sim"], cpu["strive_before"], cpu["uncertainty"], 0.5, rows=rows)
    assert float((mem.frame_score.cpu() - want["frame_score"]).abs().max()) <= 2e-6
    assert torch.equal(mask.cpu(), want["mask"])
    assert torch.equal(mem.picked.cpu().long(), want["picked"])
    got = out.reshape(b, T, 128, h * w)[..., rows.cuda()].cpu()
    ref = want["motion_features_global"].reshape(b, T, 128, h * w)[..., rows]
    hmax = float(want["hidden"][..., rows].abs().max())
    d = (got - ref).abs()
    assert float(d.max()) <= 3e-2 * hmax + 1e-6, (float(d.max()), hmax)
    assert float(d.mean()) <= 3e-3 * hmax + 1e-7
    # pixels outside `rows` of the oracle output equal motion_features; of ours they must differ (computed)
    assert torch.isfinite(out).all()


def test_headline_shape_one_query_frame_vs_flash_attn(P):
    """One full query frame (Lq = 58 880, Lk = 294 400) against flash_attn_func, the reference's call at
    ppmstereo.py:550, fed the SAME packed bf16 operands."""
    fa = pytest.importorskip("flash_attn")
    b, T, h, w, K = 1, 20, 184, 320, 5
    hw = h * w
    g = torch.Generator(device="cuda").manual_seed(2025)
    rn = lambda *s: torch.randn(*s, generator=g, device="cuda")
    q, k, v = rn(b, 128, T, h, w).bfloat16(), rn(b, 128, T, h, w).bfloat16(), rn(b, 128, T, h, w).bfloat16()
    mf = torch.zeros(b, 128, T, h, w, device="cuda")
    unc = torch.rand(b * T, 1, h, w, generator=g, device="cuda").bfloat16()
    sim = torch.rand(b, 1, T, T, generator=g, device="cuda") * 2 - 1
    mem = P.PPMMemory(q, k, None, top_k=K)
    mem.pick(sim, torch.ones_like(sim), unc)
    out = mem.attend(v, mf, torch.tensor([1.0], device="cuda"))            # mf = 0, beta = 1 -> out = hidden
    for i in (0, 13):
        qi = mem.q_packed.view(T, hw, 128)[i][None, :, None, :]
        ki = mem.k_packed.view(T, K, hw, 128)[i].reshape(1, K * hw, 1, 128)
        picked = mem.picked[0, i].long()
        vi = v[0][:, picked].permute(1, 2, 3, 0).reshape(1, K * hw, 1, 128).contiguous()
        ref = fa.flash_attn_func(qi, ki, vi, dropout_p=0.0, softmax_scale=P.softmax_scale(128), causal=False).float()
        got = out[i].reshape(128, hw).t()
        d = (got - ref[0, :, 0]).abs()
        hmax = float(ref.abs().max())
        assert float(d.max()) <= 2e-2 * hmax, (i, float(d.max()), hmax)     # one bf16 ulp of the output
        assert float(d.mean()) <= 2e-3 * hmax
