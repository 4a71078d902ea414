"""CPU oracle for the PPMStereo hot path (TEST INFRASTRUCTURE ONLY).

This module restates, from the math, what the reference computes on the path
named by BASELINE.json's north_star: correlation pyramid build, radius-r
lookup, and the Pick-and-Play Memory (PPM) read.  It is NOT part of the
product: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  The product
(``ppmstereo_b200``) never imports this file and has no CPU fallback.

Parity pinning: the reference ships no tests or golden vectors,
so this oracle is pinned against outputs of the reference itself, imported
unmodified from /root/reference in the build container by
``tests/golden/make_golden.py``; the resulting fixtures live in
``tests/golden/*.npz`` and ``tests/test_oracle_golden.py`` checks every
function below against them.  The attention arithmetic of the reference lives
in the un-vendored, un-pinned third-party ``flash_attn`` package
(reference models/core/ppmstereo.py:38-41, call site :550); its published
algorithm (softmax(scale * Q K^T) V with bf16 operands, fp32 softmax,
P rounded to bf16 before P.V, bf16 output) is restated in
``attention_bf16``; no reference test pins that function, so it is anchored on
the reference's own call site and, on the GPU box (where the CUDA-only library
can run), checked against the installed flash_attn 2.8.3 by
``tests/test_ppm_read_gpu.py::test_oracle_attention_vs_installed_flash_attn``.

Every function cites the reference file:line it follows (paths relative to
/root/reference).  All math is fp32 on CPU unless a bf16 rounding is stated.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch

# --------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------

def bf16_round(x: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest-even to bfloat16 and come back to fp32."""
    return x.to(torch.bfloat16).to(torch.float32)


def tf32_trunc(x: torch.Tensor) -> torch.Tensor:
    """Drop the 13 low mantissa bits of fp32 (what a TF32 tensor-core read keeps)."""
    return (x.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


# --------------------------------------------------------------------------
# K1: correlation pyramid build   (models/core/corr.py:56-72, 96-104)
# --------------------------------------------------------------------------

def corr_volume(fmap1: torch.Tensor, fmap2: torch.Tensor) -> torch.Tensor:
    """corr[b,h,i,j] = sum_d fmap1[b,d,h,i] * fmap2[b,d,h,j] / sqrt(D).

    Follows CorrBlock1D.corr (models/core/corr.py:96-104): an einsum over the
    channel axis per image row, then a division by sqrt(D).  fp32 throughout.
    Returns (B, H, W1, W2).
    """
    B, D, H, W1 = fmap1.shape
    a = fmap1.float().permute(0, 2, 3, 1)          # B,H,W1,D
    b = fmap2.float().permute(0, 2, 1, 3)          # B,H,D,W2
    vol = torch.matmul(a, b)                       # B,H,W1,W2
    return vol / math.sqrt(D)


def pool_level(level: torch.Tensor) -> torch.Tensor:
    """Average adjacent pairs along the last axis, dropping an odd last column.

    Follows F.avg_pool2d(corr, [1,2], stride=[1,2]) at models/core/corr.py:71.
    """
    w = level.shape[-1] // 2
    return (level[..., 0:2 * w:2] + level[..., 1:2 * w:2]) / 2


def corr_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4,
                 mode: str = "fp32") -> List[torch.Tensor]:
    """The (num_levels+1)-level pyramid of CorrBlock1D.__init__ (corr.py:56-72).

    Each level is returned as (B*H*W1, W2 >> l) fp32.

    mode "fp32": exact fp32 (the reference's CPU path).
    mode "bf16": what the reference's GPU autocast path stores
      (models/core/ppmstereo.py:617 puts the einsum under bf16 autocast):
      inputs rounded to bf16, fp32 accumulation, the einsum result rounded to
      bf16, the division by sqrt(D) rounded to bf16 again, and every pooled
      level rounded to bf16 (cascaded rounding).
    mode "tf32": inputs truncated to TF32, fp32 accumulate, fp32 output.
    """
    B, D, H, W1 = fmap1.shape
    W2 = fmap2.shape[-1]
    if mode == "fp32":
        vol = corr_volume(fmap1, fmap2)
    elif mode == "tf32":
        vol = corr_volume(tf32_trunc(fmap1.float()), tf32_trunc(fmap2.float()))
    elif mode == "bf16":
        a = bf16_round(fmap1.float()).permute(0, 2, 3, 1)
        b = bf16_round(fmap2.float()).permute(0, 2, 1, 3)
        acc = bf16_round(torch.matmul(a, b))
        vol = bf16_round(acc / torch.sqrt(torch.tensor(float(D))))
    else:
        raise ValueError(mode)
    levels = [vol.reshape(B * H * W1, W2)]
    for _ in range(num_levels):
        nxt = pool_level(levels[-1])
        if mode == "bf16":
            nxt = bf16_round(nxt)
        levels.append(nxt)
    return levels


# --------------------------------------------------------------------------
# K2: radius-r lookup   (models/core/corr.py:74-94, bilinear_sampler :10-27)
# --------------------------------------------------------------------------

def corr_lookup(levels: Sequence[torch.Tensor], flow: torch.Tensor, radius: int = 4,
                num_levels: int = 4) -> torch.Tensor:
    """out[b, l*(2r+1) + (k+r), y, x] = lerp of level l's row (b,y,x) at
    (x + flow_x[b,y,x]) / 2**l + k, taps outside [0, W_l-1] contributing zero.

    Closed form of CorrBlock1D.__call__ (corr.py:74-94): grid_sample with
    align_corners=True, zeros padding, H == 1, y pinned to 0.  Only the x
    channel of ``flow`` is used (corr.py:77).  flow: (B,2,H,W1) fp32.
    Returns (B, num_levels*(2r+1), H, W1) fp32.
    """
    B, _, H, W1 = flow.shape
    xs = torch.arange(W1, dtype=torch.float32).view(1, 1, W1)
    center = (xs + flow[:, 0].float()).reshape(-1)            # (B*H*W1,)
    rows = torch.arange(center.numel())
    out = []
    for l in range(num_levels):
        lvl = levels[l].float()
        Wl = lvl.shape[-1]
        c = center / (2 ** l)
        for k in range(-radius, radius + 1):
            x = c + k
            x0 = torch.floor(x)
            a = x - x0
            i0 = x0.long()
            i1 = i0 + 1
            v0 = torch.where((i0 >= 0) & (i0 < Wl), lvl[rows, i0.clamp(0, Wl - 1)], torch.zeros(()))
            v1 = torch.where((i1 >= 0) & (i1 < Wl), lvl[rows, i1.clamp(0, Wl - 1)], torch.zeros(()))
            out.append((1 - a) * v0 + a * v1)
    out = torch.stack(out, 0).reshape(num_levels * (2 * radius + 1), B, H, W1)
    return out.permute(1, 0, 2, 3).contiguous()


# --------------------------------------------------------------------------
# K3: Pick-and-Play Memory read   (models/core/ppmstereo.py:397-423, 443-555)
# --------------------------------------------------------------------------

def temporal_pe(T: int, c: int) -> torch.Tensor:
    """Sinusoidal table PE[t, 2i] = sin(pos_t * w_i), PE[t, 2i+1] = cos(pos_t * w_i),
    pos_t = t / (T-1), w_i = 10000^(-2i/c).  Follows
    get_temporal_positional_encoding(is_normalize=True, scale=1.)
    (models/core/ppmtereo_update.py:25-88; call site ppmstereo.py:453-461).
    Returns (T, c) fp32.  T == 1 gives NaN, as in the reference (0/0).
    """
    pos = torch.arange(T, dtype=torch.float32)
    pos = pos / pos[-1] * 1.0
    w = 1.0 / (10000.0 ** (torch.arange(0, c, 2).float() / c))
    ang = pos[:, None] * w[None, :]
    pe = torch.zeros(T, c)
    pe[:, 0::2] = torch.sin(ang)
    pe[:, 1::2] = torch.cos(ang)
    return pe


def _adaptive_bounds(n_in: int, n_out: int) -> List[Tuple[int, int]]:
    return [((i * n_in) // n_out, -((-(i + 1) * n_in) // n_out)) for i in range(n_out)]


def qk_similarity(query: torch.Tensor, key: torch.Tensor) -> torch.Tensor:
    """sim[b,0,i,j] = cos(k_[b,i], q_[b,j]) where x_ is the channel-mean of the
    adaptive max pool of frame x to (h//4, w//4).  Follows
    PPMStereo.compute_qk_similarity (models/core/ppmstereo.py:397-423), with
    F.cosine_similarity's eps = 1e-8 clamp on each norm.
    query, key: (b, c, T, h, w).  Returns (b, 1, T, T) fp32.

    dtype rule (the reference's GPU path runs under bf16 autocast, ppmstereo.py:617): q/k come out of the
    ``to_qk`` conv as bf16; ``AdaptiveMaxPool2d`` keeps bf16 (exact); ``.mean(dim=1)`` (:409) is NOT an
    autocast-fp32 op, so it returns a bf16 tensor (fp32 accumulation, ONE rounding of the mean);
    ``F.cosine_similarity`` IS an autocast-fp32 op and up-casts those bf16 means.  So: for bf16 inputs the
    pooled channel mean is rounded to bf16 here; for fp32 inputs nothing is rounded.
    """
    b, c, T, h, w = query.shape
    oh, ow = h // 4, w // 4

    def pooled(x):
        x = x.float()
        rows = []
        for (r0, r1) in _adaptive_bounds(h, oh):
            cols = []
            for (c0, c1) in _adaptive_bounds(w, ow):
                cols.append(x[..., r0:r1, c0:c1].amax(dim=(-2, -1)))
            rows.append(torch.stack(cols, -1))
        p = torch.stack(rows, -2)                  # b,c,T,oh,ow
        m = p.mean(dim=1)
        if x_is_bf16:
            m = bf16_round(m)                      # torch.mean of a bf16 tensor returns bf16 (:409)
        return m.reshape(b, T, oh * ow)

    x_is_bf16 = query.dtype == torch.bfloat16
    q_ = pooled(query)
    x_is_bf16 = key.dtype == torch.bfloat16
    k_ = pooled(key)
    dots = torch.einsum("bin,bjn->bij", k_, q_)
    nk = k_.norm(dim=-1).clamp_min(1e-8)
    nq = q_.norm(dim=-1).clamp_min(1e-8)
    return (dots / (nk[:, :, None] * nq[:, None, :])).unsqueeze(1)


def qam_frame_score(sim: torch.Tensor, strive: torch.Tensor,
                    uncertainty: torch.Tensor) -> torch.Tensor:
    """frame_score[b,0,i,j] = exp(-S_ij / (sum_j S_ij + T)) * sim_ij + conf_j with
    conf_j the mean over pixels of frame j's uncertainty map.
    Follows models/core/ppmstereo.py:501-507.
    sim, strive: (b,1,T,T); uncertainty: (b*T, 1, h, w).

    dtype rule: under the reference's bf16 autocast ``uncertainty_score`` (conv -> sigmoid) is bf16 and
    ``.mean(-1)`` (:506) returns a bf16-ROUNDED frame confidence (mean is not an autocast-fp32 op); the add at
    :507 then promotes it back to fp32.  So a bf16 uncertainty map gives conf = bf16(mean); fp32 gives fp32.
    """
    b, _, T, _ = sim.shape
    penalty = torch.exp(-strive / (strive.sum(3, keepdim=True) + T))
    conf = uncertainty.float().reshape(b, 1, 1, T, -1).mean(-1)
    if uncertainty.dtype == torch.bfloat16:
        conf = bf16_round(conf)
    return penalty * sim + conf


def topk_pick(frame_score: torch.Tensor, top_k: int = 5) -> torch.Tensor:
    """Boolean mask (b,1,T,T): per (b, i) row, the top_k largest scores.
    Follows models/core/ppmstereo.py:509-512 (argsort descending, first top_k,
    scatter into a mask).  Ties broken towards the lower frame index (the
    reference's argsort is unstable, so exact ties are undefined there).
    """
    T = frame_score.shape[-1]
    order = torch.sort(frame_score, dim=-1, descending=True, stable=True).indices
    mask = torch.zeros_like(frame_score, dtype=torch.bool)
    mask.scatter_(3, order[..., :min(top_k, T)], True)
    return mask


def attention_bf16(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: float,
                   round_p: bool = True, chunk: int = 8192) -> torch.Tensor:
    """softmax(scale * q k^T) v with FlashAttention-2 rounding: q, k, v rounded
    to bf16, scores/softmax in fp32, P rounded to bf16 before P.v (round_p),
    fp32 accumulation, normalisation by the fp32 row sum, output rounded to
    bf16 and returned as fp32.  Restates flash_attn.flash_attn_func as called
    at models/core/ppmstereo.py:550 (single head, no mask, no dropout).
    q: (Lq, d), k, v: (Lk, d).  Query-chunked so that scores stay small.
    """
    q, k, v = bf16_round(q.float()), bf16_round(k.float()), bf16_round(v.float())
    out = torch.empty(q.shape[0], v.shape[1])
    for s in range(0, q.shape[0], chunk):
        sc = (q[s:s + chunk] @ k.t()) * scale
        m = sc.amax(dim=-1, keepdim=True)
        p = torch.exp(sc - m)
        l = p.sum(-1, keepdim=True)
        if round_p:
            p = bf16_round(p)
        out[s:s + chunk] = (p @ v) / l
    return bf16_round(out)


def ppm_softmax_scale(c: int = 128) -> float:
    """scale = c^-0.5 * log_12000(2c)  (models/core/ppmstereo.py:494; the key
    carries 2c channels after the PE concat at :475)."""
    return c ** -0.5 * math.log(2 * c, 12000)


def ppm_memory_read(query, key, pe, value, motion_features, sim, strive, uncertainty,
                    beta, scale: Optional[float] = None, top_k: int = 5,
                    rows: Optional[torch.Tensor] = None):
    """One iteration of the PPM read (models/core/ppmstereo.py:501-555).

    query, key, value, motion_features: (b, c, T, h, w); pe: (T, c);
    sim, strive: (b,1,T,T) (strive is NOT modified; the new one is returned);
    uncertainty: (b*T,1,h,w); beta: float.
    rows: optional 1-D LongTensor of query pixel indices to evaluate (for
    large shapes); other pixels of the output are left equal to motion_features.

    Returns dict with frame_score, mask, strive (updated), picked (b,T,K) ascending
    int64 (-1 padded), score_norm (b,T,K), hidden (b,T,c,hw) and
    motion_features_global (b*T, c, h, w).
    """
    b, c, T, h, w = query.shape
    hw = h * w
    if scale is None:
        scale = ppm_softmax_scale(c)
    score = qam_frame_score(sim, strive, uncertainty)
    mask = topk_pick(score, top_k)
    strive_new = strive.clone()
    strive_new[mask] += 1                                     # ppmstereo.py:513
    K = min(top_k, T)
    picked = torch.full((b, T, K), -1, dtype=torch.long)
    score_norm = torch.zeros(b, T, K)
    qf = query.float().reshape(b, c, T, hw)
    kf = key.float().reshape(b, c, T, hw)
    vf = value.float().reshape(b, c, T, hw)
    mf = motion_features.float().reshape(b, c, T, hw)
    hidden = torch.zeros(b, T, c, hw)
    out = mf.permute(0, 2, 1, 3).clone()                      # b,T,c,hw
    sel = slice(None) if rows is None else rows
    for i in range(T):
        # picked set in temporal order (masked_select semantics, :524-529)
        idx = [torch.nonzero(mask[bb, 0, i]).flatten() for bb in range(b)]
        sel_scores = torch.stack([score[bb, 0, i, idx[bb]] for bb in range(b)])   # b,K
        norm = sel_scores / sel_scores.mean()                 # mean over b*K (:533)
        for bb in range(b):
            picked[bb, i] = idx[bb]
            score_norm[bb, i] = norm[bb]
            qi = (qf[bb, :, i] + pe[i][:, None]).t()          # hw,c   (:518-522)
            ks, vs = [], []
            for n, j in enumerate(idx[bb].tolist()):
                ks.append((kf[bb, :, j] * norm[bb, n] + pe[j][:, None]).t())   # (:541)
                vs.append(vf[bb, :, j].t())
            kk, vv = torch.cat(ks, 0), torch.cat(vs, 0)       # K*hw, c (:547-548)
            hid = attention_bf16(qi[sel], kk, vv, scale)      # (:550)
            hidden[bb, i][:, sel] = hid.t()
            out[bb, i][:, sel] = mf[bb, :, i][:, sel] + beta * hid.t()         # (:552)
    return {
        "frame_score": score, "mask": mask, "strive": strive_new, "picked": picked,
        "score_norm": score_norm, "hidden": hidden,
        "motion_features_global": out.reshape(b * T, c, h, w),
    }


# ---------------------------------------------------------------------------------------------
# convex upsampling (models/core/ppmstereo.py:185-197)
# ---------------------------------------------------------------------------------------------
def convex_upsample(flow: torch.Tensor, mask: torch.Tensor, rate: int = 4) -> torch.Tensor:
    """out[n,c,rate*y+i,rate*x+j] = sum_k softmax_k(mask[n,(k*rate+i)*rate+j,y,x]) * rate*flow[n,c,y+k//3-1,x+k%3-1],
    zero outside the image.  flow (N,2,H,W); mask (N,9*rate^2,H,W) fp32 or bf16.  A bf16 mask only ever reaches
    the reference inside ``with autocast(...)`` (ppmstereo.py:617), and torch.softmax is on autocast's fp32 list:
    the mask is up-cast, the weights are fp32 and are NOT rounded to bf16 (:188, :193)."""
    N, _, H, W = flow.shape
    w = torch.softmax(mask.float().reshape(N, 9, rate, rate, H, W), dim=1)
    fp = torch.nn.functional.pad(rate * flow.float(), (1, 1, 1, 1))              # zero padding (:190)
    out = torch.zeros(N, 2, rate, rate, H, W)
    for k in range(9):
        ky, kx = divmod(k, 3)
        nb = fp[:, :, ky:ky + H, kx:kx + W]                                       # (N,2,H,W) neighbour k
        out += w[:, k][:, None] * nb[:, :, None, None]
    return out.permute(0, 1, 4, 2, 5, 3).reshape(N, 2, rate * H, rate * W)


# ---------------------------------------------------------------------------------------------
# SKSepConvGRU3D (models/core/ppmtereo_update.py:254-312) -- SURVEY 8f rank 2
# ---------------------------------------------------------------------------------------------
def _sep_conv(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, axis: int, rnd) -> torch.Tensor:
    """1-D convolution of a (b, C, T, H, W) tensor along ``axis`` (2 = T, 3 = H, 4 = W), zero padding (k-1)/2, written
    as a sum of shifted channel contractions: out[p] = bias + sum_d W_d . in[p + d - (k-1)/2]  -- what nn.Conv3d with a
    (k,1,1) / (1,k,1) / (1,1,k) kernel computes (ppmtereo_update.py:257-289).  weight: (Cout, Cin, kt, kh, kw) with a
    single non-unit kernel dim.  ``rnd`` rounds operands and the result (identity in fp32 mode, bf16 in autocast mode:
    conv3d is an autocast-to-bf16 op -- input, weight and bias are cast to bf16, fp32 accumulation, bf16 result, then the
    bias added in bf16: identified against torch/cuDNN on the B200, scripts/exp/gru_rounding_probe.py)."""
    k = weight.shape[axis]
    w = rnd(weight.float()).reshape(weight.shape[0], weight.shape[1], k)            # (Cout, Cin, k)
    xin = rnd(x.float())
    pad = (k - 1) // 2
    n = x.shape[axis]
    out = torch.zeros(x.shape[0], weight.shape[0], *x.shape[2:])
    for d in range(k):
        lo, hi = max(0, pad - d), min(n, n + pad - d)          # output positions whose tap d falls inside the tensor
        if hi <= lo:
            continue
        src = xin.narrow(axis, lo + d - pad, hi - lo)
        out.narrow(axis, lo, hi - lo).add_(torch.einsum("oc,bcthw->bothw", w[:, :, d], src))
    # two roundings in bf16 mode: cuDNN stores the bf16 convolution, the (bf16) bias is added as a separate bf16 op
    return rnd(rnd(out) + rnd(bias.float()).view(1, -1, 1, 1, 1))


def sk_sep_conv_gru3d(h: torch.Tensor, x: torch.Tensor, params: dict, mode: str = "fp32",
                      hidden: str = "bf16") -> torch.Tensor:
    """SKSepConvGRU3D.forward(h, x) (ppmtereo_update.py:291-312).  h (b,128,T,H,W), x (b,384,T,H,W); ``params`` maps
    the module's state-dict names ('convz1.0.weight', ..., 'convq3.bias') to tensors.

    mode "fp32": the reference's CPU arithmetic.
    mode "bf16": the reference's GPU arithmetic under autocast (ppmstereo.py:568): every conv3d runs in bf16 (operands
      and result rounded, fp32 accumulation); gelu / sigmoid / tanh of a bf16 tensor return bf16; the products and sums
      of the gate (r*h, (1-z)*h, z*q, their sum) are bf16 ops when h is bf16 (``hidden="bf16"``; h rounded on entry)
      -- each rounds once.  ``hidden="fp32"`` is what PPMStereo actually hands over (its SST blocks return fp32, so
      tanh(net) is fp32, ppmstereo.py:631, 661): torch's type promotion then makes (1-z)*h and the final sum fp32 ops
      (the state keeps fp32 precision), r*h is an fp32 product that the next conv's autocast cast rounds to bf16, and
      the convs see bf16(h)."""
    rnd = bf16_round if mode == "bf16" else (lambda t: t)
    hrnd = rnd if hidden == "bf16" else (lambda t: t)
    gelu = torch.nn.functional.gelu
    p = params
    h = hrnd(h.float())
    x = x.float()
    for stage, axis in ((1, 4), (2, 3), (3, 2)):                                   # horizontal, vertical, time
        hx = torch.cat([h, x], dim=1)
        if stage == 1:
            z = _sep_conv(rnd(gelu(_sep_conv(hx, p["convz1.0.weight"], p["convz1.0.bias"], axis, rnd))),
                          p["convz1.2.weight"], p["convz1.2.bias"], axis, rnd)
            r = _sep_conv(rnd(gelu(_sep_conv(hx, p["convr1.0.weight"], p["convr1.0.bias"], axis, rnd))),
                          p["convr1.2.weight"], p["convr1.2.bias"], axis, rnd)
        else:
            z = _sep_conv(hx, p[f"convz{stage}.weight"], p[f"convz{stage}.bias"], axis, rnd)
            r = _sep_conv(hx, p[f"convr{stage}.weight"], p[f"convr{stage}.bias"], axis, rnd)
        z, r = rnd(torch.sigmoid(z)), rnd(torch.sigmoid(r))
        rh = rnd(r * h)
        q = rnd(torch.tanh(_sep_conv(torch.cat([rh, x], dim=1), p[f"convq{stage}.weight"], p[f"convq{stage}.bias"],
                                     axis, rnd)))
        h = hrnd(hrnd(rnd(1 - z) * h) + rnd(z * q))
    return h


def conv3d_shifted(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, rnd) -> torch.Tensor:
    """nn.Conv3d(kernel (kt,kh,kw), padding (k-1)/2, stride 1) as a sum over taps of shifted channel contractions:
    out[p] = bias + sum_{it,ih,iw} W[:, :, it, ih, iw] . in[p + (it,ih,iw) - pad], zeros outside.  ``rnd`` as in _sep_conv."""
    co, ci, kt, kh, kw = weight.shape
    w = rnd(weight.float())
    xin = rnd(x.float())
    b, _, T, H, W = x.shape
    xp = torch.nn.functional.pad(xin, ((kw - 1) // 2,) * 2 + ((kh - 1) // 2,) * 2 + ((kt - 1) // 2,) * 2)
    out = torch.zeros(b, co, T, H, W)
    for it in range(kt):
        for ih in range(kh):
            for iw in range(kw):
                out += torch.einsum("oc,bcthw->bothw", w[:, :, it, ih, iw], xp[:, :, it:it + T, ih:ih + H, iw:iw + W])
    return rnd(rnd(out) + rnd(bias.float()).view(1, -1, 1, 1, 1))


def flow_head3d(x: torch.Tensor, params: dict, mode: str = "fp32") -> torch.Tensor:
    """FlowHead3D.forward (models/core/ppmtereo_update.py:670-678): conv2(relu(conv1(x))), two 3x3x3 Conv3d.
    mode as in sk_sep_conv_gru3d (bf16: the autocast regime; each conv = bf16(bf16(acc) + bias))."""
    rnd = bf16_round if mode == "bf16" else (lambda t: t)
    y = torch.relu(conv3d_shifted(x, params["conv1.weight"], params["conv1.bias"], rnd))
    return conv3d_shifted(y, params["conv2.weight"], params["conv2.bias"], rnd)


# ---------------------------------------------------------------------------------------------
# BasicMotionEncoder_v2.convc1 = PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]) + the encoder's outer GELU
# (models/core/ppmtereo_update.py:91-116, :466) -- SURVEY 8f rank 1 (lookup -> convc1 fusion)
# ---------------------------------------------------------------------------------------------
def _conv1x1(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, rnd) -> torch.Tensor:
    """1x1 Conv2d as a channel contraction.  bf16 mode: bf16(bf16(acc) + bf16 bias) -- what cuDNN + torch produce
    (identified on the B200, scripts/exp/pcblock_rounding_probe.py: 99.999 %)."""
    w = rnd(weight.float()).reshape(weight.shape[0], weight.shape[1])
    acc = torch.einsum("oc,bchw->bohw", w, rnd(x.float()))
    return rnd(rnd(acc) + rnd(bias.float()).view(1, -1, 1, 1))


def _depthwise(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, rnd) -> torch.Tensor:
    """Depthwise k x k Conv2d (groups = C, zero padding k//2) as a sum of shifted per-channel products.  bf16 mode:
    ONE rounding, bf16(acc + bf16 bias) (cuDNN's depthwise path fuses the bias; probe: 99.995-100 %)."""
    C, _, k, _ = weight.shape
    w = rnd(weight.float())
    xin = torch.nn.functional.pad(rnd(x.float()), (k // 2,) * 4)
    H, W = x.shape[-2:]
    acc = torch.zeros_like(x, dtype=torch.float32)
    for ky in range(k):
        for kx in range(k):
            acc += w[:, 0, ky, kx].view(1, C, 1, 1) * xin[:, :, ky:ky + H, kx:kx + W]
    return rnd(acc + rnd(bias.float()).view(1, -1, 1, 1))


def motion_corr_front(corr: torch.Tensor, params: dict, mode: str = "fp32") -> torch.Tensor:
    """gelu(PCBlock4_Deep_nopool_res(corr)): the ``cor = F.gelu(self.convc1(corr))`` line of
    BasicMotionEncoder_v2.forward (ppmtereo_update.py:466) with the block's forward (:110-116) written out.
    corr (B, 36, h, w) fp32 (the lookup's output); params: the block's state dict ('ffn1.0.weight', 'conv_list.1.bias',
    'pw.weight', 'ffn2.2.weight', ...).  mode "bf16" = the autocast regime: every conv output is a bf16 tensor (rounding
    rules above), the residual stream ``x + branch`` and its GELUs stay fp32 (fp32 + bf16 promotes to fp32), GELU of a
    bf16 tensor is bf16."""
    rnd = bf16_round if mode == "bf16" else (lambda t: t)
    gelu = torch.nn.functional.gelu
    p = params
    x = corr.float()
    t = _conv1x1(rnd(gelu(_conv1x1(x, p["ffn1.0.weight"], p["ffn1.0.bias"], rnd))), p["ffn1.2.weight"], p["ffn1.2.bias"], rnd)
    x = gelu(x + t)
    for i in range(2):                                           # conv_list: depthwise k = 1, then k = 7
        x = gelu(x + _depthwise(x, p[f"conv_list.{i}.weight"], p[f"conv_list.{i}.bias"], rnd))
    x = gelu(x + _conv1x1(x, p["pw.weight"], p["pw.bias"], rnd))
    u = rnd(gelu(_conv1x1(x, p["ffn2.0.weight"], p["ffn2.0.bias"], rnd)))
    out = _conv1x1(u, p["ffn2.2.weight"], p["ffn2.2.bias"], rnd)
    return rnd(gelu(out))


def motion_encoder(flow: torch.Tensor, corr: torch.Tensor, motion_hidden_state: torch.Tensor, params: dict,
                   mode: str = "fp32") -> Tuple[torch.Tensor, torch.Tensor]:
    """BasicMotionEncoder_v2.forward (models/core/ppmtereo_update.py:458-482) with ``motion_hidden_state`` given
    (init_conv, :462-464, is the caller's on the first iteration).  flow (B,2,h,w), corr (B,36,h,w),
    motion_hidden_state (B,64,h,w); params: the encoder's state dict.  Every Conv2d is written as a sum over its kernel
    offsets of shifted channel contractions (conv3d_shifted with a one-frame time axis); mode "bf16" = the autocast
    regime, each convolution bf16(bf16(acc) + bias).  Returns (motion_features (B,128,h,w) fp32 -- the cat with the fp32
    flow promotes --, new motion_hidden_state (B,64,h,w))."""
    rnd = bf16_round if mode == "bf16" else (lambda t: t)
    relu = torch.relu

    def conv(x, name):
        w, b = params[name + ".weight"], params[name + ".bias"]
        return conv3d_shifted(x.unsqueeze(2), w.unsqueeze(2), b, rnd).squeeze(2)

    pc = {k[len("convc1."):]: v for k, v in params.items() if k.startswith("convc1.")}
    cor = motion_corr_front(corr, pc, mode)
    cor = relu(conv(cor, "convc2"))
    flo = relu(conv(flow.float(), "convf1"))
    flo = relu(conv(flo, "convf2"))
    out = relu(conv(torch.cat([cor, flo, rnd(motion_hidden_state.float())], dim=1), "final_conv"))
    return torch.cat([out[:, :126], flow.float()], dim=1), out[:, 126:190]


def value_uncertainty(motion_features: torch.Tensor, net: torch.Tensor, to_v_weight: torch.Tensor, unc_params: dict,
                      mode: str = "fp32") -> Tuple[torch.Tensor, torch.Tensor]:
    """``value = aggregator.to_v(motion_features)`` (Conv2d(128,128,1,bias=False), ppmtereo_update.py:646, called at
    :948) and ``uncertainty = update_block.uncertainty(cat[net, value])`` (Conv2d(256,128,3) ReLU Conv2d(128,1,1)
    Sigmoid, :889-893; called at models/core/ppmstereo.py:495).  unc_params: the Sequential's state dict ('0.weight',
    '0.bias', '2.weight', '2.bias').  mode "bf16": each conv bf16(bf16(acc) + bias), sigmoid evaluated in fp32 on the bf16
    conv output and rounded once."""
    rnd = bf16_round if mode == "bf16" else (lambda t: t)

    def conv(x, w, b):
        b = torch.zeros(w.shape[0]) if b is None else b
        return conv3d_shifted(x.unsqueeze(2), w.unsqueeze(2), b, rnd).squeeze(2)

    value = conv(motion_features.float(), to_v_weight, None)
    hid = torch.relu(conv(torch.cat([rnd(net.float()), value], dim=1), unc_params["0.weight"], unc_params["0.bias"]))
    return value, rnd(torch.sigmoid(conv(hid, unc_params["2.weight"], unc_params["2.bias"])))
