"""Pick-and-Play Memory read: the block at reference models/core/ppmstereo.py:501-555 as one call.

``PPMMemory`` holds what is constant for one resolution level (packed queries, raw keys, the
temporal PE table, the packed-operand workspace); ``read`` performs one iteration: QAM frame
score -> top-k pick (+ strive_time update) -> key modulation -> softmax aggregation over the
picked frames -> ``motion_features + beta * hidden``.  All arithmetic runs in libppm_sm100.so.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional, Tuple

import torch

from . import _lib

C = 128   # q/k/v channels (reference: dim_head = 128, models/core/ppmstereo.py:82-85)


def temporal_pe_table(T: int, c: int, device) -> torch.Tensor:
    """(T, c) table equal to get_temporal_positional_encoding(T, c, is_normalize=True, scale=1.)
    (models/core/ppmtereo_update.py:25-88) without the (b,c,T,h,w) broadcast copy."""
    pos = torch.arange(T, device=device)
    pos = pos / pos[-1] * 1.0
    div = 1.0 / (10000.0 ** (torch.arange(0, c, 2, device=device).float() / c))
    ang = pos.unsqueeze(1) * div
    pe = torch.zeros(T, c, device=device)
    pe[:, 0::2] = torch.sin(ang)
    pe[:, 1::2] = torch.cos(ang)
    return pe


def softmax_scale(c: int = C) -> float:
    """models/core/ppmstereo.py:494 with key.shape[1] == 2c after the PE concat (:475)."""
    return c ** -0.5 * math.log(2 * c, 12000)


def _bcthw(t: torch.Tensor, name: str) -> Tuple[torch.Tensor, int, int, int]:
    """Return (tensor, sb, sc, st) for a (b, c, T, h, w) tensor whose pixels are contiguous."""
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (ppmstereo_b200 has no CPU path)")
    if t.dim() != 5 or t.shape[1] != C:
        raise RuntimeError(f"{name} must be (b, {C}, T, h, w), got {tuple(t.shape)}")
    if t.dtype not in (torch.float32, torch.bfloat16):
        t = t.float()
    w = t.shape[4]
    if t.stride(4) != 1 or t.stride(3) != w:
        t = t.contiguous()
    return t, t.stride(0), t.stride(1), t.stride(2)


class PPMMemory:
    def __init__(self, query: torch.Tensor, key: torch.Tensor, pe: Optional[torch.Tensor] = None,
                 top_k: int = 5):
        self.lib = _lib.load()
        self.query, *self.q_strides = _bcthw(query, "query")
        self.key, *self.k_strides = _bcthw(key, "key")
        b, _, T, h, w = self.query.shape
        self.b, self.T, self.h, self.w, self.hw = b, T, h, w, h * w
        self.top_k = top_k
        dev = query.device
        self.pe = (temporal_pe_table(T, C, dev) if pe is None else pe.to(dev).float()).contiguous()
        if self.pe.shape != (T, C):
            raise RuntimeError(f"pe must be ({T}, {C}), got {tuple(self.pe.shape)}")
        qb, kb, vb = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
        _lib.check(self.lib.ppm_pack_bytes(b, T, top_k, self.hw, ctypes.byref(qb), ctypes.byref(kb),
                                           ctypes.byref(vb)), "ppm_pack_bytes")
        self.q_packed = torch.empty(qb.value // 2, dtype=torch.bfloat16, device=dev)
        self.k_packed = torch.empty(kb.value // 2, dtype=torch.bfloat16, device=dev)
        self._v_buf = torch.empty(vb.value // 2, dtype=torch.bfloat16, device=dev)
        self.v_packed = self._v_buf
        self.frame_score = torch.empty(b, 1, T, T, dtype=torch.float32, device=dev)
        self.picked = torch.empty(b, T, top_k, dtype=torch.int32, device=dev)
        self.score_norm = torch.empty(b, T, top_k, dtype=torch.float32, device=dev)
        self.mask = torch.empty(b, 1, T, T, dtype=torch.uint8, device=dev)
        self.conf = torch.empty(b * T, dtype=torch.float32, device=dev)
        _lib.check(self.lib.ppm_pack_query(self.query.data_ptr(), _lib.dtype_code(self.query), *self.q_strides,
                                           self.pe.data_ptr(), b, T, self.hw, self.q_packed.data_ptr(),
                                           _lib.stream_ptr()), "ppm_pack_query")

    def similarity(self) -> torch.Tensor:
        """QAM frame similarity of this level's raw query / key (PPMStereo.compute_qk_similarity,
        models/core/ppmstereo.py:397-423): (b, 1, T, T) fp32, sim[b,0,i,j] = cos(k_i, q_j)."""
        b, T, h, w = self.b, self.T, self.h, self.w
        if self.query.dtype != self.key.dtype or self.q_strides != self.k_strides:
            raise RuntimeError("query and key must share dtype and layout")
        ws = torch.empty(2 * b * T * (h // 4) * (w // 4), dtype=torch.float32, device=self.query.device)
        sim = torch.empty(b, 1, T, T, dtype=torch.float32, device=self.query.device)
        _lib.check(self.lib.ppm_qk_similarity(self.query.data_ptr(), self.key.data_ptr(), _lib.dtype_code(self.query),
                                              *self.q_strides, b, T, h, w, ws.data_ptr(), sim.data_ptr(),
                                              _lib.stream_ptr()), "ppm_qk_similarity")
        return sim

    def pick(self, sim: torch.Tensor, strive: torch.Tensor, uncertainty: torch.Tensor) -> torch.Tensor:
        """QAM score + top-k; updates ``strive`` in place; returns the bool mask (b,1,T,T)."""
        b, T = self.b, self.T
        if tuple(sim.shape) != (b, 1, T, T) or tuple(strive.shape) != (b, 1, T, T):
            raise RuntimeError("sim_score / strive_time must be (b, 1, T, T)")
        if strive.dtype != torch.float32 or not strive.is_contiguous():
            raise RuntimeError("strive_time must be a contiguous float32 tensor (it is updated in place)")
        sim = sim.float().contiguous()
        unc = uncertainty if uncertainty.dtype in (torch.float32, torch.bfloat16) else uncertainty.float()
        unc = unc.contiguous()
        if unc.numel() != b * T * self.hw:
            raise RuntimeError(f"uncertainty_score must hold b*T*h*w = {b * T * self.hw} values")
        _lib.check(self.lib.ppm_qam_pick(sim.data_ptr(), strive.data_ptr(), unc.data_ptr(), _lib.dtype_code(unc),
                                         b, T, self.hw, self.top_k, self.frame_score.data_ptr(),
                                         self.picked.data_ptr(), self.score_norm.data_ptr(),
                                         self.mask.data_ptr(), self.conf.data_ptr(), _lib.stream_ptr()),
                   "ppm_qam_pick")
        return self.mask.bool()

    def pack_kv(self, value: torch.Tensor) -> None:
        """Per-iteration operand prep: modulated keys for the current pick and bf16 values."""
        b, T, hw = self.b, self.T, self.hw
        value, *vs = _bcthw(value, "value")
        if value.shape != self.query.shape:
            raise RuntimeError("value must have the query's (b, c, T, h, w) shape")
        st = _lib.stream_ptr()
        _lib.check(self.lib.ppm_pack_key(self.key.data_ptr(), _lib.dtype_code(self.key), *self.k_strides,
                                         self.pe.data_ptr(), self.picked.data_ptr(), self.score_norm.data_ptr(),
                                         b, T, self.top_k, hw, self.k_packed.data_ptr(), st), "ppm_pack_key")
        # SURVEY 8f rank 1 (`to_v` -> V pack): the packed V layout (frame, 128, round_up(hw, 8)) bf16 IS the NCHW layout
        # of a contiguous (b*T, 128, h, w) bf16 tensor when hw % 8 == 0 -- which is exactly what `aggregator.to_v` writes
        # under autocast (reference ppmtereo_update.py:945-950, rearranged as a view at ppmstereo.py:498).  Then there is
        # nothing to pack: the conv's output is the attention kernel's V operand (value must stay unmodified until
        # aggregate() has been enqueued on the same stream).
        frame_major = (vs[1] == hw and vs[2] == C * hw and (b == 1 or vs[0] == T * C * hw))
        if value.dtype == torch.bfloat16 and hw % 8 == 0 and frame_major and value.data_ptr() % 16 == 0:
            self.v_packed = value.permute(0, 2, 1, 3, 4).reshape(-1)          # a view, no copy
            assert self.v_packed.data_ptr() == value.data_ptr()
        else:
            self.v_packed = self._v_buf
            _lib.check(self.lib.ppm_pack_value(value.data_ptr(), _lib.dtype_code(value), *vs, b, T, hw,
                                               self.v_packed.data_ptr(), st), "ppm_pack_value")

    def aggregate(self, motion_features: torch.Tensor, beta: torch.Tensor,
                  scale: Optional[float] = None, out: Optional[torch.Tensor] = None,
                  band: Optional[Tuple[int, int]] = None) -> torch.Tensor:
        """ppm_memory_attend on the packed operands; returns (b*T, 128, h, w) fp32.
        ``band = (first_pixel, n_pixels)`` restricts the computed/stored queries (row-band split); the
        other pixels of ``out`` are left untouched."""
        b, T, hw = self.b, self.T, self.hw
        mf, *ms = _bcthw(motion_features, "motion_features")
        if mf.shape != self.query.shape:
            raise RuntimeError("motion_features must have the query's (b, c, T, h, w) shape")
        if not torch.is_tensor(beta):
            beta = torch.tensor([float(beta)], device=mf.device)
        beta = beta.detach().float().reshape(-1)[:1].contiguous().to(mf.device)
        scale = softmax_scale(C) if scale is None else float(scale)
        if out is None:
            out = torch.empty(b * T, C, self.h, self.w, dtype=torch.float32, device=mf.device)
        q_begin, q_count = (0, hw) if band is None else (int(band[0]), int(band[1]))
        _lib.check(self.lib.ppm_memory_attend_band(self.q_packed.data_ptr(), self.k_packed.data_ptr(),
                                                   self.v_packed.data_ptr(), self.picked.data_ptr(), mf.data_ptr(),
                                                   _lib.dtype_code(mf), *ms, beta.data_ptr(), scale, b, T,
                                                   self.top_k, hw, q_begin, q_count, out.data_ptr(),
                                                   _lib.stream_ptr()),
                   "ppm_memory_attend_band")
        return out

    def attend(self, value: torch.Tensor, motion_features: torch.Tensor, beta: torch.Tensor,
               scale: Optional[float] = None) -> torch.Tensor:
        """Modulate + aggregate with the current pick; returns (b*T, 128, h, w) fp32."""
        self.pack_kv(value)
        return self.aggregate(motion_features, beta, scale)

    def read(self, value, motion_features, sim, strive, uncertainty, beta, scale=None):
        mask = self.pick(sim, strive, uncertainty)
        return self.attend(value, motion_features, beta, scale), mask


def ppm_memory_read(query, key, pe, value, motion_features, sim_score, strive_time, uncertainty_score,
                    beta, scale: Optional[float] = None, top_k: int = 5):
    """One-shot form (SURVEY.md §8b surface 2).  Returns (motion_features_global, topk_mask);
    ``strive_time`` is updated in place as at models/core/ppmstereo.py:513."""
    mem = PPMMemory(query, key, pe, top_k)
    return mem.read(value, motion_features, sim_score, strive_time, uncertainty_score, beta, scale)
