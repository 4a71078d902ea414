"""Lookup fused into the motion encoder's front end (csrc/corr_lookup.cu, K6).

``MotionCorrFront(pcblock)`` computes ``F.gelu(pcblock(corr))`` -- the first line of the correlation branch of the
reference's ``BasicMotionEncoder_v2.forward`` (models/core/ppmtereo_update.py:466) -- in one C-ABI call, either from a
materialised lookup tensor or, fused, from a ``LookupRequest`` (the correlation block + the flow), in which case the
(B, 36, h, w) lookup result is never written.  ``encoder_forward`` is the replacement ``install()`` puts on the encoder
class: the reference's forward with that one line exchanged.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn.functional as F

from . import _lib

A_FLOATS, B_FLOATS = 4124, 19212          # PPM_MF_A_FLOATS / PPM_MF_B_FLOATS (include/ppm_b200.h)
_ATTR = "_ppmstereo_b200_front"
_ATTR_VU = "_ppmstereo_b200_value_uncertainty"


class LookupRequest:
    """What ``forward_update_block`` hands to the encoder instead of ``corr_fn(flow)`` when the front end is fused."""

    def __init__(self, block, flow: torch.Tensor):
        self.block, self.flow = block, flow

    def materialise(self) -> torch.Tensor:
        return self.block(self.flow)


def _b16(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.bfloat16).float()


class MotionCorrFront:
    def __init__(self, pcblock):
        self.m = pcblock
        self.lib = _lib.load()
        self._versions = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def supported(self) -> bool:
        m = self.m
        try:
            return (tuple(m.ffn1[0].weight.shape) == (54, 36, 1, 1) and tuple(m.ffn2[2].weight.shape) == (256, 54, 1, 1)
                    and [tuple(c.weight.shape) for c in m.conv_list] == [(36, 1, 1, 1), (36, 1, 7, 7)]
                    and tuple(m.pw.weight.shape) == (36, 36, 1, 1))
        except (AttributeError, IndexError):
            return False

    def _params(self):
        m = self.m
        convs = [m.ffn1[0], m.ffn1[2], m.conv_list[0], m.conv_list[1], m.pw, m.ffn2[0], m.ffn2[2]]
        versions = tuple((c.weight._version, c.weight.data_ptr(), c.bias._version, c.bias.data_ptr()) for c in convs)
        if versions != self._versions:
            dev = m.pw.weight.device
            a = torch.zeros(A_FLOATS, device=dev)
            a[0:1944] = _b16(m.ffn1[0].weight).reshape(-1)
            a[1944:1998] = _b16(m.ffn1[0].bias)
            a[2000:4016].view(36, 56)[:, :54] = _b16(m.ffn1[2].weight).reshape(36, 54)
            a[4016:4052] = _b16(m.ffn1[2].bias)
            a[4052:4088] = _b16(m.conv_list[0].weight).reshape(-1)
            a[4088:4124] = _b16(m.conv_list[0].bias)
            b = torch.zeros(B_FLOATS, device=dev)
            b[0:1764] = _b16(m.conv_list[1].weight).reshape(-1)
            b[1764:1800] = _b16(m.conv_list[1].bias)
            b[1800:3096] = _b16(m.pw.weight).reshape(-1)
            b[3096:3132] = _b16(m.pw.bias)
            b[3132:5076] = _b16(m.ffn2[0].weight).reshape(-1)
            b[5076:5130] = _b16(m.ffn2[0].bias)
            b[5132:18956] = _b16(m.ffn2[2].weight).reshape(256, 54).t().contiguous().reshape(-1)
            b[18956:19212] = _b16(m.ffn2[2].bias)
            self.wa, self.wb, self._versions = a, b, versions
        return self.wa, self.wb

    def __call__(self, corr) -> torch.Tensor:
        wa, wb = self._params()
        level0 = flow = corr_ptr = None
        keep = None
        if isinstance(corr, LookupRequest):
            blk = corr.block
            if getattr(blk, "_l0_only", False):
                keep = (blk._level0, corr.flow.detach().float().contiguous())
                level0, flow = keep[0].data_ptr(), keep[1].data_ptr()
                B, _, H, W1 = keep[1].shape
                W2 = blk._w2
            else:
                corr = corr.materialise()
        if level0 is None:
            if not corr.is_cuda or corr.dim() != 4 or corr.shape[1] != 36:
                raise RuntimeError(f"corr must be a CUDA tensor (B, 36, h, w), got {tuple(corr.shape)}")
            keep = corr.float().contiguous()
            corr_ptr = keep.data_ptr()
            B, _, H, W1 = keep.shape
            W2 = 0
        dev = wa.device
        key = (B, H, W1)
        ws = self._ws.get(key)
        if ws is None:
            self._ws.clear()
            ws = self._ws[key] = torch.empty(B * 36 * H * W1, dtype=torch.float32, device=dev)
        out = torch.empty(B, 256, H, W1, dtype=torch.bfloat16, device=dev)
        _lib.check(self.lib.ppm_motion_corr_front(level0, flow, corr_ptr, wa.data_ptr(), wb.data_ptr(), B, H, W1, W2,
                                                  ws.data_ptr(), out.data_ptr(), _lib.stream_ptr()),
                   "ppm_motion_corr_front")
        return out


def _pack3x3(conv, n_pad: int):
    w = conv.weight.detach()
    co, ci = w.shape[:2]
    out = torch.zeros(9, n_pad, ci, dtype=torch.bfloat16, device=w.device)
    out[:, :co] = w.reshape(co, ci, 9).permute(2, 0, 1).to(torch.bfloat16)            # (tap = ih*3 + iw, Cout, Cin)
    b = torch.zeros(n_pad, device=w.device)
    b[:co] = _b16(conv.bias)
    return out.contiguous(), b


class MotionEncoder:
    """``BasicMotionEncoder_v2.forward`` (models/core/ppmtereo_update.py:458-482) as one C-ABI call
    (``ppm_motion_encoder``, K7); init_conv (first iteration only) stays on cuDNN."""

    def __init__(self, encoder):
        self.m = encoder
        self.front = MotionCorrFront(encoder.convc1)
        self.lib = self.front.lib
        self._versions = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def supported(self) -> bool:
        m = self.m
        try:
            return (self.front.supported() and tuple(m.convc2.weight.shape) == (192, 256, 3, 3)
                    and tuple(m.convf2.weight.shape) == (64, 128, 3, 3)
                    and tuple(m.final_conv.weight.shape) == (190, 320, 3, 3)
                    and tuple(m.convf1.weight.shape) == (128, 2, 7, 7))
        except AttributeError:
            return False

    def _params(self):
        m = self.m
        convs = (m.convc1.ffn2[2], m.convf1, m.convc2, m.convf2, m.final_conv)
        versions = tuple((c.weight._version, c.weight.data_ptr(), c.bias._version, c.bias.data_ptr()) for c in convs)
        if versions != self._versions:
            self.wd = torch.zeros(1, 256, 64, dtype=torch.bfloat16, device=m.convf1.weight.device)
            self.wd[0, :, :54] = m.convc1.ffn2[2].weight.detach().reshape(256, 54).to(torch.bfloat16)
            self.wf1 = torch.zeros(1, 128, 128, dtype=torch.bfloat16, device=m.convf1.weight.device)
            self.wf1[0, :, :98] = m.convf1.weight.detach().reshape(128, 98).to(torch.bfloat16)   # column = c*49 + ih*7 + iw
            self.bf1 = _b16(m.convf1.bias).contiguous()
            self.wc2, self.bc2 = _pack3x3(m.convc2, 192)
            self.wf2, self.bf2 = _pack3x3(m.convf2, 64)
            self.wfin, self.bfin = _pack3x3(m.final_conv, 192)
            self._versions = versions

    def __call__(self, flow, corr, motion_hidden_state, inp):
        import ctypes
        m = self.m
        if motion_hidden_state is None:
            motion_hidden_state = m.init_conv(inp)
        wa, wb = self.front._params()
        self._params()
        flow32 = flow.detach().float().contiguous()
        level0 = corr_ptr = None
        W2 = 0
        keep = None
        if isinstance(corr, LookupRequest):
            blk = corr.block
            if getattr(blk, "_l0_only", False) and corr.flow is flow:      # the entry point has ONE flow: lookup and convf1
                keep = blk._level0
                level0, W2 = keep.data_ptr(), blk._w2
            else:
                corr = corr.materialise()
        if level0 is None:
            if not corr.is_cuda or corr.dim() != 4 or corr.shape[1] != 36:
                raise RuntimeError(f"corr must be a CUDA tensor (B, 36, h, w), got {tuple(corr.shape)}")
            keep = corr.float().contiguous()
            corr_ptr = keep.data_ptr()
        B, _, H, W = flow32.shape
        mhs = motion_hidden_state
        if mhs.dtype not in (torch.float32, torch.bfloat16):
            mhs = mhs.float()
        mhs = mhs.contiguous()
        if tuple(mhs.shape) != (B, 64, H, W):
            raise RuntimeError(f"motion_hidden_state must be ({B}, 64, {H}, {W}), got {tuple(mhs.shape)}")
        dev = flow32.device
        key = (B, H, W, dev)
        ws = self._ws.get(key)
        if ws is None:
            n = ctypes.c_size_t()
            _lib.check(self.lib.ppm_motion_encoder_workspace(B, H, W, ctypes.byref(n)), "ppm_motion_encoder_workspace")
            self._ws.clear()
            ws = self._ws[key] = torch.empty(n.value, dtype=torch.uint8, device=dev)
        feat = torch.empty(B, 128, H, W, dtype=torch.float32, device=dev)
        mhs_out = torch.empty(B, 64, H, W, dtype=torch.bfloat16, device=dev)
        _lib.check(self.lib.ppm_motion_encoder(level0, flow32.data_ptr(), corr_ptr, W2,
                                               wa.data_ptr(), wb.data_ptr(), self.wd.data_ptr(), self.wf1.data_ptr(), self.bf1.data_ptr(),
                                               mhs.data_ptr(), _lib.dtype_code(mhs), self.wc2.data_ptr(),
                                               self.bc2.data_ptr(), self.wf2.data_ptr(), self.bf2.data_ptr(),
                                               self.wfin.data_ptr(), self.bfin.data_ptr(), B, H, W, ws.data_ptr(),
                                               ws.numel(), feat.data_ptr(), mhs_out.data_ptr(), _lib.stream_ptr()),
                   "ppm_motion_encoder")
        feat[:, 126:] = flow32
        return feat, mhs_out


class ValueUncertainty:
    """``value = update_block.aggregator.to_v(motion_features)`` and
    ``uncertainty = update_block.uncertainty(torch.cat([net, value], dim=1))`` (models/core/ppmstereo.py:492-495;
    modules at ppmtereo_update.py:646 and :889-893) as one C-ABI call (``ppm_value_uncertainty``, K8)."""

    def __init__(self, update_block):
        self.blk = update_block
        self.lib = _lib.load()
        self._versions = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def supported(self) -> bool:
        try:
            tv, u = self.blk.aggregator.to_v, self.blk.uncertainty
            return (isinstance(tv, torch.nn.Conv2d) and tuple(tv.weight.shape) == (128, 128, 1, 1) and tv.bias is None
                    and len(u) == 4 and tuple(u[0].weight.shape) == (128, 256, 3, 3) and isinstance(u[1], torch.nn.ReLU)
                    and tuple(u[2].weight.shape) == (1, 128, 1, 1) and isinstance(u[3], torch.nn.Sigmoid))
        except (AttributeError, TypeError, IndexError):
            return False

    def _params(self):
        tv, u = self.blk.aggregator.to_v, self.blk.uncertainty
        ps = (tv.weight, u[0].weight, u[0].bias, u[2].weight, u[2].bias)
        versions = tuple((q._version, q.data_ptr()) for q in ps)
        if versions != self._versions:
            dev = tv.weight.device
            self.wv = tv.weight.detach().reshape(1, 128, 128).to(torch.bfloat16).contiguous()
            self.bv = torch.zeros(128, device=dev)
            self.wu1, self.bu1 = _pack3x3(u[0], 128)
            self.wu2 = torch.zeros(1, 64, 128, dtype=torch.bfloat16, device=dev)
            self.wu2[0, 0] = u[2].weight.detach().reshape(128).to(torch.bfloat16)
            self.bu2 = torch.zeros(64, device=dev)
            self.bu2[0] = _b16(u[2].bias)[0]
            self._versions = versions

    def __call__(self, motion_features: torch.Tensor, net: torch.Tensor):
        import ctypes
        ts = []
        for name, x in (("motion_features", motion_features), ("net", net)):
            if not x.is_cuda or x.dim() != 4 or x.shape[1] != 128:
                raise RuntimeError(f"{name} must be a CUDA tensor (B, 128, h, w), got {tuple(x.shape)}")
            if x.dtype not in (torch.float32, torch.bfloat16):
                x = x.float()
            ts.append(x.detach().contiguous())
        mf, net = ts
        if mf.shape != net.shape:
            raise RuntimeError("motion_features and net must share their shape")
        self._params()
        B, _, H, W = mf.shape
        dev = mf.device
        key = (B, H, W, dev)
        ws = self._ws.get(key)
        if ws is None:
            n = ctypes.c_size_t()
            _lib.check(self.lib.ppm_value_uncertainty_workspace(B, H, W, ctypes.byref(n)), "ppm_value_uncertainty_workspace")
            self._ws.clear()
            ws = self._ws[key] = torch.empty(n.value, dtype=torch.uint8, device=dev)
        value = torch.empty(B, 128, H, W, dtype=torch.bfloat16, device=dev)
        unc = torch.empty(B, 1, H, W, dtype=torch.bfloat16, device=dev)
        _lib.check(self.lib.ppm_value_uncertainty(mf.data_ptr(), _lib.dtype_code(mf), net.data_ptr(), _lib.dtype_code(net),
                                                  self.wv.data_ptr(), self.bv.data_ptr(), self.wu1.data_ptr(),
                                                  self.bu1.data_ptr(), self.wu2.data_ptr(), self.bu2.data_ptr(), B, H, W,
                                                  ws.data_ptr(), ws.numel(), value.data_ptr(), unc.data_ptr(),
                                                  _lib.stream_ptr()), "ppm_value_uncertainty")
        return value, unc


def value_and_uncertainty(update_block, flow, corr, motion_hidden_state, inp, net):
    """The head of one update iteration as ``forward_update_block`` runs it (models/core/ppmstereo.py:492-495):
    ``get_motion_and_value`` + ``get_uncertainty``.  With the encoder class installed and in the bf16 autocast regime
    the projection and the uncertainty head are one call (K8); otherwise the block's own methods."""
    vu = update_block.__dict__.get(_ATTR_VU)
    if vu is None:
        vu = update_block.__dict__[_ATTR_VU] = ValueUncertainty(update_block)
    if (hasattr(type(update_block.encoder), "_ppm_reference_forward") and _kernel_regime(net) and vu.supported()
            and net.dtype in (torch.float32, torch.bfloat16)):
        motion_features, motion_hidden_state = update_block.encoder(flow, corr, motion_hidden_state, inp)
        if motion_features.dtype in (torch.float32, torch.bfloat16):
            value, uncertainty = vu(motion_features, net)
            return motion_features, motion_hidden_state, value, uncertainty
        value = update_block.aggregator.to_v(motion_features)
    else:
        motion_features, motion_hidden_state, value = update_block.get_motion_and_value(flow, corr, motion_hidden_state, inp)
    return motion_features, motion_hidden_state, value, update_block.get_uncertainty(torch.cat([net, value], dim=1))


def _kernel_regime(t: torch.Tensor) -> bool:
    return (t.is_cuda and torch.is_autocast_enabled() and torch.get_autocast_gpu_dtype() == torch.bfloat16
            and not torch.is_grad_enabled())


def encoder_forward(self, flow, corr, motion_hidden_state, inp):
    """Replacement for ``BasicMotionEncoder_v2.forward`` (models/core/ppmtereo_update.py:458-482): one ``ppm_motion_encoder``
    call (K7) in the bf16 autocast regime with an fp32 flow; otherwise the reference's lines with
    ``F.gelu(self.convc1(corr))`` fused (K6) where that applies.  ``corr`` may be a LookupRequest."""
    enc = self.__dict__.get(_ATTR)
    if enc is None:
        enc = self.__dict__[_ATTR] = MotionEncoder(self)
    regime = _kernel_regime(flow)
    if regime and flow.dtype == torch.float32 and enc.supported():
        return enc(flow, corr, motion_hidden_state, inp)
    front = enc.front
    if motion_hidden_state is None:
        motion_hidden_state = self.init_conv(inp)
    if regime and front.supported():
        cor = front(corr)
    else:
        if isinstance(corr, LookupRequest):
            corr = corr.materialise()
        cor = F.gelu(self.convc1(corr))
    cor = F.relu(self.convc2(cor))
    flo = F.relu(self.convf1(flow))
    flo = F.relu(self.convf2(flo))
    cor_flo = torch.cat([cor, flo, motion_hidden_state], dim=1)
    out = F.relu(self.final_conv(cor_flo))
    out, motion_hidden_state = torch.split(out, [126, 64], dim=1)
    return torch.cat([out, flow], dim=1), motion_hidden_state
