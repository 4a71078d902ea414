"""SKSepConvGRU3D on the B200: the reference module's ``forward(h, x)``
(models/core/ppmtereo_update.py:291-312) as one C-ABI call (``ppm_sepconv_gru3d``, csrc/sepconv_gru3d.cu).

``SepConvGRU3D(module)`` reads the nine Conv3d layers of a reference ``SKSepConvGRU3D`` instance (state-dict names
unchanged), packs their weights once into the kernel's [tap][Cout][Cin] bf16 layout (re-packed when a parameter
changes) and keeps a workspace per input shape.  bf16 autocast regime only -- the regime ``PPMStereoModel`` hard-codes
(models/ppm_stereo_model.py:26); ``install()`` leaves the module's own forward in place outside it.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Tuple

import torch

from . import _lib

HIDDEN, INPUT = 128, 384
_LAYERS = [("convz1.0", "convr1.0"), ("convz1.2",), ("convr1.2",), ("convq1",), ("convz2", "convr2"), ("convq2",),
           ("convz3", "convr3"), ("convq3",)]


def _layer(module, name):
    obj = module
    for part in name.split("."):
        obj = obj[int(part)] if part.isdigit() else getattr(obj, part)
    return obj


def _pack_weight(conv: torch.nn.Conv3d) -> torch.Tensor:
    w = conv.weight.detach()
    co, ci = w.shape[:2]
    return w.reshape(co, ci, -1).permute(2, 0, 1).to(torch.bfloat16).contiguous()          # (taps, Cout, Cin)


def _strides5(t: torch.Tensor, name: str, channels: int) -> Tuple[torch.Tensor, int, int, int]:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (ppmstereo_b200 has no CPU path)")
    if t.dim() != 5 or t.shape[1] != channels:
        raise RuntimeError(f"{name} must be (b, {channels}, T, H, W), got {tuple(t.shape)}")
    if t.dtype not in (torch.float32, torch.bfloat16):
        t = t.float()
    if t.stride(4) != 1 or t.stride(3) != t.shape[4]:
        t = t.contiguous()
    return t, t.stride(0), t.stride(1), t.stride(2)


class SepConvGRU3D:
    def __init__(self, module):
        self.module = module
        self.lib = _lib.load()
        self._versions = None
        self._weights = self._biases = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def _params(self):
        convs = [[_layer(self.module, n) for n in names] for names in _LAYERS]
        versions = tuple((c.weight._version, c.weight.data_ptr(), c.bias._version, c.bias.data_ptr())
                         for group in convs for c in group)
        if versions != self._versions:
            self._weights = [torch.cat([_pack_weight(c) for c in group], dim=1).contiguous() for group in convs]
            # conv3d under autocast casts the bias to bf16 as well
            self._biases = [torch.cat([c.bias.detach().to(torch.bfloat16).float() for c in group]).contiguous()
                            for group in convs]
            self._wptr = _lib.ptr_array(self._weights)
            self._bptr = _lib.ptr_array(self._biases)
            self._versions = versions
        return self._wptr, self._bptr

    def __call__(self, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        h, hsb, hsc, hst = _strides5(h, "h", HIDDEN)
        x, xsb, xsc, xst = _strides5(x, "x", INPUT)
        b, _, T, H, W = h.shape
        if tuple(x.shape) != (b, INPUT, T, H, W):
            raise RuntimeError(f"x must be ({b}, {INPUT}, {T}, {H}, {W}), got {tuple(x.shape)}")
        wptr, bptr = self._params()
        key = (b, T, H, W, h.device)
        ws = self._ws.get(key)
        if ws is None:
            n = ctypes.c_size_t()
            _lib.check(self.lib.ppm_sepconv_gru3d_workspace(b, T, H, W, ctypes.byref(n)), "ppm_sepconv_gru3d_workspace")
            self._ws.clear()                                     # one shape at a time (levels alternate rarely)
            ws = self._ws[key] = torch.empty(n.value, dtype=torch.uint8, device=h.device)
        out = torch.empty(b, HIDDEN, T, H, W, dtype=h.dtype, device=h.device)     # the reference's type promotion
        _lib.check(self.lib.ppm_sepconv_gru3d(h.data_ptr(), _lib.dtype_code(h), hsb, hsc, hst, x.data_ptr(),
                                              _lib.dtype_code(x), xsb, xsc, xst, wptr, bptr, b, T, H, W, ws.data_ptr(),
                                              ws.numel(), out.data_ptr(), _lib.dtype_code(out), _lib.stream_ptr()),
                   "ppm_sepconv_gru3d")
        _remember_state(out, ws, self.lib)
        return out


_ATTR = "_ppmstereo_b200_gru"


def gru_forward(self, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """Replacement for ``SKSepConvGRU3D.forward`` installed by ``ppmstereo_b200.install``: the sm_100a kernel in the
    reference's bf16 autocast regime on a CUDA device (what PPMStereoModel runs; hidden state fp32 or bf16); the
    module's own forward otherwise (fp32 models, training)."""
    if (h.is_cuda and h.dtype in (torch.bfloat16, torch.float32) and torch.is_autocast_enabled()
            and torch.get_autocast_gpu_dtype() == torch.bfloat16 and not torch.is_grad_enabled()):
        impl = self.__dict__.get(_ATTR)
        if impl is None:
            impl = SepConvGRU3D(self)
            self.__dict__[_ATTR] = impl
        return impl(h, x)
    return type(self)._ppm_reference_forward(self, h, x)


# ---------------------------------------------------------------------------------------------------------------
# FlowHead3D (models/core/ppmtereo_update.py:670-678) on the same kernel
# ---------------------------------------------------------------------------------------------------------------
_LAST_STATE = {}          # set by SepConvGRU3D.__call__: the result tensor and where its channels-last bf16 copy lives


def _remember_state(out: torch.Tensor, ws: torch.Tensor, lib) -> None:
    import weakref
    b, _, T, H, W = out.shape
    off = ctypes.c_size_t()
    _lib.check(lib.ppm_sepconv_gru3d_state_offset(b, T, H, W, ctypes.byref(off)), "ppm_sepconv_gru3d_state_offset")
    _LAST_STATE.update(ref=weakref.ref(out), version=out._version, cl_ptr=ws.data_ptr() + off.value, ws=ws,
                       stream=_lib.stream_ptr())


def _state_channels_last(x: torch.Tensor) -> int:
    """Device address of the bf16 channels-last copy of ``x`` when ``x`` IS the tensor object K5 just returned
    (identity, not address: a later tensor may reuse the address), unmodified since, on the same stream; else 0."""
    s = _LAST_STATE
    if s and s["ref"]() is x and s["version"] == x._version and s["stream"] == _lib.stream_ptr():
        return s["cl_ptr"]
    return 0


class FlowHead3D:
    def __init__(self, module):
        self.module = module
        self.lib = _lib.load()
        self._versions = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def _params(self):
        c1, c2 = self.module.conv1, self.module.conv2
        versions = tuple((c.weight._version, c.weight.data_ptr(), c.bias._version, c.bias.data_ptr()) for c in (c1, c2))
        if versions != self._versions:
            self.w1 = _pack_weight(c1)                                                   # (27, 256, 128)
            w2 = _pack_weight(c2)                                                        # (27, 2, 256)
            # second layer: ONE weight tile whose row tap*2 + c is W2[c, :, tap] (54 rows + zero padding): the kernel
            # contracts the 256 channels once per position for all 27 x 2 (offset, output) pairs, then sums shifted planes
            self.w2 = torch.zeros(1, 64, w2.shape[2], dtype=torch.bfloat16, device=w2.device)
            self.w2[0, :54] = w2.reshape(54, w2.shape[2])
            self.b1 = c1.bias.detach().to(torch.bfloat16).float().contiguous()
            self.b2 = torch.zeros(64, device=w2.device)
            self.b2[:w2.shape[1]] = c2.bias.detach().to(torch.bfloat16).float()
            self.n_out = w2.shape[1]
            self._versions = versions

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        self._params()
        if self.n_out != 2 or tuple(self.w1.shape) != (27, 256, HIDDEN):
            raise RuntimeError("FlowHead3D kernel: expects Conv3d(128, 256, 3) -> Conv3d(256, 2, 3)")
        cl = _state_channels_last(x)
        x, sb, sc, st = _strides5(x, "x", HIDDEN)
        b, _, T, H, W = x.shape
        key = (b, T, H, W, x.device)
        ws = self._ws.get(key)
        if ws is None:
            n = ctypes.c_size_t()
            _lib.check(self.lib.ppm_flow_head3d_workspace(b, T, H, W, ctypes.byref(n)), "ppm_flow_head3d_workspace")
            self._ws.clear()
            ws = self._ws[key] = torch.empty(n.value, dtype=torch.uint8, device=x.device)
        out = torch.empty(b, 2, T, H, W, dtype=torch.bfloat16, device=x.device)      # conv3d is an autocast-to-bf16 op
        _lib.check(self.lib.ppm_flow_head3d(x.data_ptr(), _lib.dtype_code(x), sb, sc, st, cl or None, self.w1.data_ptr(),
                                            self.b1.data_ptr(), self.w2.data_ptr(), self.b2.data_ptr(), b, T, H, W,
                                            ws.data_ptr(), ws.numel(), out.data_ptr(), _lib.PPM_BF16, _lib.stream_ptr()),
                   "ppm_flow_head3d")
        return out


def flow_head_forward(self, x: torch.Tensor) -> torch.Tensor:
    """Replacement for ``FlowHead3D.forward`` installed by ``ppmstereo_b200.install`` (same regime rule as the GRU)."""
    if (x.is_cuda and x.dim() == 5 and x.dtype in (torch.bfloat16, torch.float32) and torch.is_autocast_enabled()
            and torch.get_autocast_gpu_dtype() == torch.bfloat16 and not torch.is_grad_enabled()):
        impl = self.__dict__.get(_ATTR)
        if impl is None:
            impl = FlowHead3D(self)
            self.__dict__[_ATTR] = impl
        return impl(x)
    return type(self)._ppm_reference_forward(self, x)


# ---------------------------------------------------------------------------------------------------------------
# SequenceUpdateBlock3D.forward (models/core/ppmtereo_update.py:971-1003) as one call (K5c)
# ---------------------------------------------------------------------------------------------------------------
class UpdateBlock3D:
    """GRU + flow head + 2-D mask head of one reference ``SequenceUpdateBlock3D`` without update attention."""

    def __init__(self, block):
        self.block = block
        self.lib = _lib.load()
        self.gru = SepConvGRU3D(block.gru)
        self.head = FlowHead3D(block.flow_head)
        self._mask_versions = None
        self._ws: Dict[Tuple, torch.Tensor] = {}

    def supported(self) -> bool:
        blk = self.block
        try:
            m1, m2 = blk.mask_2d[0], blk.mask_2d[2]
            return (getattr(blk, "attention_type", None) is None and not blk.use_convex_3d
                    and tuple(m1.weight.shape) == (256, HIDDEN, 3, 3) and m2.weight.shape[1:] == (256, 1, 1)
                    and m2.weight.shape[0] <= 192 and tuple(blk.flow_head.conv1.weight.shape) == (256, HIDDEN, 3, 3, 3)
                    and tuple(blk.flow_head.conv2.weight.shape) == (2, 256, 3, 3, 3))
        except (AttributeError, IndexError):
            return False

    def _mask_params(self):
        m1, m2 = self.block.mask_2d[0], self.block.mask_2d[2]
        versions = tuple((c.weight._version, c.weight.data_ptr(), c.bias._version, c.bias.data_ptr()) for c in (m1, m2))
        if versions != self._mask_versions:
            self.mw1 = _pack_weight(m1)                                                  # (9, 256, 128)
            w2 = _pack_weight(m2)                                                        # (1, n_mask, 256)
            self.n_mask = w2.shape[1]
            self.mw2 = torch.zeros(1, 192, 256, dtype=torch.bfloat16, device=w2.device)
            self.mw2[:, :self.n_mask] = w2
            self.mb1 = m1.bias.detach().to(torch.bfloat16).float().contiguous()
            self.mb2 = torch.zeros(192, device=w2.device)
            self.mb2[:self.n_mask] = m2.bias.detach().to(torch.bfloat16).float()
            self._mask_versions = versions

    def __call__(self, net, inp, motion_features, motion_features_global, t: int):
        parts = []
        for name, x in (("net", net), ("inp", inp), ("motion_features", motion_features),
                        ("motion_features_global", motion_features_global)):
            if not x.is_cuda or x.dim() != 4 or x.shape[1] != HIDDEN:
                raise RuntimeError(f"{name} must be a CUDA tensor (b*T, {HIDDEN}, h, w), got {tuple(x.shape)}")
            if x.dtype not in (torch.float32, torch.bfloat16):
                x = x.float()
            parts.append(x.contiguous())
        net, xs = parts[0], parts[1:]
        bt, _, H, W = net.shape
        b = bt // t
        if any(tuple(x.shape) != tuple(net.shape) for x in xs):
            raise RuntimeError("net, inp, motion_features, motion_features_global must share their shape")
        wptr, bptr = self.gru._params()
        self.head._params()
        self._mask_params()
        key = (b, t, H, W, net.device)
        ws = self._ws.get(key)
        if ws is None:
            n = ctypes.c_size_t()
            _lib.check(self.lib.ppm_update_block3d_workspace(b, t, H, W, ctypes.byref(n)), "ppm_update_block3d_workspace")
            self._ws.clear()
            ws = self._ws[key] = torch.empty(n.value, dtype=torch.uint8, device=net.device)
        net_out = torch.empty_like(net)
        mask = torch.empty(bt, self.n_mask, H, W, dtype=torch.bfloat16, device=net.device)
        dflow = torch.empty(bt, 2, H, W, dtype=torch.bfloat16, device=net.device)
        xp = _lib.ptr_array(xs)
        xd = (ctypes.c_int * 3)(*[_lib.dtype_code(x) for x in xs])
        _lib.check(self.lib.ppm_update_block3d(net.data_ptr(), _lib.dtype_code(net), xp, xd, wptr, bptr,
                                               self.head.w1.data_ptr(), self.head.b1.data_ptr(), self.head.w2.data_ptr(),
                                               self.head.b2.data_ptr(), self.mw1.data_ptr(), self.mb1.data_ptr(),
                                               self.mw2.data_ptr(), self.mb2.data_ptr(), self.n_mask, b, t, H, W,
                                               ws.data_ptr(), ws.numel(), net_out.data_ptr(), mask.data_ptr(),
                                               dflow.data_ptr(), _lib.stream_ptr()), "ppm_update_block3d")
        return net_out, mask, dflow


def update_block_forward(self, net, inp, motion_features, motion_features_global, t=1):
    """Replacement for ``SequenceUpdateBlock3D.forward`` installed by ``ppmstereo_b200.install``: one call for blocks
    without update attention in the bf16 autocast regime; the block's own forward otherwise (it then still reaches the
    patched GRU / flow head classes)."""
    if (net.is_cuda and net.dtype in (torch.bfloat16, torch.float32) and torch.is_autocast_enabled()
            and torch.get_autocast_gpu_dtype() == torch.bfloat16 and not torch.is_grad_enabled()):
        impl = self.__dict__.get(_ATTR)
        if impl is None:
            impl = UpdateBlock3D(self)
            self.__dict__[_ATTR] = impl
        if impl.supported():
            return impl(net, inp, motion_features, motion_features_global, t)
    return type(self)._ppm_reference_forward(self, net, inp, motion_features, motion_features_global, t=t)
