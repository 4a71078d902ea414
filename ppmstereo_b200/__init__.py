"""ppmstereo_b200 — B200 (sm_100a) drop-in for PPMStereo's correlation + PPM-memory hot path.

Public surface (mirrors the reference's call sites):
  CorrBlock1D            models/core/corr.py:55
  PPMMemory / ppm_memory_read   the block at models/core/ppmstereo.py:501-555
  install(ppmstereo_module)     rewires a loaded reference model onto the two above
  convex_upsample               PPMStereo.convex_upsample (models/core/ppmstereo.py:185-197) as one kernel
  StagedRunner                  overlapped host<->device staging of independent windows / levels
  forward_batch_test            PPMStereo.forward_batch_test (models/core/ppmstereo.py:238-320): same windows and
                                stitching, per-frame encoder cache, staged copies (window.py)
There is no CPU or PyTorch fallback: every op fails loudly without libppm_sm100.so + CUDA.
"""
from .corr import CorrBlock1D, build_pyramid, lookup, lookup_l0
from .memory import PPMMemory, ppm_memory_read, softmax_scale, temporal_pe_table
from .install import install, uninstall, forward_update_block
from .pipeline import StagedRunner
from . import shard
from .upsample import convex_upsample
from .gru import FlowHead3D, SepConvGRU3D, UpdateBlock3D
from .motion import LookupRequest, MotionCorrFront, MotionEncoder, ValueUncertainty
from .rowband import RowBandMemory, gather_bands, row_bands
from .window import FrameFeatureCache, forward_batch_test
from . import window
from . import rowband

__all__ = ["CorrBlock1D", "build_pyramid", "lookup", "lookup_l0", "PPMMemory", "ppm_memory_read", "softmax_scale",
           "temporal_pe_table", "install", "uninstall", "forward_update_block", "StagedRunner", "shard", "convex_upsample", "SepConvGRU3D", "FlowHead3D", "UpdateBlock3D", "MotionCorrFront", "MotionEncoder", "ValueUncertainty", "LookupRequest", "RowBandMemory", "gather_bands", "row_bands",
           "FrameFeatureCache", "forward_batch_test", "window"]
