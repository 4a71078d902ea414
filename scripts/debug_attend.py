"""GPU debugging aid: PPM read error statistics for P-in-TMEM / P-in-smem and several lazy-rescale thresholds."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import ppmstereo_b200 as P
from ppmstereo_b200 import _lib
from oracle import ppm_oracle as O
from test_ppm_read_gpu import make_case, run_read

lib = _lib.load()
lib.ppm_debug_set_p_in_tmem.argtypes = [ctypes.c_int]
lib.ppm_debug_set_rescale_threshold.argtypes = [ctypes.c_float]

for name, t, scale in (("peaky", make_case(1, 7, 16, 24, seed=12, peaky=True), 1.0),
                       ("ragged", make_case(2, 8, 13, 23, seed=11), None),
                       ("ragged_s0.5", make_case(2, 8, 13, 23, seed=11), 0.5)):
    b, c, T, h, w = t["query"].shape
    want = O.ppm_memory_read(t["query"], t["key"], O.temporal_pe(T, c), t["value"], t["motion_features"],
                             t["sim"], t["strive_before"], t["uncertainty"], 0.5, scale=scale)
    ref = want["motion_features_global"].reshape(b, T, c, h * w)
    for tm in (1, 0):
        for thr in (8.0, 8.0, 8.0, 4.0, 0.0):
            lib.ppm_debug_set_p_in_tmem(tm)
            lib.ppm_debug_set_rescale_threshold(thr)
            mem, out, mask, strive = run_read(P, t, scale=scale)
            d = (out.reshape(b, T, c, h * w) - ref).abs()
            per = d.amax(dim=(0, 2, 3))
            print(f"{name} tmem={tm} thr={thr:g}: max {float(d.max()):.4f} mean {float(d.mean()):.2e} "
                  f"hmax {float(want['hidden'].abs().max()):.3f} per-frame max {[round(float(x), 3) for x in per]}"
                  f" picked0 {mem.picked.cpu()[0].tolist()}")
lib.ppm_debug_set_p_in_tmem(1)
lib.ppm_debug_set_rescale_threshold(8.0)
