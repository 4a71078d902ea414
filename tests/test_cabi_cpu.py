"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
that include/ppm_b200.h declares, argument errors are reported without a GPU, and the host
logic mirrors the reference.  No compute calls."""
import ctypes
import os
import re

import pytest
import torch

import ppmstereo_b200 as P
from ppmstereo_b200 import _lib, build
from oracle import ppm_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "ppm_b200.h")).read()
    declared = set(re.findall(r"^PPM_API (?:int|const char\*)\s+(ppm_\w+)\s*\(", header, flags=re.M))
    assert len(declared) >= 10
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ppm_abi_version() == 1


def test_exports_nothing_but_the_declared_symbols(lib):
    """The converse: the dynamic symbol table of the library holds exactly the header's functions -- no debug
    hooks, no C++ internals (-fvisibility=hidden) -- and the library reads no environment variable."""
    import subprocess
    header = open(os.path.join(ROOT, "include", "ppm_b200.h")).read()
    declared = set(re.findall(r"^PPM_API (?:int|const char\*)\s+(ppm_\w+)\s*\(", header, flags=re.M))
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-2] in ("T", "t", "W", "V", "D", "B")}
    assert exported == declared, exported ^ declared
    # (libcudart_static itself imports getenv, so the import table cannot be used): no switch name in the binary,
    # and every getenv in the sources sits behind PPM_EXPERIMENTS
    blob = open(_lib.LIB_PATH, "rb").read()
    assert b"PPM_CB_" not in blob and b"ppm_debug" not in blob
    csrc = os.path.join(ROOT, "ppmstereo_b200", "csrc")
    for fn in os.listdir(csrc):
        src = open(os.path.join(csrc, fn)).read()
        for m in re.finditer(r"\bgetenv\(", src):
            guard = src.rfind("#ifdef PPM_EXPERIMENTS", 0, m.start())
            assert guard >= 0 and "#else" not in src[guard:m.start()], (fn, "getenv outside PPM_EXPERIMENTS")


def test_argument_errors_are_reported(lib):
    rc = lib.ppm_corr_build(None, None, 0, 1, 1, 1, 8, 8, 8, 8, None, 4, 0, None)
    assert rc == 1 and b"null" in lib.ppm_last_error()
    buf = ctypes.create_string_buffer(64)
    arr = (ctypes.c_void_p * 5)(*[ctypes.addressof(buf)] * 5)
    rc = lib.ppm_corr_build(ctypes.addressof(buf), ctypes.addressof(buf), 1, 1, 8, 1, 8, 8, 8, 8, arr, 4, 1, None)
    assert rc == 1 and b"too small" in lib.ppm_last_error()      # reference: W/16 >= 1 (corr.py:71)
    rc = lib.ppm_corr_lookup(arr, 0, ctypes.addressof(buf), 1, 1, 8, 16, 5, 4, ctypes.addressof(buf), None)
    assert rc == 3 and b"radius" in lib.ppm_last_error()
    rc = lib.ppm_qam_pick(*[ctypes.addressof(buf)] * 3, 0, 1, 65, 4, 5, *[ctypes.addressof(buf)] * 5, None)
    assert rc == 3 and b"T = 65" in lib.ppm_last_error()
    a = ctypes.addressof(buf)
    rc = lib.ppm_memory_attend_band(a, a, a, a, a, 0, 1, 1, 1, a, 0.05, 1, 2, 2, 100, 90, 20, a, None)
    assert rc == 1 and b"query band" in lib.ppm_last_error()      # [90, 110) outside [0, 100)
    rc = lib.ppm_convex_upsample(a, a, 0, 1, 4, 4, 8, a, None)
    assert rc == 3 and b"rate" in lib.ppm_last_error()            # only the reference's rate = 4
    rc = lib.ppm_qk_similarity(a, a, 0, 1, 1, 1, 1, 2, 3, 8, a, a, None)
    assert rc == 1 and b">= 4" in lib.ppm_last_error()            # h // 4 == 0: the reference's pool would be empty
    q, k, v = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    assert lib.ppm_pack_bytes(1, 20, 5, 3680, ctypes.byref(q), ctypes.byref(k), ctypes.byref(v)) == 0
    assert (q.value, k.value, v.value) == (20 * 3680 * 256, 100 * 3680 * 256, 20 * 128 * 3680 * 2)
    assert lib.ppm_pack_bytes(1, 1, 1, 5, ctypes.byref(q), None, ctypes.byref(v)) == 0 and v.value == 128 * 8 * 2


def test_update_iteration_entries_check_arguments_and_size_their_workspaces(lib):
    """K5 / K5c / K7 / K8: workspace formulas (host-only functions) and the checks that fire before any CUDA call."""
    n = ctypes.c_size_t()
    P = 2 * 3 * 5 * 7
    assert lib.ppm_sepconv_gru3d_workspace(2, 3, 5, 7, ctypes.byref(n)) == 0
    assert n.value == P * ((384 + 128 + 128 + 128 + 256) * 2 + 128 * 4)
    assert lib.ppm_update_block3d_workspace(2, 3, 5, 7, ctypes.byref(n)) == 0 and n.value > P * 2560
    assert lib.ppm_motion_encoder_workspace(6, 5, 7, ctypes.byref(n)) == 0
    assert n.value == 6 * 5 * 7 * (36 * 4 + (256 + 128 + 320) * 2)
    assert lib.ppm_value_uncertainty_workspace(6, 5, 7, ctypes.byref(n)) == 0 and n.value == 6 * 5 * 7 * 4 * 128 * 2
    for fn in (lib.ppm_motion_encoder_workspace, lib.ppm_value_uncertainty_workspace):
        assert fn(0, 5, 7, ctypes.byref(n)) == 1 and b"bad argument" in lib.ppm_last_error()
        assert fn(1, 5, 7, None) == 1
    buf = ctypes.create_string_buffer(4096)
    a = (ctypes.addressof(buf) + 255) & ~255
    # motion encoder: null weight, then a workspace that is too small
    rc = lib.ppm_motion_encoder(None, a, a, 0, a, a, a, None, a, a, 0, a, a, a, a, a, a, 1, 4, 4, a, 1 << 20, a, a, None)
    assert rc == 1 and b"null" in lib.ppm_last_error()
    rc = lib.ppm_motion_encoder(None, a, a, 0, a, a, a, a, a, a, 0, a, a, a, a, a, a, 1, 4, 4, a, 16, a, a, None)
    assert rc == 1 and b"workspace of 16 bytes" in lib.ppm_last_error()
    rc = lib.ppm_motion_encoder(None, a, a, 0, a, a, a, a, a, a, 7, a, a, a, a, a, a, 1, 4, 4, a, 1 << 20, a, a, None)
    assert rc == 1 and b"dtype" in lib.ppm_last_error()
    # value + uncertainty: bad dtype, misaligned workspace
    rc = lib.ppm_value_uncertainty(a, 5, a, 0, a, a, a, a, a, a, 1, 4, 4, a, 1 << 20, a, a, None)
    assert rc == 1 and b"dtype" in lib.ppm_last_error()
    rc = lib.ppm_value_uncertainty(a, 0, a, 0, a, a, a, a, a, a, 1, 4, 4, a + 8, 1 << 20, a, a, None)
    assert rc == 1 and b"256-byte aligned" in lib.ppm_last_error()


def test_cpu_tensors_are_rejected():
    f = torch.zeros(1, 8, 2, 16)
    with pytest.raises(RuntimeError, match="CUDA"):
        P.CorrBlock1D(f, f)
    with pytest.raises(RuntimeError, match="CUDA"):
        P.PPMMemory(torch.zeros(1, 128, 2, 2, 2), torch.zeros(1, 128, 2, 2, 2))


def test_host_helpers_match_oracle():
    assert torch.allclose(P.temporal_pe_table(20, 128, "cpu"), O.temporal_pe(20, 128), atol=1e-7)
    assert abs(P.softmax_scale(128) - O.ppm_softmax_scale(128)) < 1e-15
    assert abs(P.softmax_scale(128) - 0.05218212637175694) < 1e-15


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "ppmstereo_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("no oracle", ""), fn
