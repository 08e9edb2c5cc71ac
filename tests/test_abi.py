"""The C-ABI library loads and exports exactly what include/mbots.h declares; argument errors are reported
without a GPU; and the product does not fall back to the CPU when there is no device (CPU only)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mbots.h")
PKG = os.path.join(ROOT, "madrona-bots_b200")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mbots_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_manager_surface():
    syms = declared_symbols()
    for need in ("mbots_create", "mbots_destroy", "mbots_step", "mbots_shift_observations", "mbots_tensor",
                 "mbots_set_action", "mbots_agent_offset_for_world", "mbots_num_agents", "mbots_last_error"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    from madrona_bots import _native
    lib = _native.lib()
    syms = declared_symbols()
    assert sorted(_native.SYMBOLS) == syms, "madrona_bots/_native.py SYMBOLS out of sync with include/mbots.h"
    for s in syms:
        assert hasattr(lib, s), f"libmbots.so does not export {s}"
    out = subprocess.check_output(["nm", "-D", "--defined-only", _native.LIB_PATH], text=True)
    exported = set(re.findall(r"\bT (mbots_[a-z0-9_]+)", out))
    assert set(syms) <= exported
    assert lib.mbots_version().startswith(b"mbots-b200")


def test_export_ids_match_the_header():
    from madrona_bots._native import EXPORT
    src = open(HEADER).read()
    body = src[src.index("enum mbots_export_id"):src.index("enum mbots_dtype")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names, val = {}, -1
    for m in re.finditer(r"\b(MBOTS_(?:EXPORT|DEBUG)_[A-Z_]+)(?:\s*=\s*(\d+))?", body):
        val = int(m.group(2)) if m.group(2) else val + 1
        names[m.group(1)] = val
    for k, v in EXPORT.items():
        key = "MBOTS_" + k if k.startswith("DEBUG_") else "MBOTS_EXPORT_" + k
        assert names[key] == v, key
    assert names["MBOTS_EXPORT_SPECIES_COUNT"] == 21


def test_argument_errors_need_no_gpu():
    from madrona_bots import _native
    lib = _native.lib()
    h = ctypes.c_void_p()
    assert lib.mbots_create_ex(None, ctypes.byref(h)) == -1
    assert b"null" in lib.mbots_last_error()
    cfg = _native.Config(0, 0, 0, 32, 0, 0)
    assert lib.mbots_create_ex(ctypes.byref(cfg), ctypes.byref(h)) == -1
    assert b"num_worlds" in lib.mbots_last_error()
    cfg = _native.Config(0, 4, 0, 64, 32, 0)     # capacity below the initial population
    assert lib.mbots_create_ex(ctypes.byref(cfg), ctypes.byref(h)) == -1
    cfg = _native.Config(0, 4, 0, 32, 4096, 0)   # capacity above one CTA
    assert lib.mbots_create_ex(ctypes.byref(cfg), ctypes.byref(h)) == -1
    assert lib.mbots_step(None, None) == -1
    assert lib.mbots_destroy(None) == -1
    n = ctypes.c_uint32()
    assert lib.mbots_num_agents(None, ctypes.byref(n)) == -1


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from madrona_bots import SimManager
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
        SimManager(0, 4, 0, 32)


def test_product_sources_do_not_touch_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use oracle/."""
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower() or f == "sim_prims.cuh" or f == "sim_kernels.cu" or f == "Makefile", \
                    f"{f} mentions the oracle"
                assert "libmbots_oracle" not in text and "mbo_" not in text, f


def test_viewer_is_declared_unavailable():
    from madrona_bots import ScriptBotsViewer
    with pytest.raises(RuntimeError, match="unavailable"):
        ScriptBotsViewer()
