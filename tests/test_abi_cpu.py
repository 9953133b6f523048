"""CPU suite: the C-ABI library loads without a GPU, exports every symbol include/vktb/backend.h declares
and fails LOUDLY (no CPU fallback) when asked to compute."""
import ctypes
import os
import re

import pytest

from volkit_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vktb", "backend.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vktb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(capi.SIGNATURES.keys())


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(capi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert capi.load().vktb_abi_version() == 1


def test_no_libcuda_link_dependency():
    # cudart is static and libcuda is resolved lazily: the .so must load on a driverless machine
    import subprocess
    out = subprocess.run(["ldd", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "libcuda.so" not in out and "libcudart" not in out


def test_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = capi.load()
    p = ctypes.c_void_p()
    assert lib.vktb_alloc(ctypes.byref(p), 1024) != 0
    assert b"no CPU fallback" in lib.vktb_last_error()
    v = capi.vol(0x1000, (4, 4, 4), capi.UINT8)
    assert lib.vktb_fill_range(ctypes.byref(v), capi.I3(0, 0, 0), capi.I3(4, 4, 4), 0.5) != 0
    with pytest.raises(capi.VktbError):
        capi.check(lib.vktb_sync(), "sync")


def test_slab_planner_matches_oracle_coordinates():
    """vktb_resample_window_src_range is host logic: check it against the coordinate expression."""
    import numpy as np
    lib = capi.load()
    for D, S in ((3072, 2048), (1024, 2048), (24, 16), (100, 37), (7, 50)):
        for n in (1, 2, 3, 4, 8):
            for g in range(n):
                dz0, dz1 = g * D // n, (g + 1) * D // n
                if dz1 <= dz0:
                    continue
                a, b = ctypes.c_int32(), ctypes.c_int32()
                for filt in (0, 1, 2):
                    assert lib.vktb_resample_window_src_range(filt, dz0, dz1, D, S, ctypes.byref(a), ctypes.byref(b)) == 0
                    zs = [int(np.float32(np.float32(z) / np.float32(D)) * np.float32(S)) for z in range(dz0, dz1)]
                    assert a.value == min(zs)
                    extra = (0, 2, 1)[filt]
                    assert b.value == min(max(zs) + 1 + extra, S)
