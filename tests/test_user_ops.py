"""User-supplied Transform ops under the GPU policy (the reference's contract is an arbitrary
TransformUnaryOp / TransformBinaryOp, include/cpp/vkt/Transform.hpp:16-25).

tests/user_ops/ops.cu is ONE source: g++ turns it into host functions, which the UNMODIFIED reference runs through
vkt::TransformRange (oracle/_ref, ref_transform_fn1/2) -- that is the oracle; nvcc (and NVRTC) turn it into sm_100a
kernels registered with vktb_register_*_op_image / _source and run by vktb_transform_range1/2.  Bit-exact."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UDIR = os.path.join(ROOT, "tests", "user_ops")
HOST_SO = os.path.join(UDIR, "build", "libuser_ops_host.so")
CUBIN = os.path.join(UDIR, "build", "user_ops.cubin")


def build_ops():
    subprocess.run(["make", "-C", UDIR], check=True, stdout=subprocess.DEVNULL)
    return ctypes.CDLL(HOST_SO)


def fn_addr(lib, name):
    return ctypes.cast(getattr(lib, name), ctypes.c_void_p).value


def test_one_source_builds_for_host_and_device():
    """the header is self-contained: g++ gives the three host functions, nvcc (cross-compiling) the three kernels"""
    lib = build_ops()
    for name in ("checkered3", "ramp16", "sort_pair"):
        assert fn_addr(lib, name)
    elf = subprocess.run(["cuobjdump", "-elf", CUBIN], capture_output=True, text=True).stdout
    for k in ("vktb_unary_checkered3", "vktb_unary_ramp16", "vktb_binary_sort_pair"):
        assert k in elf
    usage = subprocess.run(["cuobjdump", "-res-usage", CUBIN], capture_output=True, text=True).stdout
    assert "STACK:0" in usage and "LOCAL:0" in usage          # the op is inlined: the 8-byte scratch lives in registers


@pytest.mark.needs_ref
def test_host_twin_runs_in_the_unmodified_reference():
    """checkered3 restates the reference example's MakeCheckered<3> (src/examples/Arithmetic.cpp:8-21): on the 32^3
    UInt8 volume of that example the reference, running the host twin, produces the 8^3 checkerboard of 0 / 128"""
    lib = build_ops()
    v = po.Vol(np.full((32, 32, 32), 7, np.uint8), po.UINT8)
    po.reference().transform_fn1(v, (0, 0, 0), (32, 32, 32), fn_addr(lib, "checkered3"))
    z, y, x = np.indices((32, 32, 32)) >> 3
    lin = z * 1024 + y * 32 + x
    want = np.where(((y % 2 == z % 2) & (lin % 2 == 0)) | ((y % 2 != z % 2) & (lin % 2 == 1)), 128, 0).astype(np.uint8)
    assert v.a.tobytes() == want.tobytes()


@pytest.mark.needs_ref
def test_restatement_with_a_user_op_is_pinned_on_the_reference():
    """oracle/vkt_oracle.c: vko_transform_fn1 == the reference's TransformRange for the same host function, and its
    z-window form reproduces the planes of the whole-volume walk"""
    lib = build_ops()
    for name, fmt in (("checkered3", po.UINT8), ("ramp16", po.UINT16), ("ramp16", po.UINT32)):
        dims, first, last = (29, 18, 21), (2, 1, 3), (27, 17, 20)
        src = po.synth(dims, fmt, 4)
        a, b = po.Vol(src.copy(), fmt), po.Vol(src.copy(), fmt)
        po.reference().transform_fn1(a, first, last, fn_addr(lib, name))
        po.oracle().transform_fn1(b, first, last, fn_addr(lib, name))
        assert a.a.tobytes() == b.a.tobytes(), name
        win = po.Vol(src[9:12].copy(), fmt)
        po.oracle().transform_fn1(win, (first[0], first[1], 0), (last[0], last[1], 3), fn_addr(lib, name), 9)
        assert win.a.tobytes() == a.a[9:12].tobytes(), name


def test_registration_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    from volkit_b200 import capi
    build_ops()
    lib = capi.load()
    img = open(CUBIN, "rb").read()
    h = ctypes.c_int(-1)
    assert lib.vktb_register_unary_op_image(img, len(img), b"checkered3", ctypes.byref(h)) != 0
    assert b"no usable CUDA device" in lib.vktb_last_error() or b"cuda" in lib.vktb_last_error().lower()


def _gpu_unary(capi, lib, handle, a, fmt, first, last, lo=0.0, hi=1.0):
    from volkit_b200.device import DeviceVolume
    d = DeviceVolume.from_numpy(a, fmt, lo, hi)
    capi.check(lib.vktb_transform_range1(ctypes.byref(d.desc()), capi.I3(*first), capi.I3(*last), handle), "user transform")
    out = d.download()
    d.free()
    return out


@pytest.mark.gpu
@pytest.mark.needs_ref
def test_user_ops_match_the_reference_bit_for_bit():
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    host = build_ops()
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    img = open(CUBIN, "rb").read()
    ref = po.reference()
    handles = {}
    for name in ("checkered3", "ramp16"):
        h = ctypes.c_int(-1)
        capi.check(lib.vktb_register_unary_op_image(img, len(img), name.encode(), ctypes.byref(h)), "register image")
        assert h.value >= 1000
        handles[name] = h.value
        # binding the host function makes the unchanged vkt::Transform(volume, hostFn) resolve it
        capi.check(lib.vktb_bind_unary_op(fn_addr(host, name), h.value))
        got = ctypes.c_int(-1)
        capi.check(lib.vktb_lookup_unary_op(fn_addr(host, name), ctypes.byref(got)))
        assert got.value == h.value
    # shapes: aligned rows, odd rows (partial head / tail chunks), a sub-range with absolute coordinates
    for dims, first, last in (((64, 32, 32), (0, 0, 0), (64, 32, 32)), ((37, 19, 11), (0, 0, 0), (37, 19, 11)),
                              ((50, 33, 21), (3, 5, 7), (47, 30, 20))):
        for name, fmt, pfmt in (("checkered3", capi.UINT8, po.UINT8), ("checkered3", capi.UINT16, po.UINT16),
                                ("ramp16", capi.UINT16, po.UINT16), ("ramp16", capi.UINT32, po.UINT32)):
            src = po.synth(dims, pfmt, 5)
            want = po.Vol(src.copy(), pfmt)
            ref.transform_fn1(want, first, last, fn_addr(host, name))
            got = _gpu_unary(capi, lib, handles[name], src, fmt, first, last)
            assert got.tobytes() == want.a.tobytes(), (name, dims, fmt)
    # binary op, both volumes modified; same and mixed formats
    h2 = ctypes.c_int(-1)
    capi.check(lib.vktb_register_binary_op_image(img, len(img), b"sort_pair", ctypes.byref(h2)), "register binary image")
    for f1, f2 in ((po.UINT16, po.UINT16), (po.UINT8, po.UINT16)):
        dims, first, last = (41, 17, 9), (1, 2, 3), (40, 16, 8)
        a, b = po.synth(dims, f1, 1), po.synth(dims, f2, 2)
        wa, wb = po.Vol(a.copy(), f1), po.Vol(b.copy(), f2)
        ref.transform_fn2(wa, wb, first, last, fn_addr(host, "sort_pair"))
        da, db = DeviceVolume.from_numpy(a, f1), DeviceVolume.from_numpy(b, f2)
        capi.check(lib.vktb_transform_range2(ctypes.byref(da.desc()), ctypes.byref(db.desc()), capi.I3(*first), capi.I3(*last),
                                             h2.value), "user transform2")
        assert da.download().tobytes() == wa.a.tobytes() and db.download().tobytes() == wb.a.tobytes(), (f1, f2)
        da.free(); db.free()
    # a unary handle is not a binary op and vice versa; an image without the kernel is refused with a message
    v = DeviceVolume((8, 8, 8), capi.UINT8)
    assert lib.vktb_transform_range1(ctypes.byref(v.desc()), capi.I3(0, 0, 0), capi.I3(8, 8, 8), h2.value) != 0
    bad = ctypes.c_int(-1)
    assert lib.vktb_register_unary_op_image(img, len(img), b"no_such_op", ctypes.byref(bad)) != 0
    assert b"vktb_unary_no_such_op" in lib.vktb_last_error()
    v.free()


@pytest.mark.gpu
@pytest.mark.needs_ref
def test_user_op_from_source_through_nvrtc():
    from volkit_b200 import capi
    host = build_ops()
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    if not lib.vktb_nvrtc_available():
        pytest.skip("libnvrtc cannot be loaded on this machine")
    source = open(os.path.join(UDIR, "ops.cu"), "rb").read()
    h = ctypes.c_int(-1)
    capi.check(lib.vktb_register_unary_op_source(source, b"ramp16", ctypes.byref(h)), "register source")
    dims = (45, 23, 13)
    src = po.synth(dims, po.UINT16, 9)
    want = po.Vol(src.copy(), po.UINT16)
    po.reference().transform_fn1(want, (0, 0, 0), dims, fn_addr(host, "ramp16"))
    assert _gpu_unary(capi, lib, h.value, src, capi.UINT16, (0, 0, 0), dims).tobytes() == want.a.tobytes()
    # a source that does not compile comes back with NVRTC's log
    assert lib.vktb_register_unary_op_source(b"#include <vktb/transform_op.cuh>\nVKTB_UNARY_OP(broken, x, y, z, v) { nonsense; }",
                                             b"broken", ctypes.byref(h)) != 0
    assert b"nonsense" in lib.vktb_last_error()
