"""The host mirror's C API (include/vkt/*.h, volkit_b200/lib/libvkt_b200.so) behind the pyoracle interface, so that
tests/fuzz_sweeps.py can drive the WHOLE stack the way a reference user does: volumes are created and filled under the CPU
policy, the thread's policy is switched, the vkt*SV call runs (under Device::GPU that is migrate() -> C-ABI -> kernels),
the result is read back after switching to the CPU policy again.

Covered: Fill, Copy, the ten arithmetic ops, Resample, Flip, Scan, Scale, Rotate, ComputeAggregates (CPU policy) and
Transform with the host ops 0..2 of the oracle's catalogue written as Python callbacks (bound to the built-in device ops
by name for the GPU policy).  Histogram, ApplyFilter and BrickDecompose have no plain-C entry point taking one volume and
are left to the other tests: the sweeps skip what an adapter does not implement."""
import ctypes
import os

import numpy as np

from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "volkit_b200", "lib", "libvkt_b200.so")


class Policy(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int), ("hostApi", ctypes.c_int), ("deviceApi", ctypes.c_int), ("printPerformance", ctypes.c_uint8)]


class Vec3f(ctypes.Structure):
    _fields_ = [("x", ctypes.c_float), ("y", ctypes.c_float), ("z", ctypes.c_float)]


class VoxelView(ctypes.Structure):
    _fields_ = [("bytes", ctypes.POINTER(ctypes.c_uint8)), ("dataFormat", ctypes.c_int), ("mappingLo", ctypes.c_float),
                ("mappingHi", ctypes.c_float)]


UNARY = ctypes.CFUNCTYPE(None, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, VoxelView)
BPV = {po.UINT8: 1, po.UINT16: 2, po.UINT32: 4, po.FLOAT32: 4}
ARITH = ["Sum", "Diff", "Prod", "Quot", "AbsDiff", "SafeSum", "SafeDiff", "SafeProd", "SafeQuot", "SafeAbsDiff"]


def _op_xor_coord(x, y, z, v):
    v.bytes[0] ^= (x + y + z) & 0xff


def _op_invert(x, y, z, v):
    for i in range(BPV[v.dataFormat]):
        v.bytes[i] = ~v.bytes[i] & 0xff


def _op_checker8(x, y, z, v):
    if (((x >> 3) + (y >> 3) + (z >> 3)) & 1) == 0:
        for i in range(BPV[v.dataFormat]):
            v.bytes[i] = 0


class HostMirror:
    """device: "cpu" or "gpu" -- the policy the algorithm calls run under"""

    def __init__(self, device):
        self.lib = ctypes.CDLL(LIB)
        L = self.lib
        self.gpu = device == "gpu"
        L.vktGetThreadExecutionPolicy.restype = Policy
        L.vktSetThreadExecutionPolicy.argtypes = [Policy]
        L.vktStructuredVolumeGetData.restype = ctypes.c_void_p
        L.vktStructuredVolumeGetData.argtypes = [ctypes.c_void_p]
        L.vktStructuredVolumeCreate.argtypes = [ctypes.POINTER(ctypes.c_void_p)] + [ctypes.c_int32] * 3 + [ctypes.c_int] + [ctypes.c_float] * 5
        L.vktStructuredVolumeDestroy.argtypes = [ctypes.c_void_p]
        i9, i6 = [ctypes.c_int32] * 9, [ctypes.c_int32] * 6
        L.vktFillRangeSV.argtypes = [ctypes.c_void_p] + i6 + [ctypes.c_float]
        L.vktCopyRangeSV.argtypes = [ctypes.c_void_p] * 2 + i9
        L.vktScanRangeSV.argtypes = [ctypes.c_void_p] * 2 + i9
        L.vktFlipRangeSV.argtypes = [ctypes.c_void_p] * 2 + i9 + [ctypes.c_int]
        L.vktResampleSV.argtypes = [ctypes.c_void_p] * 2 + [ctypes.c_int]
        L.vktTransformRangeSV1.argtypes = [ctypes.c_void_p] + i6 + [UNARY]
        L.vktBindDeviceOp1.argtypes = [UNARY, ctypes.c_char_p]
        L.vktScaleSV.argtypes = [ctypes.c_void_p] * 2 + [Vec3f, Vec3f]
        L.vktRotateSV.argtypes = [ctypes.c_void_p] * 2 + [Vec3f, ctypes.c_float, Vec3f]
        L.vktComputeAggregatesRangeSV.argtypes = [ctypes.c_void_p, ctypes.c_void_p] + i6
        for name in ARITH:
            getattr(L, f"vkt{name}RangeSV").argtypes = [ctypes.c_void_p] * 3 + i9
        self.cpu_policy = L.vktGetThreadExecutionPolicy()
        self.run_policy = Policy(1 if self.gpu else 0, self.cpu_policy.hostApi, self.cpu_policy.deviceApi, 0)
        # the callbacks must outlive every call that may invoke them
        self.ops = [UNARY(_op_xor_coord), UNARY(_op_invert), UNARY(_op_checker8)]
        if self.gpu:
            for fn, name in zip(self.ops, (b"xor_coord", b"invert", b"checker8")):
                assert L.vktBindDeviceOp1(fn, name) == 0, name

    # ---- plumbing
    def _make(self, v: po.Vol):
        h = ctypes.c_void_p()
        Z, Y, X = v.a.shape
        self.lib.vktStructuredVolumeCreate(ctypes.byref(h), X, Y, Z, v.fmt, 1.0, 1.0, 1.0, v.lo, v.hi)
        ctypes.memmove(self.lib.vktStructuredVolumeGetData(h), v.a.ctypes.data, v.a.nbytes)     # CPU policy: a host pointer
        return h

    def _back(self, h, v: po.Vol):
        ctypes.memmove(v.a.ctypes.data, self.lib.vktStructuredVolumeGetData(h), v.a.nbytes)     # migrates to the host first

    def _call(self, fn, vols, out, *args):
        """vols: Vol objects in argument order (the same object twice = one handle twice); out: the Vols to read back"""
        L = self.lib
        L.vktSetThreadExecutionPolicy(self.cpu_policy)
        handles = {}
        for v in vols:
            if id(v) not in handles:
                handles[id(v)] = self._make(v)
        L.vktSetThreadExecutionPolicy(self.run_policy)
        rc = fn(*[handles[id(v)] for v in vols], *args)
        L.vktSetThreadExecutionPolicy(self.cpu_policy)
        for v in out:
            self._back(handles[id(v)], v)
        for h in handles.values():
            L.vktStructuredVolumeDestroy(h)
        assert rc == 0, (fn.__name__, rc)

    # ---- the pyoracle interface
    def fill_range(self, vol, first, last, value):
        self._call(self.lib.vktFillRangeSV, [vol], [vol], *first, *last, value)

    def copy_range(self, dst, src, first, last, off=(0, 0, 0)):
        self._call(self.lib.vktCopyRangeSV, [dst, src], [dst], *first, *last, *off)

    def arith_range(self, op, dst, s1, s2, first, last, off=(0, 0, 0)):
        self._call(getattr(self.lib, f"vkt{ARITH[op]}RangeSV"), [dst, s1, s2], [dst], *first, *last, *off)

    def transform_range1(self, vol, first, last, op):
        if op > 2:
            return False                 # op 3 needs the codec inside the callback: left to the fixed tests
        self._call(self.lib.vktTransformRangeSV1, [vol], [vol], *first, *last, self.ops[op])
        return True

    def resample(self, dst, src, filt):
        self._call(self.lib.vktResampleSV, [dst, src], [dst], filt)

    def flip_range(self, dst, src, first, last, off=(0, 0, 0), axis=0):
        self._call(self.lib.vktFlipRangeSV, [dst, src], [dst], *first, *last, *off, axis)

    def scan_range(self, dst, src, first, last, off=(0, 0, 0)):
        self._call(self.lib.vktScanRangeSV, [dst, src], [dst], *first, *last, *off)

    def scale(self, dst, src, factor, center):
        self._call(self.lib.vktScaleSV, [dst, src], [dst], Vec3f(*factor), Vec3f(*center))

    def aggregates_range(self, vol, first, last):
        if self.gpu:
            raise NotImplementedError("the CUDA path follows a tolerance contract: tests/cases.py check_aggregates")
        a = po.Aggregates()
        self._call(self.lib.vktComputeAggregatesRangeSV, [vol], [], ctypes.byref(a), *first, *last)
        return a

    def rotate(self, dst, src, axis, angle, center):
        self._call(self.lib.vktRotateSV, [dst, src], [dst], Vec3f(*axis), ctypes.c_float(angle), Vec3f(*center))
