"""Runs the case catalogue (tests/cases.py) on the CUDA product THROUGH THE C-ABI: every call uploads the
numpy inputs with vktb_memcpy, runs one vktb_* algorithm on the device and downloads the outputs.
Exposes the same method set as oracle.pyoracle._Lib so cases.run_case() works unchanged."""
from __future__ import annotations

import ctypes

from oracle import pyoracle as po
from volkit_b200 import capi
from volkit_b200.device import DeviceVolume

I3 = capi.I3


class GpuLib:
    kind = "cuda"

    def __init__(self):
        self.lib = capi.load()

    @staticmethod
    def _up(v: po.Vol) -> DeviceVolume:
        return DeviceVolume.from_numpy(v.a, v.fmt, v.lo, v.hi)

    @staticmethod
    def _down(dv: DeviceVolume, v: po.Vol) -> None:
        v.a[...] = dv.download()
        dv.free()

    def fill_range(self, vol, first, last, value):
        d = self._up(vol)
        capi.check(self.lib.vktb_fill_range(ctypes.byref(d.desc()), I3(*first), I3(*last), value), "fill")
        self._down(d, vol)

    def copy_range(self, dst, src, first, last, off=(0, 0, 0)):
        d, s = self._up(dst), self._up(src)
        capi.check(self.lib.vktb_copy_range(ctypes.byref(d.desc()), ctypes.byref(s.desc()), I3(*first), I3(*last),
                                            I3(*off)), "copy")
        self._down(d, dst)
        s.free()

    def scan_range(self, dst, src, first, last, off=(0, 0, 0)):
        d = self._up(dst)
        s = d if src is dst else self._up(src)
        capi.check(self.lib.vktb_scan_range(ctypes.byref(d.desc()), ctypes.byref(s.desc()), I3(*first), I3(*last),
                                            I3(*off)), "scan")
        self._down(d, dst)
        if s is not d:
            s.free()

    def copy_range_batch(self, dsts, src, firsts, lasts, offs):
        """n CopyRange calls from one source in ONE vktb_copy_range_batch launch; dsts are updated in place"""
        n = len(dsts)
        s = self._up(src)
        dv = [self._up(d) for d in dsts]
        descs = (capi.VktbVol * n)(*[d.desc() for d in dv])
        flat = lambda t: (ctypes.c_int32 * (3 * n))(*[c for v in t for c in v])
        capi.check(self.lib.vktb_copy_range_batch(n, descs, ctypes.byref(s.desc()), flat(firsts), flat(lasts), flat(offs)),
                   "copy batch")
        for d, v in zip(dv, dsts):
            self._down(d, v)
        s.free()

    def brick_decompose(self, src, brick, halo_neg=(0, 0, 0), halo_pos=(0, 0, 0), prefill_seed=None):
        """BrickDecomposeResize's layout (host logic, taken from the oracle) + one batched launch for all bricks"""
        sdims = src.a.shape[::-1]
        nb, dims = po.oracle().brick_layout(sdims, brick, halo_neg, halo_pos)
        bricks = [po.Vol(po.synth(d, src.fmt, 7 + i), src.fmt, src.lo, src.hi) for i, d in enumerate(dims)]
        firsts, lasts = [], []
        for z in range(nb[2]):
            for y in range(nb[1]):
                for x in range(nb[0]):
                    f = [i * b for i, b in zip((x, y, z), brick)]
                    lasts.append(tuple(min(f_ + b, d) + p for f_, b, d, p in zip(f, brick, sdims, halo_pos)))
                    firsts.append(tuple(f_ - n_ for f_, n_ in zip(f, halo_neg)))
        self.copy_range_batch(bricks, src, firsts, lasts, [(0, 0, 0)] * len(bricks))
        return [b.a for b in bricks]

    def arith_range(self, op, dst, s1, s2, first, last, off=(0, 0, 0)):
        d, a, b = self._up(dst), self._up(s1), self._up(s2)
        capi.check(self.lib.vktb_arith_range(op, ctypes.byref(d.desc()), ctypes.byref(a.desc()), ctypes.byref(b.desc()),
                                             I3(*first), I3(*last), I3(*off)), "arith")
        self._down(d, dst)
        a.free(); b.free()

    def transform_range1(self, vol, first, last, op):
        d = self._up(vol)
        capi.check(self.lib.vktb_transform_range1(ctypes.byref(d.desc()), I3(*first), I3(*last), op), "transform1")
        self._down(d, vol)

    def transform_range2(self, v1, v2, first, last, op):
        a, b = self._up(v1), self._up(v2)
        capi.check(self.lib.vktb_transform_range2(ctypes.byref(a.desc()), ctypes.byref(b.desc()), I3(*first),
                                                  I3(*last), op), "transform2")
        self._down(a, v1)
        self._down(b, v2)

    def resample(self, dst, src, filt):
        d, s = self._up(dst), self._up(src)
        capi.check(self.lib.vktb_resample(ctypes.byref(d.desc()), ctypes.byref(s.desc()), filt), "resample")
        self._down(d, dst)
        s.free()

    def flip_range(self, dst, src, first, last, off=(0, 0, 0), axis=0):
        d = self._up(dst)
        s = d if src is dst else self._up(src)
        capi.check(self.lib.vktb_flip_range(ctypes.byref(d.desc()), ctypes.byref(s.desc()), I3(*first), I3(*last),
                                            I3(*off), axis), "flip")
        if s is not d:
            s.free()
        self._down(d, dst)

    def _affine(self, dst, src, m, center):
        d, s = self._up(dst), self._up(src)
        F9, F3 = ctypes.c_float * 9, ctypes.c_float * 3
        capi.check(self.lib.vktb_affine_range(ctypes.byref(d.desc()), ctypes.byref(s.desc()), I3(0, 0, 0),
                                              I3(*src.a.shape[::-1]), F9(*m), F3(*center)), "affine")
        self._down(d, dst)
        s.free()

    def scale(self, dst, src, factor, center):
        """the matrix comes from the host side of the boundary (vkt::Scale builds it); here: the oracle's twin"""
        F3, F9 = ctypes.c_float * 3, ctypes.c_float * 9
        m = F9()
        po.oracle().lib.vko_scale_matrix(F3(*factor), m)
        self._affine(dst, src, list(m), center)

    def rotate(self, dst, src, axis, angle, center):
        F3, F9 = ctypes.c_float * 3, ctypes.c_float * 9
        m = F9()
        po.oracle().lib.vko_rotate_matrix(F3(*axis), ctypes.c_float(angle), m)
        self._affine(dst, src, list(m), center)

    def apply_filter(self, dst, src, weights):
        import numpy as np
        w = np.ascontiguousarray(weights, np.float32)
        d = self._up(dst)
        s = d if src is dst else self._up(src)
        capi.check(self.lib.vktb_apply_filter(ctypes.byref(d.desc()), ctypes.byref(s.desc()),
                                              w.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), I3(*w.shape[::-1]), 2),
                   "apply_filter")
        self._down(d, dst)
        if s is not d:
            s.free()

    def aggregates_range(self, vol, first, last):
        """vktb_aggregates_range -> po.Aggregates (same 13 words)"""
        v = self._up(vol)
        a = capi.VktbAggregates()
        capi.check(self.lib.vktb_aggregates_range(ctypes.byref(v.desc()), I3(*first), I3(*last), ctypes.byref(a)),
                   "aggregates")
        v.free()
        return po.Aggregates.from_buffer_copy(bytes(a))

    def histogram_range(self, vol, first, last, num_bins):
        import numpy as np
        v = self._up(vol)
        bins = ctypes.c_void_p()
        capi.check(self.lib.vktb_alloc(ctypes.byref(bins), num_bins * 8), "alloc bins")
        dropped = ctypes.c_uint64(0)
        capi.check(self.lib.vktb_histogram_range(ctypes.byref(v.desc()), I3(*first), I3(*last), bins, num_bins,
                                                 ctypes.byref(dropped)), "histogram")
        out = np.zeros(num_bins, np.uint64)
        capi.check(self.lib.vktb_memcpy(out.ctypes.data, bins, num_bins * 8, capi.D2H), "download bins")
        capi.check(self.lib.vktb_free(bins))
        v.free()
        return out, int(dropped.value)

    def resample_window(self, dst, src, filt, dz0, dst_full, sz0, src_full):
        d, s = self._up(dst), self._up(src)
        capi.check(self.lib.vktb_resample_window(ctypes.byref(d.desc()), ctypes.byref(s.desc()), filt, dz0,
                                                 I3(*dst_full), sz0, I3(*src_full)), "resample_window")
        self._down(d, dst)
        s.free()


class HostedLib:
    """The vktb_hosted_* entry points: inputs live in pinned host memory (vktb_host_alloc), the library streams
    z-slabs through the device.  Same method names as GpuLib for the three algorithms it covers."""
    kind = "cuda-hosted"

    def __init__(self):
        self.lib = capi.load()

    def _pin(self, v: po.Vol):
        import numpy as np
        p = ctypes.c_void_p()
        n = max(1, v.a.nbytes)
        capi.check(self.lib.vktb_host_alloc(ctypes.byref(p), n), "vktb_host_alloc")
        view = np.ctypeslib.as_array(ctypes.cast(p.value, ctypes.POINTER(ctypes.c_uint8)), shape=(n,))
        view[:v.a.nbytes] = np.frombuffer(v.a.tobytes(), dtype=np.uint8)
        x, y, z = v.a.shape[2], v.a.shape[1], v.a.shape[0]
        return p, view, capi.vol(p.value, (x, y, z), v.fmt, v.lo, v.hi)

    def _unpin(self, pinned, v: po.Vol, write_back: bool):
        import numpy as np
        p, view, _ = pinned
        if write_back:
            v.a[...] = np.frombuffer(view[:v.a.nbytes].tobytes(), dtype=v.a.dtype).reshape(v.a.shape)
        capi.check(self.lib.vktb_host_free(p), "vktb_host_free")

    def arith_range(self, op, dst, s1, s2, first, last, off=(0, 0, 0)):
        d = self._pin(dst)
        a = d if s1 is dst else self._pin(s1)
        b = d if s2 is dst else (a if s2 is s1 else self._pin(s2))
        capi.check(self.lib.vktb_hosted_arith_range(op, ctypes.byref(d[2]), ctypes.byref(a[2]), ctypes.byref(b[2]),
                                                    I3(*first), I3(*last), I3(*off)), "hosted arith")
        if b is not d and b is not a:
            self._unpin(b, s2, False)
        if a is not d:
            self._unpin(a, s1, False)
        self._unpin(d, dst, True)

    def transform_range1(self, vol, first, last, op):
        d = self._pin(vol)
        capi.check(self.lib.vktb_hosted_transform_range1(ctypes.byref(d[2]), I3(*first), I3(*last), op), "hosted transform1")
        self._unpin(d, vol, True)

    def resample(self, dst, src, filt):
        d, s = self._pin(dst), self._pin(src)
        capi.check(self.lib.vktb_hosted_resample(ctypes.byref(d[2]), ctypes.byref(s[2]), filt), "hosted resample")
        self._unpin(s, src, False)
        self._unpin(d, dst, True)
