"""Host-side convenience over the C-ABI: a device-resident volume handle.

Mirrors what vkt::StructuredVolume + ManagedBuffer do around the backend (reference
include/cpp/vkt/ManagedBuffer.hpp:168-223): allocate with vktb_alloc, migrate with vktb_memcpy, describe
with a vktb_vol.  numpy arrays are (Z, Y, X)-shaped host mirrors.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi

NP_DTYPES = {capi.UINT8: np.uint8, capi.UINT16: np.uint16, capi.UINT32: np.uint32, capi.FLOAT32: np.float32,
             capi.INT16: np.int16}


def launch_count() -> int:
    return int(capi.load().vktb_launch_count())


def sync() -> None:
    capi.check(capi.load().vktb_sync(), "vktb_sync")


class DeviceVolume:
    """dims = (X, Y, Z); memory lives on the calling thread's device."""

    def __init__(self, dims, fmt: int, lo: float = 0.0, hi: float = 1.0):
        self.dims = tuple(int(d) for d in dims)
        self.fmt, self.lo, self.hi = fmt, float(lo), float(hi)
        self.nbytes = self.dims[0] * self.dims[1] * self.dims[2] * capi.BYTES_PER_VOXEL[fmt]
        p = ctypes.c_void_p()
        capi.check(capi.load().vktb_alloc(ctypes.byref(p), self.nbytes), "vktb_alloc")
        self.ptr = p.value

    @classmethod
    def from_numpy(cls, a: np.ndarray, fmt: int, lo: float = 0.0, hi: float = 1.0) -> "DeviceVolume":
        v = cls(a.shape[::-1], fmt, lo, hi)
        v.upload(a)
        return v

    def desc(self, z0: int = 0, nz: int | None = None) -> capi.VktbVol:
        """vktb_vol of the whole volume or of the z-window [z0, z0+nz)."""
        nz = self.dims[2] - z0 if nz is None else nz
        off = z0 * self.dims[0] * self.dims[1] * capi.BYTES_PER_VOXEL[self.fmt]
        return capi.vol(self.ptr + off, (self.dims[0], self.dims[1], nz), self.fmt, self.lo, self.hi)

    def upload(self, a: np.ndarray) -> None:
        a = np.ascontiguousarray(a)
        assert a.nbytes == self.nbytes, (a.nbytes, self.nbytes)
        capi.check(capi.load().vktb_memcpy(self.ptr, a.ctypes.data, self.nbytes, capi.H2D), "vktb_memcpy H2D")

    def download(self) -> np.ndarray:
        out = np.empty((self.dims[2], self.dims[1], self.dims[0]), NP_DTYPES[self.fmt])
        capi.check(capi.load().vktb_memcpy(out.ctypes.data, self.ptr, self.nbytes, capi.D2H), "vktb_memcpy D2H")
        return out

    def free(self) -> None:
        if self.ptr:
            capi.check(capi.load().vktb_free(self.ptr), "vktb_free")
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
