"""TEST INFRASTRUCTURE ONLY -- ctypes access to the two CPU checkers.

* ``oracle()``    -> oracle/libvkt_oracle.so, the plain-C restatement (oracle/vkt_oracle.c)
* ``reference()`` -> oracle/_ref/libvkt_ref.so, the UNMODIFIED reference serial path compiled from
                    /root/reference by oracle/Makefile (prebuilt file travels to the GPU box)

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.  The product
package (volkit_b200/) must never do so.

Volumes are numpy arrays shaped (Z, Y, X) (x fastest, as reference src/vkt/StructuredVolume.cpp:311-317)
wrapped in ``Vol`` together with DataFormat and the (lo, hi) voxel mapping.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# numeric values of vkt::DataFormat (reference include/cpp/vkt/common.hpp:53-66)
INT8, INT16, INT32, UINT8, UINT16, UINT32, FLOAT32 = 1, 2, 3, 4, 5, 6, 7
DTYPES = {INT8: np.int8, INT16: np.int16, INT32: np.int32, UINT8: np.uint8, UINT16: np.uint16,
          UINT32: np.uint32, FLOAT32: np.float32}
FORMAT_NAMES = {UINT8: "UInt8", UINT16: "UInt16", UINT32: "UInt32", FLOAT32: "Float32", INT16: "Int16"}

# arithmetic op ids, order of reference include/cpp/vkt/Arithmetic.hpp
SUM, DIFF, PROD, QUOT, ABSDIFF, SAFESUM, SAFEDIFF, SAFEPROD, SAFEQUOT, SAFEABSDIFF = range(10)
ARITH_NAMES = ["Sum", "Diff", "Prod", "Quot", "AbsDiff", "SafeSum", "SafeDiff", "SafeProd", "SafeQuot",
               "SafeAbsDiff"]
NEAREST, LINEAR, TRILINEAR = 0, 1, 2


class CVol(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dims", ctypes.c_int32 * 3), ("format", ctypes.c_int32),
                ("lo", ctypes.c_float), ("hi", ctypes.c_float)]


I3 = ctypes.c_int32 * 3


class Aggregates(ctypes.Structure):
    """vkt::Aggregates (reference include/cpp/vkt/Aggregates.hpp:14-25)"""
    _fields_ = [("min", ctypes.c_float), ("max", ctypes.c_float), ("mean", ctypes.c_float), ("stddev", ctypes.c_float),
                ("var", ctypes.c_float), ("sum", ctypes.c_float), ("prod", ctypes.c_float),
                ("argmin", ctypes.c_int32 * 3), ("argmax", ctypes.c_int32 * 3)]

    def words(self) -> bytes:
        return bytes(self)


@dataclass
class Vol:
    a: np.ndarray          # (Z, Y, X), C-contiguous, dtype DTYPES[fmt]
    fmt: int
    lo: float = 0.0
    hi: float = 1.0

    def c(self) -> CVol:
        assert self.a.flags["C_CONTIGUOUS"] and self.a.ndim == 3
        v = CVol()
        v.data = self.a.ctypes.data
        v.dims[:] = self.a.shape[::-1]
        v.format = self.fmt
        v.lo = self.lo
        v.hi = self.hi
        return v

    def copy(self) -> "Vol":
        return Vol(self.a.copy(), self.fmt, self.lo, self.hi)


def empty(dims, fmt, lo=0.0, hi=1.0) -> Vol:
    """dims = (X, Y, Z) like the reference API."""
    return Vol(np.zeros((dims[2], dims[1], dims[0]), DTYPES[fmt]), fmt, lo, hi)


class _Lib:
    def __init__(self, path: str, prefix: str, kind: str):
        self.path, self.prefix, self.kind = path, prefix, kind
        self.lib = ctypes.CDLL(path)
        p = ctypes.POINTER(CVol)
        f = self._f
        f("map_voxel", None, [ctypes.c_void_p, ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_float])
        f("unmap_voxel", ctypes.c_float, [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_float])
        f("fill_range", ctypes.c_int, [p, I3, I3, ctypes.c_float])
        f("copy_range", ctypes.c_int, [p, p, I3, I3, I3])
        f("arith_range", ctypes.c_int, [ctypes.c_int, p, p, p, I3, I3, I3])
        f("transform_range1", ctypes.c_int, [p, I3, I3, ctypes.c_int])
        f("transform_range2", ctypes.c_int, [p, p, I3, I3, ctypes.c_int])
        f("resample", ctypes.c_int, [p, p, ctypes.c_int])
        f("flip_range", ctypes.c_int, [p, p, I3, I3, I3, ctypes.c_int])
        f("histogram_range", ctypes.c_int, [p, I3, I3, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p])
        f("scan_range", ctypes.c_int, [p, p, I3, I3, I3])
        F3 = ctypes.c_float * 3
        f("scale", ctypes.c_int, [p, p, F3, F3])
        if prefix == "vko_":
            f("affine_range", ctypes.c_int, [p, p, I3, I3, ctypes.c_float * 9, F3])
        f("rotate", ctypes.c_int, [p, p, F3, ctypes.c_float, F3])
        f("apply_filter", ctypes.c_int, [p, p, ctypes.POINTER(ctypes.c_float), I3])
        if kind == "reference":
            f("transform_fn1", ctypes.c_int, [p, I3, I3, ctypes.c_void_p])
            f("transform_fn2", ctypes.c_int, [p, p, I3, I3, ctypes.c_void_p])
            f("write_range", ctypes.c_int, [p, I3, I3, ctypes.c_char_p])
            f("read_range", ctypes.c_int, [p, I3, I3, ctypes.c_char_p])
            f("write", ctypes.c_int, [p, ctypes.c_char_p])
            f("read", ctypes.c_int, [p, ctypes.c_char_p])
        else:
            f("rows_gather", ctypes.c_size_t, [p, I3, I3, ctypes.c_void_p])
            f("rows_scatter", ctypes.c_size_t, [p, I3, I3, ctypes.c_void_p])
        f("rawfile_parse", None, [ctypes.c_char_p, I3, ctypes.POINTER(ctypes.c_int32)])
        f("aggregates_range", ctypes.c_int, [p, I3, I3, ctypes.POINTER(Aggregates)])
        i32p = ctypes.POINTER(ctypes.c_int32)
        f("brick_layout", ctypes.c_size_t, [I3, I3, I3, I3, I3, i32p])
        f("brick_decompose", ctypes.c_int, [p, I3, I3, I3, ctypes.c_void_p, ctypes.c_size_t])
        if prefix == "vko_":
            F9 = ctypes.c_float * 9
            f("transform_range1_window", ctypes.c_int, [p, I3, I3, ctypes.c_int, ctypes.c_int32])
            f("transform_fn1", ctypes.c_int, [p, I3, I3, ctypes.c_void_p, ctypes.c_int32])
            f("affine_window", ctypes.c_int, [p, ctypes.c_int32, p, ctypes.c_int32, I3, I3, I3, F9, F3])
            f("apply_filter_window", ctypes.c_int, [p, ctypes.c_int32, p, ctypes.c_int32, I3,
                                                    ctypes.POINTER(ctypes.c_float), I3])
            f("scale_matrix", None, [F3, F9])
            f("rotate_matrix", None, [F3, ctypes.c_float, F9])
            f("resample_window", ctypes.c_int, [p, p, ctypes.c_int, ctypes.c_int32, I3, ctypes.c_int32, I3])
            f("synth_fill", None, [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64])

    def _f(self, name, res, args):
        fn = getattr(self.lib, self.prefix + name)
        fn.restype, fn.argtypes = res, args
        setattr(self, "_" + name, fn)

    # ---- scalar codec
    def map_voxel(self, value: float, fmt: int, lo=0.0, hi=1.0) -> bytes:
        b = np.zeros(8, np.uint8)
        self._map_voxel(b.ctypes.data, value, fmt, lo, hi)
        return bytes(b[: np.dtype(DTYPES[fmt]).itemsize])

    def unmap_voxel(self, raw: bytes, fmt: int, lo=0.0, hi=1.0) -> float:
        b = np.frombuffer(bytes(raw) + b"\0" * 8, np.uint8).copy()
        return float(self._unmap_voxel(b.ctypes.data, fmt, lo, hi))

    # ---- algorithms (in place on the numpy buffers)
    def fill_range(self, vol: Vol, first, last, value: float):
        assert self._fill_range(ctypes.byref(vol.c()), I3(*first), I3(*last), value) == 0

    def copy_range(self, dst: Vol, src: Vol, first, last, off=(0, 0, 0)):
        assert self._copy_range(ctypes.byref(dst.c()), ctypes.byref(src.c()), I3(*first), I3(*last), I3(*off)) == 0

    def arith_range(self, op: int, dst: Vol, s1: Vol, s2: Vol, first, last, off=(0, 0, 0)):
        assert self._arith_range(op, ctypes.byref(dst.c()), ctypes.byref(s1.c()), ctypes.byref(s2.c()),
                                 I3(*first), I3(*last), I3(*off)) == 0

    def transform_range1(self, vol: Vol, first, last, op: int):
        assert self._transform_range1(ctypes.byref(vol.c()), I3(*first), I3(*last), op) == 0

    def transform_range2(self, v1: Vol, v2: Vol, first, last, op: int):
        assert self._transform_range2(ctypes.byref(v1.c()), ctypes.byref(v2.c()), I3(*first), I3(*last), op) == 0

    def transform_fn1(self, vol: Vol, first, last, fn_ptr: int, z_origin: int = 0):
        """TransformRange with a caller-supplied host function (address of a TransformUnaryOp); z_origin (restatement
        only) makes `vol` a z-window of a larger volume"""
        if self.kind == "reference":
            assert z_origin == 0
            assert self._transform_fn1(ctypes.byref(vol.c()), I3(*first), I3(*last), fn_ptr) == 0
        else:
            assert self._transform_fn1(ctypes.byref(vol.c()), I3(*first), I3(*last), fn_ptr, z_origin) == 0

    def transform_fn2(self, v1: Vol, v2: Vol, first, last, fn_ptr: int):
        assert self._transform_fn2(ctypes.byref(v1.c()), ctypes.byref(v2.c()), I3(*first), I3(*last), fn_ptr) == 0

    def resample(self, dst: Vol, src: Vol, filter: int):
        assert self._resample(ctypes.byref(dst.c()), ctypes.byref(src.c()), filter) == 0

    def flip_range(self, dst: Vol, src: Vol, first, last, off=(0, 0, 0), axis: int = 0):
        """src may be dst itself (in place)"""
        assert self._flip_range(ctypes.byref(dst.c()), ctypes.byref(src.c()), I3(*first), I3(*last), I3(*off), axis) == 0

    def histogram_range(self, vol: Vol, first, last, num_bins: int):
        """(bins as uint64 array, voxels dropped because their bin index fell outside [0, num_bins))"""
        bins = np.zeros(num_bins, np.uint64)
        dropped = np.zeros(1, np.uint64)
        assert self._histogram_range(ctypes.byref(vol.c()), I3(*first), I3(*last), bins.ctypes.data, num_bins,
                                     dropped.ctypes.data) == 0
        return bins, int(dropped[0])

    def scale(self, dst: Vol, src: Vol, factor, center):
        F3 = ctypes.c_float * 3
        assert self._scale(ctypes.byref(dst.c()), ctypes.byref(src.c()), F3(*factor), F3(*center)) == 0

    def rotate(self, dst: Vol, src: Vol, axis, angle: float, center):
        F3 = ctypes.c_float * 3
        assert self._rotate(ctypes.byref(dst.c()), ctypes.byref(src.c()), F3(*axis), angle, F3(*center)) == 0

    def apply_filter(self, dst: Vol, src: Vol, weights: np.ndarray):
        """weights: float32 array shaped (fz, fy, fx); src may be dst itself"""
        w = np.ascontiguousarray(weights, np.float32)
        assert self._apply_filter(ctypes.byref(dst.c()), ctypes.byref(src.c()),
                                  w.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), I3(*w.shape[::-1])) == 0

    # ---- stream I/O.  The restatement works on byte buffers, the reference on real files.
    def rows_gather(self, vol: Vol, first, last) -> bytes:
        """the bytes OutputStream::writeRange emits"""
        n = max(0, last[0] - first[0]) * max(0, last[1] - first[1]) * max(0, last[2] - first[2]) * vol.a.dtype.itemsize
        buf = np.zeros(max(n, 1), np.uint8)
        got = self._rows_gather(ctypes.byref(vol.c()), I3(*first), I3(*last), buf.ctypes.data)
        return buf[:got].tobytes()

    def rows_scatter(self, vol: Vol, first, last, data: bytes):
        """what InputStream::readRange does with the stream `data`"""
        buf = np.frombuffer(data, np.uint8).copy()
        self._rows_scatter(ctypes.byref(vol.c()), I3(*first), I3(*last), buf.ctypes.data)

    def write_range(self, vol: Vol, first, last, path: str) -> int:
        return self._write_range(ctypes.byref(vol.c()), I3(*first), I3(*last), path.encode())

    def read_range(self, vol: Vol, first, last, path: str) -> int:
        return self._read_range(ctypes.byref(vol.c()), I3(*first), I3(*last), path.encode())

    def write(self, vol: Vol, path: str) -> int:
        return self._write(ctypes.byref(vol.c()), path.encode())

    def read(self, vol: Vol, path: str) -> int:
        return self._read(ctypes.byref(vol.c()), path.encode())

    def rawfile_parse(self, name: str):
        """(dims, DataFormat) RawFile's constructor parses out of a file name (the reference needs the file to exist:
        its destructor fclose()s a null FILE* otherwise)"""
        dims, fmt = I3(), ctypes.c_int32()
        self._rawfile_parse(name.encode(), dims, ctypes.byref(fmt))
        return tuple(dims), fmt.value

    def aggregates_range(self, vol: Vol, first, last) -> "Aggregates":
        a = Aggregates()
        assert self._aggregates_range(ctypes.byref(vol.c()), I3(*first), I3(*last), ctypes.byref(a)) == 0
        return a

    def scan_range(self, dst: Vol, src: Vol, first, last, off=(0, 0, 0)):
        """src may be dst itself (in place; then first > last on an axis scans right to left)"""
        assert self._scan_range(ctypes.byref(dst.c()), ctypes.byref(src.c()), I3(*first), I3(*last), I3(*off)) == 0

    def brick_layout(self, sdims, brick, halo_neg=(0, 0, 0), halo_pos=(0, 0, 0)):
        """BrickDecomposeResize: (num_bricks (x, y, z), [brick dims (x, y, z), ...] in x-fastest brick order)"""
        nb = I3()
        self._brick_layout(I3(*sdims), I3(*brick), I3(*halo_neg), I3(*halo_pos), nb, None)
        n = nb[0] * nb[1] * nb[2]
        dims = (ctypes.c_int32 * (3 * n))()
        self._brick_layout(I3(*sdims), I3(*brick), I3(*halo_neg), I3(*halo_pos), nb, dims)
        return tuple(nb), [tuple(dims[3 * i: 3 * i + 3]) for i in range(n)]

    def brick_decompose(self, src: Vol, brick, halo_neg=(0, 0, 0), halo_pos=(0, 0, 0)):
        """BrickDecomposeResize + BrickDecompose: the bricks as a list of (Z, Y, X) arrays, x-fastest brick order"""
        sdims = src.a.shape[::-1]
        _, dims = self.brick_layout(sdims, brick, halo_neg, halo_pos)
        out = np.zeros(sum(d[0] * d[1] * d[2] for d in dims), DTYPES[src.fmt])
        assert self._brick_decompose(ctypes.byref(src.c()), I3(*brick), I3(*halo_neg), I3(*halo_pos), out.ctypes.data,
                                     out.nbytes) == 0
        bricks, at = [], 0
        for d in dims:
            n = d[0] * d[1] * d[2]
            bricks.append(out[at: at + n].reshape(d[2], d[1], d[0]))
            at += n
        return bricks

    # ---- z-window forms (restatement only): plane-by-plane checks of full-size volumes
    def transform_range1_window(self, vol: Vol, first, last, op: int, z_origin: int):
        assert self._transform_range1_window(ctypes.byref(vol.c()), I3(*first), I3(*last), op, z_origin) == 0

    def scale_matrix(self, factor):
        m = (ctypes.c_float * 9)()
        self._scale_matrix((ctypes.c_float * 3)(*factor), m)
        return list(m)

    def rotate_matrix(self, axis, angle):
        m = (ctypes.c_float * 9)()
        self._rotate_matrix((ctypes.c_float * 3)(*axis), angle, m)
        return list(m)

    def affine_range(self, dst: Vol, src: Vol, first, last, m, center):
        assert self._affine_range(ctypes.byref(dst.c()), ctypes.byref(src.c()), I3(*first), I3(*last),
                                  (ctypes.c_float * 9)(*m), (ctypes.c_float * 3)(*center)) == 0

    def affine_window(self, dst: Vol, dz0: int, src: Vol, sz0: int, src_full, first, last, m, center):
        assert self._affine_window(ctypes.byref(dst.c()), dz0, ctypes.byref(src.c()), sz0, I3(*src_full), I3(*first),
                                   I3(*last), (ctypes.c_float * 9)(*m), (ctypes.c_float * 3)(*center)) == 0

    def apply_filter_window(self, dst: Vol, dz0: int, src: Vol, sz0: int, full, weights: np.ndarray):
        w = np.ascontiguousarray(weights, np.float32)
        rc = self._apply_filter_window(ctypes.byref(dst.c()), dz0, ctypes.byref(src.c()), sz0, I3(*full),
                                       w.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), I3(*w.shape[::-1]))
        assert rc == 0, rc

    def resample_window(self, dst: Vol, src: Vol, filter: int, dz0: int, dst_full, sz0: int, src_full):
        assert self._resample_window(ctypes.byref(dst.c()), ctypes.byref(src.c()), filter, dz0, I3(*dst_full),
                                     sz0, I3(*src_full)) == 0


_cache: dict = {}


class _RenderLib:
    """the headless ray-marcher of either checker: oracle/libvkt_oracle.so (vko_render_raymarch) or the reference's own
    kernel compiled into oracle/_ref/libvkt_ref_render.so (ref_render_raymarch)"""

    def __init__(self, lib, prefix):
        F3 = ctypes.c_float * 3
        self.fn = getattr(lib, prefix + "render_raymarch")
        self.fn.restype = ctypes.c_int
        self.fn.argtypes = [ctypes.POINTER(CVol), ctypes.c_void_p, ctypes.c_int, F3, F3, F3, ctypes.c_float, ctypes.c_int,
                            ctypes.c_int, ctypes.c_float, ctypes.c_int, F3, ctypes.c_void_p, ctypes.c_void_p]

    def raymarch(self, vol: Vol, lut, eye, center, up, fovy_deg, w, h, dt=1.0, srgb=True, dist=(1.0, 1.0, 1.0)):
        """(rgba32f (h, w, 4), rgba8 (h, w, 4)); lut = (n, 4) float32 array or None"""
        F3 = ctypes.c_float * 3
        f32, u8 = np.zeros((h, w, 4), np.float32), np.zeros((h, w, 4), np.uint8)
        l = None if lut is None else np.ascontiguousarray(lut, np.float32)
        rc = self.fn(ctypes.byref(vol.c()), None if l is None else l.ctypes.data, 0 if l is None else l.shape[0], F3(*eye),
                     F3(*center), F3(*up), fovy_deg, w, h, dt, 1 if srgb else 0, F3(*dist), f32.ctypes.data, u8.ctypes.data)
        assert rc == 0, rc
        return f32, u8


def render_oracle() -> _RenderLib:
    return _RenderLib(oracle().lib, "vko_")


def have_reference_render() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libvkt_ref_render.so"))


def render_reference() -> _RenderLib:
    if "rr" not in _cache:
        _cache["rr"] = ctypes.CDLL(os.path.join(HERE, "_ref", "libvkt_ref_render.so"))
    return _RenderLib(_cache["rr"], "ref_")


def default_camera(dims, w, h, dist=(1.0, 1.0, 1.0), reference=False):
    """(eye, center, up) of the viewer's default camera (45 degrees, view_all)"""
    F3 = ctypes.c_float * 3
    e, c, u = F3(), F3(), F3()
    if reference:
        render_reference()
        _cache["rr"].ref_render_default_camera(I3(*dims), F3(*dist), w, h, e, c, u)
    else:
        oracle().lib.vko_render_default_camera(I3(*dims), F3(*dist), e, c, u)
    return tuple(e), tuple(c), tuple(u)


def synth(n_or_dims, fmt: int, seed: int, first_voxel: int = 0) -> np.ndarray:
    """SURVEY.md 8(d) counter-based inputs: element i <- splitmix64(i ^ seed*phi)."""
    lib = oracle()
    if isinstance(n_or_dims, (tuple, list)):
        x, y, z = n_or_dims
        out = np.empty((z, y, x), DTYPES[fmt])
    else:
        out = np.empty(int(n_or_dims), DTYPES[fmt])
    lib._synth_fill(out.ctypes.data, first_voxel, out.size, fmt, seed)
    return out


def synth_numpy(n: int, fmt: int, seed: int, first_voxel: int = 0, index: np.ndarray | None = None) -> np.ndarray:
    """Vectorised numpy twin of vko_synth_fill (fast enough for 10^9 voxels): voxels [first_voxel, first_voxel + n),
    or the voxels whose linear indices `index` (any shape, uint64) names."""
    M = np.uint64(0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        k = np.uint64((seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
        i = np.arange(first_voxel, first_voxel + n, dtype=np.uint64) if index is None else index.astype(np.uint64)
        x = (i ^ k) + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    _ = M
    if fmt == FLOAT32:
        return ((x >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)
    return x.astype(DTYPES[fmt])




def build(ref: bool = True) -> None:
    """Build the checkers (building the checker is not using it)."""
    target = "all" if ref else "oracle"
    subprocess.run(["make", "-C", HERE, target], check=True, stdout=subprocess.DEVNULL)


def oracle() -> _Lib:
    if "o" not in _cache:
        path = os.path.join(HERE, "libvkt_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _cache["o"] = _Lib(path, "vko_", "port")
    return _cache["o"]


def have_reference() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libvkt_ref.so"))


def reference() -> _Lib:
    if "r" not in _cache:
        _cache["r"] = _Lib(os.path.join(HERE, "_ref", "libvkt_ref.so"), "ref_", "reference")
    return _cache["r"]
