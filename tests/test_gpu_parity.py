"""GPU parity tests proper: the CUDA path, called through the C-ABI, against (a) the frozen outputs of the
reference build (tests/golden/golden.json) and (b) the C restatement on larger seeded inputs.
Bit-exact for every integer format and for Float32 up to NaN payload."""
import ctypes
import hashlib
import json
import os

import numpy as np
import pytest

import cases
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
ALL = cases.cases()


@pytest.fixture(scope="module")
def gpu():
    from gpu_adapter import GpuLib
    from volkit_b200 import launch_count
    g = GpuLib()
    before = launch_count()
    yield g
    assert launch_count() > before, "no CUDA kernel was launched: the tests ran on something else"


def same_bits(got: np.ndarray, ref: np.ndarray) -> bool:
    if got.tobytes() == ref.tobytes():
        return True
    if got.dtype != np.float32:
        return False
    # Float32: NaN payload/sign may differ (x86 propagates payloads, the GPU returns the canonical NaN)
    gn, rn = np.isnan(got), np.isnan(ref)
    if not (gn == rn).all():
        return False
    return got[~gn].tobytes() == ref[~rn].tobytes()


@pytest.mark.parametrize("kind", sorted({c[1] for c in ALL}))
def test_golden(gpu, kind):
    o = po.oracle()
    for case in ALL:
        if case[1] != kind:
            continue
        outs = cases.run_case(gpu, case)
        if kind == "agg":
            # min/max/argmin/argmax are frozen; the accumulated fields follow a tolerance contract (test_aggregates)
            assert hashlib.sha256(outs[0].tobytes()).hexdigest() == GOLD["sha256"][case[0]][0], case[0]
            continue
        digest = [hashlib.sha256(a.tobytes()).hexdigest() for a in outs]
        if digest != GOLD["sha256"][case[0]]:
            # only Float32 NaN payloads may differ from the frozen reference bytes
            refs = cases.run_case(o, case)
            for g, r in zip(outs, refs):
                assert same_bits(g, r), case[0]
                assert g.dtype == np.float32, case[0]


def test_codec_probe(gpu):
    """device map/unmap vs the oracle over edge values (wrap, not saturate)."""
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    o = po.oracle()
    vals = np.array([0.0, -0.0, 0.1, 0.5, 0.999, 1.0, 1.5, -0.25, 2.0, -1.0, 1e-30, 1e10, -1e10, 3e38, -3e38,
                     np.inf, -np.inf, np.nan, 0.99999994, 32767.5 / 65536, 255.0 / 255.999, 2999.0, -999.5, 4e9,
                     2147483520.0, 2147483648.0, -2147483648.0, 1e19, -1e19] + list(np.linspace(-2, 3, 997)), np.float32)
    n = vals.size
    dv = DeviceVolume.from_numpy(vals.reshape(1, 1, n), capi.FLOAT32)
    dm = DeviceVolume((n, 1, 1), capi.UINT32)
    du = DeviceVolume((n, 1, 1), capi.FLOAT32)
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32, po.INT16):
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0), (0.25, 0.75), (1.0, 1.0), (1.0, 0.0), (0.0, 0.3)):
            capi.check(gpu.lib.vktb_debug_map_unmap(fmt, lo, hi, dv.ptr, dm.ptr, du.ptr, n), "probe")
            mapped = dm.download().ravel()
            unm = du.download().ravel()
            nb = np.dtype(po.DTYPES[fmt]).itemsize
            for i in range(n):
                raw = o.map_voxel(float(vals[i]), fmt, lo, hi)
                ref = int.from_bytes(raw, "little")
                if fmt == po.FLOAT32 and np.isnan(np.frombuffer(raw, np.float32)[0]):
                    assert np.isnan(np.uint32(mapped[i]).view(np.float32))
                    continue
                assert int(mapped[i]) == ref, (fmt, lo, hi, float(vals[i]))
                back = np.float32(o.unmap_voxel(raw, fmt, lo, hi))
                assert np.float32(unm[i]).tobytes() == back.tobytes() or (np.isnan(back) and np.isnan(unm[i])), \
                    (fmt, lo, hi, float(vals[i]), nb)


def test_arith_exhaustive_u8_pairs(gpu):
    o = po.oracle()
    g = np.arange(256, dtype=np.uint8)
    A = np.repeat(g, 256).reshape(1, 256, 256).copy()
    B = np.tile(g, 256).reshape(1, 256, 256).copy()
    for op in range(10):
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
            a, b = po.empty((256, 256, 1), po.UINT8, lo, hi), po.empty((256, 256, 1), po.UINT8, lo, hi)
            o.arith_range(op, a, po.Vol(A, po.UINT8), po.Vol(B, po.UINT8, lo, hi), (0, 0, 0), (256, 256, 1))
            gpu.arith_range(op, b, po.Vol(A, po.UINT8), po.Vol(B, po.UINT8, lo, hi), (0, 0, 0), (256, 256, 1))
            assert a.a.tobytes() == b.a.tobytes(), (op, lo, hi)


def test_arith_u8_pair_table_stream(gpu):
    """UInt8 streams of >= 2^24 voxels take the 64 KiB pair-table kernel: all ten ops, two mappings, in-place use, and an
    unaligned contiguous range (scalar head / tail)."""
    o = po.oracle()
    dims = (512, 256, 128)
    s1, s2 = cases.mk(dims, po.UINT8, 1), cases.mk(dims, po.UINT8, 2, -1000.0, 3000.0)
    for op in range(10):
        lo, hi = ((0.0, 1.0), (-1000.0, 3000.0))[op % 2]
        a, b = po.empty(dims, po.UINT8, lo, hi), po.empty(dims, po.UINT8, lo, hi)
        o.arith_range(op, a, s1, s2, (0, 0, 0), dims)
        gpu.arith_range(op, b, s1, s2, (0, 0, 0), dims)
        assert a.a.tobytes() == b.a.tobytes(), op
    # identity mappings everywhere (the IDENT template), dst aliasing source 1
    t2 = cases.mk(dims, po.UINT8, 2)
    a, b = s1.copy(), s1.copy()
    o.arith_range(po.SAFESUM, a, a, t2, (0, 0, 0), dims)
    gpu.arith_range(po.SAFESUM, b, b, t2, (0, 0, 0), dims)
    assert a.a.tobytes() == b.a.tobytes()
    # one long row: contiguous but starting at an odd byte -> scalar head and tail around the vector body
    long = (dims[0] * dims[1] * dims[2], 1, 1)
    r1, r2 = po.Vol(s1.a.reshape(1, 1, -1), po.UINT8), po.Vol(t2.a.reshape(1, 1, -1), po.UINT8)
    a, b = po.empty(long, po.UINT8), po.empty(long, po.UINT8)
    o.arith_range(po.SAFEDIFF, a, r1, r2, (3, 0, 0), (long[0] - 5, 1, 1))
    gpu.arith_range(po.SAFEDIFF, b, r1, r2, (3, 0, 0), (long[0] - 5, 1, 1))
    assert a.a.tobytes() == b.a.tobytes()


def test_arith_u8_pair_table_rows(gpu):
    """UInt8 *Range calls of >= 2^24 voxels on rows in one 16-byte phase take the pair-table rows kernel: sources at
    the dst phase and off it, partial head / tail chunks, two mappings, a range that ends on the volume's last voxel."""
    o = po.oracle()
    sd, dd = (528, 200, 170), (560, 204, 172)
    s1, s2 = cases.mk(sd, po.UINT8, 1), cases.mk(sd, po.UINT8, 2, -1000.0, 3000.0)
    for k, (first, last, off) in enumerate((((0, 0, 0), sd, (16, 2, 1)), ((0, 0, 0), sd, (3, 1, 2)),
                                            ((5, 1, 0), sd, (0, 0, 0)), ((0, 0, 0), (523, 200, 170), (37, 4, 2)))):
        for op in (po.SAFESUM, po.PROD, po.SAFEABSDIFF)[k % 2:][:2]:
            lo, hi = ((0.0, 1.0), (-1000.0, 3000.0))[k % 2]
            a, b = cases.mk(dd, po.UINT8, 3, lo, hi), cases.mk(dd, po.UINT8, 3, lo, hi)
            o.arith_range(op, a, s1, s2, first, last, off)
            gpu.arith_range(op, b, s1, s2, first, last, off)
            assert a.a.tobytes() == b.a.tobytes(), (first, last, off, op)


def test_arith_u16_edge_pairs(gpu):
    """2^16-stride sweep + all edge pairs of UInt16 x UInt16 for each op."""
    o = po.oracle()
    edge = np.array([0, 1, 2, 255, 256, 32767, 32768, 32769, 65534, 65535], np.uint16)
    sweep = np.arange(0, 65536, 97, dtype=np.uint16)
    vals = np.concatenate([edge, sweep])
    A = np.repeat(vals, vals.size).reshape(1, vals.size, vals.size).copy()
    B = np.tile(vals, vals.size).reshape(1, vals.size, vals.size).copy()
    dims = (vals.size, vals.size, 1)
    for op in range(10):
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
            a, b = po.empty(dims, po.UINT16, lo, hi), po.empty(dims, po.UINT16, lo, hi)
            o.arith_range(op, a, po.Vol(A, po.UINT16, lo, hi), po.Vol(B, po.UINT16), (0, 0, 0), dims)
            gpu.arith_range(op, b, po.Vol(A, po.UINT16, lo, hi), po.Vol(B, po.UINT16), (0, 0, 0), dims)
            assert a.a.tobytes() == b.a.tobytes(), (op, lo, hi)


def test_config2_512cube_uint16(gpu):
    """BASELINE config 2 AS WRITTEN: 512^3 UInt16 SafeSum / SafeDiff / CopyRange with dstOffset on one B200 vs the CPU
    path, bit-exact over the WHOLE volume.  The oracle runs on 16-plane z-slabs in a thread pool (ctypes releases the
    GIL); a slab of a dense volume is a contiguous sub-volume, so slab results concatenate to the whole-volume result."""
    from concurrent.futures import ThreadPoolExecutor
    o = po.oracle()
    E = 512
    dims = (E, E, E)
    s1 = po.Vol(po.synth_numpy(E ** 3, po.UINT16, 1).reshape(E, E, E), po.UINT16)
    s2 = po.Vol(po.synth_numpy(E ** 3, po.UINT16, 2).reshape(E, E, E), po.UINT16)
    slabs = [(z, min(z + 16, E)) for z in range(0, E, 16)]
    pool = ThreadPoolExecutor(min(32, os.cpu_count() or 1))
    for op in (po.SAFESUM, po.SAFEDIFF):
        want = po.empty(dims, po.UINT16)

        def one(zz, op=op, want=want):
            z0, z1 = zz
            o.arith_range(op, po.Vol(want.a[z0:z1], po.UINT16), po.Vol(s1.a[z0:z1], po.UINT16), po.Vol(s2.a[z0:z1], po.UINT16),
                          (0, 0, 0), (E, E, z1 - z0))
        list(pool.map(one, slabs))
        got = po.empty(dims, po.UINT16)
        gpu.arith_range(op, got, s1, s2, (0, 0, 0), dims)
        assert got.a.tobytes() == want.a.tobytes(), po.ARITH_NAMES[op]
    # CopyRange first (0,0,0), last (504,504,504), dstOffset (8,8,8), dst pre-filled with seed 3
    m = E - 8
    pre = po.synth_numpy(E ** 3, po.UINT16, 3).reshape(E, E, E)
    want, got = po.Vol(pre.copy(), po.UINT16), po.Vol(pre.copy(), po.UINT16)

    def copy_one(zz):
        z0, z1 = zz[0], min(zz[1], m)
        if z1 > z0:
            o.copy_range(po.Vol(want.a[8 + z0:8 + z1], po.UINT16), po.Vol(s1.a[z0:z1], po.UINT16), (0, 0, 0), (m, m, z1 - z0), (8, 8, 0))
    list(pool.map(copy_one, slabs))
    gpu.copy_range(got, s1, (0, 0, 0), (m, m, m), (8, 8, 8))
    assert got.a.tobytes() == want.a.tobytes()
    assert got.a[:8].tobytes() == pre[:8].tobytes() and got.a[100, :8].tobytes() == pre[100, :8].tobytes()   # untouched border
    # the same with the clamped source border: first < 0
    want, got = po.Vol(pre.copy(), po.UINT16), po.Vol(pre.copy(), po.UINT16)

    def clamp_one(zz):
        # dst plane 8 + (pz + 3) <- src plane clamp(pz), pz in [-3, 400): slab by dst planes
        for pz in range(zz[0] - 3, min(zz[1], 403) - 3):
            sz = min(max(pz, 0), E - 1)
            o.copy_range(po.Vol(want.a[pz + 11:pz + 12], po.UINT16), po.Vol(s1.a[sz:sz + 1], po.UINT16), (-3, -2, 0), (400, 410, 1), (8, 8, 0))
    list(pool.map(clamp_one, slabs))
    gpu.copy_range(got, s1, (-3, -2, -3), (400, 410, 400), (8, 8, 8))
    assert got.a.tobytes() == want.a.tobytes()
    pool.shutdown()


@pytest.mark.parametrize("fmt", cases.FMT4)
def test_config2_256cube(gpu, fmt):
    """BASELINE config 2 shape at 256^3: SafeSum/SafeDiff whole volume, CopyRange with dstOffset."""
    o = po.oracle()
    dims = (256, 256, 256)
    s1, s2 = cases.mk(dims, fmt, 1), cases.mk(dims, fmt, 2)
    for op in (po.SAFESUM, po.SAFEDIFF):
        a, b = po.empty(dims, fmt), po.empty(dims, fmt)
        o.arith_range(op, a, s1, s2, (0, 0, 0), dims)
        gpu.arith_range(op, b, s1, s2, (0, 0, 0), dims)
        assert same_bits(b.a, a.a), op
    a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
    o.copy_range(a, s1, (0, 0, 0), (248, 248, 248), (8, 8, 8))
    gpu.copy_range(b, s1, (0, 0, 0), (248, 248, 248), (8, 8, 8))
    assert a.a.tobytes() == b.a.tobytes()
    # src and dst rows in the same non-zero 16-byte phase (uniform-phase row kernel, partial head and tail chunks)
    a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
    o.copy_range(a, s1, (9, 5, 7), (254, 250, 251), (1, 2, 3))
    gpu.copy_range(b, s1, (9, 5, 7), (254, 250, 251), (1, 2, 3))
    assert a.a.tobytes() == b.a.tobytes()
    a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
    o.arith_range(po.SAFEDIFF, a, s1, s2, (3, 5, 7), (255, 254, 253), (8, 1, 2))
    gpu.arith_range(po.SAFEDIFF, b, s1, s2, (3, 5, 7), (255, 254, 253), (8, 1, 2))
    assert same_bits(b.a, a.a)
    t1, t2, u1, u2 = cases.mk(dims, fmt, 4), cases.mk(dims, fmt, 5), cases.mk(dims, fmt, 4), cases.mk(dims, fmt, 5)
    for op in (0, 1):
        o.transform_range2(t1, t2, (3, 5, 7), (255, 254, 253), op)
        gpu.transform_range2(u1, u2, (3, 5, 7), (255, 254, 253), op)
        assert same_bits(u1.a, t1.a) and same_bits(u2.a, t2.a), op
    for op in range(4):
        o.transform_range1(t1, (5, 1, 2), (250, 255, 250), op)
        gpu.transform_range1(u1, (5, 1, 2), (250, 255, 250), op)
        assert same_bits(u1.a, t1.a), op
    # range with odd offsets (unaligned rows) and in-place use
    a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
    o.arith_range(po.SAFESUM, a, s1, s2, (3, 5, 7), (255, 254, 253), (1, 2, 3))
    gpu.arith_range(po.SAFESUM, b, s1, s2, (3, 5, 7), (255, 254, 253), (1, 2, 3))
    assert same_bits(b.a, a.a)


def test_range_ops_shifted_rows(gpu):
    """*Range calls whose source rows sit at another 16-byte phase than the destination rows (funnel-shifted 128-bit
    loads): pitches that are / are not multiples of 16 bytes, ranges that end on the very last voxel of the source
    (the one chunk whose second vector would leave the buffer), clamped Copy borders, in-place use."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32):
        for sdims, ddims in (((64, 16, 8), (96, 24, 12)), ((67, 13, 9), (83, 21, 15)), ((48, 5, 4), (48, 5, 4))):
            s1, s2 = cases.mk(sdims, fmt, 1), cases.mk(sdims, fmt, 2)
            X, Y, Z = sdims
            for first, last, off in (((0, 0, 0), sdims, (0, 0, 0)), ((1, 0, 0), sdims, (0, 0, 0)),
                                     ((3, 1, 2), sdims, (2, 3, 1)), ((5, 2, 1), (X - 1, Y, Z), (7, 0, 2)),
                                     ((0, 0, 0), (X - 3, Y - 1, Z), (13, 1, 0))):
                fits = all(l + f_ - f_ + o_ <= d for f_, l, o_, d in zip(first, last, off, ddims))
                if not fits or any(f_ + o_ < 0 for f_, o_ in zip(first, off)):
                    continue
                a, b = cases.mk(ddims, fmt, 3), cases.mk(ddims, fmt, 3)
                for op in (po.SAFESUM, po.DIFF):
                    o.arith_range(op, a, s1, s2, first, last, off)
                    gpu.arith_range(op, b, s1, s2, first, last, off)
                    assert same_bits(b.a, a.a), (fmt, sdims, first, last, off, op)
                a, b = cases.mk(ddims, fmt, 3), cases.mk(ddims, fmt, 3)
                cl = tuple(l - f_ for f_, l in zip(first, last))   # Copy writes dst[p - first + off]
                if all(c + o_ <= d for c, o_, d in zip(cl, off, ddims)):
                    o.copy_range(a, s1, first, last, off)
                    gpu.copy_range(b, s1, first, last, off)
                    assert a.a.tobytes() == b.a.tobytes(), (fmt, sdims, first, last, off)
            # clamped borders: first < 0 and last > dims, dst offset at an odd phase
            a, b = cases.mk(ddims, fmt, 3), cases.mk(ddims, fmt, 3)
            first, last = (-3, -1, -2), (X + 2, Y + 1, Z + 1)
            if all(l - f_ + 1 <= d for f_, l, d in zip(first, last, ddims)):
                o.copy_range(a, s1, first, last, (1, 0, 0))
                gpu.copy_range(b, s1, first, last, (1, 0, 0))
                assert a.a.tobytes() == b.a.tobytes(), (fmt, sdims, "clamped")


def test_range_ops_shifted_rows_wide(gpu):
    """Rows wide enough for whole warps (the neighbour-lane exchange kernel): every byte phase between source and
    destination rows, ranges that start inside the row and end on its last voxel, row counts that leave the last
    pair of rows half empty, sources of different widths."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32):
        s1, s2 = cases.mk((528, 5, 3), fmt, 1), cases.mk((544, 6, 3), fmt, 2)
        for xo in range(0, 17):
            for first, last in (((0, 0, 0), (528, 5, 3)), ((3, 1, 0), (528, 4, 3)), ((16, 0, 1), (509, 5, 2))):
                a, b = cases.mk((560, 7, 4), fmt, 3), cases.mk((560, 7, 4), fmt, 3)
                off = (xo, 2, 1)
                o.arith_range(po.SAFESUM, a, s1, s2, first, last, off)
                gpu.arith_range(po.SAFESUM, b, s1, s2, first, last, off)
                assert same_bits(b.a, a.a), (fmt, first, last, off)
                a, b = cases.mk((560, 7, 4), fmt, 3), cases.mk((560, 7, 4), fmt, 3)
                o.copy_range(a, s1, first, last, off)
                gpu.copy_range(b, s1, first, last, off)
                assert a.a.tobytes() == b.a.tobytes(), (fmt, first, last, off, "copy")
    # long rows, several CTAs along x
    s1, s2 = cases.mk((4112, 3, 2), po.UINT16, 1), cases.mk((4112, 3, 2), po.UINT16, 2)
    for off in ((1, 0, 0), (5, 0, 0), (7, 0, 0)):
        a, b = cases.mk((4128, 3, 2), po.UINT16, 3), cases.mk((4128, 3, 2), po.UINT16, 3)
        o.arith_range(po.DIFF, a, s1, s2, (0, 0, 0), (4112, 3, 2), off)
        gpu.arith_range(po.DIFF, b, s1, s2, (0, 0, 0), (4112, 3, 2), off)
        assert same_bits(b.a, a.a), off
        a, b = cases.mk((4128, 3, 2), po.UINT16, 3), cases.mk((4128, 3, 2), po.UINT16, 3)
        o.copy_range(a, s1, (0, 0, 0), (4112, 3, 2), off)
        gpu.copy_range(b, s1, (0, 0, 0), (4112, 3, 2), off)
        assert a.a.tobytes() == b.a.tobytes(), (off, "copy")


def test_range_ops_walking_rows(gpu, monkeypatch):
    """The row kernels with a small grid (VKTB_MAX_GRID_Y): every thread walks many rows, its row pointers cross
    plane boundaries (y overflow) several times, and the two-rows-in-flight pipeline drains on a half-empty pair."""
    o = po.oracle()
    for cap in ("3", "7"):
        monkeypatch.setenv("VKTB_MAX_GRID_Y", cap)
        for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
            s1, s2 = cases.mk((272, 37, 29), fmt, 1), cases.mk((272, 37, 29), fmt, 2)
            for first, last, off in (((1, 2, 0), (270, 37, 29), (4, 1, 2)), ((0, 0, 0), (272, 37, 29), (0, 3, 1)),
                                     ((0, 0, 3), (272, 36, 28), (9, 0, 0))):
                a, b = cases.mk((288, 40, 31), fmt, 3), cases.mk((288, 40, 31), fmt, 3)
                o.arith_range(po.SAFEDIFF, a, s1, s2, first, last, off)
                gpu.arith_range(po.SAFEDIFF, b, s1, s2, first, last, off)
                assert same_bits(b.a, a.a), (cap, fmt, first, last, off)
                a, b = cases.mk((288, 40, 31), fmt, 3), cases.mk((288, 40, 31), fmt, 3)
                o.copy_range(a, s1, first, last, off)
                gpu.copy_range(b, s1, first, last, off)
                assert a.a.tobytes() == b.a.tobytes(), (cap, fmt, first, last, off, "copy")


def test_scan_kernels(gpu):
    """ScanRange (3-D prefix sum of the raw voxel integers): the catalogue's directions / offsets / in-place cases,
    then shapes that take the warp-per-row x pass and the 16-byte column walk (pitches multiples of 16 bytes), ranges
    at odd phases (partial head / tail chunks), and sizes whose sums wrap many times."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32):
        for ddims, sdims, first, last, off, in_place in cases.SCAN_CASES:
            src = cases.mk(sdims, fmt, 2, -1000.0, 3000.0)
            if in_place:
                a, b = src.copy(), src.copy()
                o.scan_range(a, a, first, last, off)
                gpu.scan_range(b, b, first, last, off)
            else:
                a, b = cases.mk(ddims, fmt, 5, -1000.0, 3000.0), cases.mk(ddims, fmt, 5, -1000.0, 3000.0)
                o.scan_range(a, src, first, last, off)
                gpu.scan_range(b, src, first, last, off)
            assert a.a.tobytes() == b.a.tobytes(), (fmt, ddims, first, last, off, in_place)
        for dims, first, last in (((128, 40, 24), (0, 0, 0), (128, 40, 24)), ((128, 40, 24), (3, 1, 2), (125, 39, 23)),
                                  ((1040, 6, 5), (0, 0, 0), (1040, 6, 5)), ((1040, 6, 5), (7, 1, 0), (1033, 6, 4)),
                                  ((67, 13, 9), (0, 0, 0), (67, 13, 9)), ((67, 13, 9), (5, 2, 1), (66, 12, 9)),
                                  ((64, 32, 16), (63, 31, 15), (-1, -1, -1)), ((64, 32, 16), (2, 31, 3), (60, 0, 14))):
            src = cases.mk(dims, fmt, 4)
            a, b = src.copy(), src.copy()
            o.scan_range(a, a, first, last)
            gpu.scan_range(b, b, first, last)
            assert a.a.tobytes() == b.a.tobytes(), (fmt, dims, first, last)
            if all(f < l for f, l in zip(first, last)):
                a, b = cases.mk(dims, fmt, 6), cases.mk(dims, fmt, 6)   # dst != src: copy, then scan
                o.scan_range(a, src, first, last)
                gpu.scan_range(b, src, first, last)
                assert a.a.tobytes() == b.a.tobytes(), (fmt, dims, first, last, "copy")
    # the fused x+y pass (one CTA per plane; needs at least two planes per SM): one, two and four vectors per thread,
    # rows shorter than the CTA, a range inside the volume that keeps the 16-byte phase, z and y walked downwards
    for fmt, dims, first, last in ((po.UINT16, (64, 9, 320), (0, 0, 0), (64, 9, 320)),
                                   (po.UINT16, (2048, 5, 300), (0, 0, 0), (2048, 5, 300)),
                                   (po.UINT8, (4112, 3, 300), (0, 0, 0), (4112, 3, 300)),
                                   (po.UINT8, (8208, 3, 300), (0, 0, 0), (8208, 3, 300)),
                                   (po.FLOAT32, (1032, 4, 310), (0, 0, 0), (1032, 4, 310)),
                                   (po.UINT16, (64, 12, 330), (8, 2, 5), (56, 11, 325)),
                                   (po.UINT32, (32, 12, 330), (4, 10, 320), (28, 1, 3))):
        src = cases.mk(dims, fmt, 4)
        a, b = src.copy(), src.copy()
        o.scan_range(a, a, first, last)
        gpu.scan_range(b, b, first, last)
        assert a.a.tobytes() == b.a.tobytes(), ("fused xy", fmt, dims, first, last)
    # converting copy in front of the scan (UInt16 source into a UInt8 / Float32 destination)
    src = cases.mk((40, 12, 9), po.UINT16, 1)
    for dfmt in (po.UINT8, po.FLOAT32):
        a, b = cases.mk((48, 16, 12), dfmt, 3, -1.0, 2.0), cases.mk((48, 16, 12), dfmt, 3, -1.0, 2.0)
        o.scan_range(a, src, (1, 0, 2), (40, 12, 9), (4, 3, 1))
        gpu.scan_range(b, src, (1, 0, 2), (40, 12, 9), (4, 3, 1))
        assert a.a.tobytes() == b.a.tobytes(), dfmt


@pytest.mark.parametrize("sweep,n", [("sweep_fill_transform", 1000), ("sweep_copy_range", 1000), ("sweep_arithmetic", 1200),
                                     ("sweep_resample", 1000), ("sweep_resample_aligned", 120), ("sweep_flip_scan_histogram_bricks", 1000),
                                     ("sweep_scale_rotate_filter", 600)])
def test_random_sweep_against_the_oracle(gpu, sweep, n):
    """tests/fuzz_sweeps.py through the C-ABI: the same seeded random cases that pin the restatement on the compiled
    reference (tests/test_oracle_fuzz.py), CUDA path against restatement -- including the corners the reference leaves
    undefined and the restatement defines (dropped histogram voxels, the tap one voxel past the source buffer)."""
    import fuzz_sweeps as fz
    getattr(fz, sweep)(po.oracle(), gpu, n, False)


def test_brick_decompose_batch(gpu):
    """BrickDecompose as one vktb_copy_range_batch launch: cropped border bricks, halos clamped at the volume border,
    brick rows at every 16-byte phase; plus a batch that mixes bytewise and converting boxes and empty boxes."""
    o = po.oracle()
    extra = [((130, 70, 33), (64, 32, 16), (1, 1, 1), (1, 1, 1)), ((96, 64, 40), (32, 32, 32), (0, 0, 0), (0, 0, 0)),
             ((200, 9, 5), (67, 4, 2), (3, 0, 1), (2, 2, 0))]
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        for sdims, brick, neg, pos in cases.BRICK_CASES + extra:
            v = cases.mk(sdims, fmt, 3, -1000.0, 3000.0)
            want = o.brick_decompose(v, brick, neg, pos)
            got = gpu.brick_decompose(v, brick, neg, pos)
            assert len(want) == len(got)
            for i, (a, b) in enumerate(zip(want, got)):
                assert a.shape == b.shape and a.tobytes() == b.tobytes(), (fmt, sdims, brick, neg, pos, i)
    # mixed batch: same-format boxes (batched), a converting box (own launch), an empty box, a dst offset
    src = cases.mk((40, 20, 12), po.UINT16, 1)
    specs = [(po.UINT16, (24, 12, 8), (3, 2, 1), (23, 12, 8), (1, 0, 0)), (po.UINT8, (24, 12, 8), (0, 0, 0), (24, 12, 8), (0, 0, 0)),
             (po.UINT16, (8, 8, 8), (5, 5, 5), (5, 9, 9), (0, 0, 0)), (po.UINT16, (50, 25, 14), (-4, -2, -1), (44, 22, 13), (1, 1, 0)),
             (po.FLOAT32, (10, 10, 10), (30, 10, 2), (40, 20, 12), (0, 0, 0))]
    want = [cases.mk(d, f, 9 + i) for i, (f, d, *_r) in enumerate(specs)]
    got = [w.copy() for w in want]
    for w, (_f, _d, first, last, off) in zip(want, specs):
        o.copy_range(w, src, first, last, off)
    gpu.copy_range_batch(got, src, [s[2] for s in specs], [s[3] for s in specs], [s[4] for s in specs])
    for i, (w, g) in enumerate(zip(want, got)):
        assert same_bits(g.a, w.a), i


def test_brick_decompose_tma_staged_and_direct(gpu, monkeypatch):
    """The TMA-staged brick kernel (copy_bricks.cu) and the direct-load kernel it replaces, both against the oracle:
    sources the tensor map can describe (row pitch a multiple of 16 bytes), halos wider than a vector and wider than
    the brick, bricks cropped at the far border, halo voxels clamped on all six sides, rows shorter than a vector,
    planes-per-CTA counts that do and do not divide the brick depth."""
    o = po.oracle()
    cfgs = [((64, 48, 40), (32, 16, 16), (1, 1, 1), (1, 1, 1)), ((64, 48, 40), (24, 20, 12), (5, 3, 2), (9, 4, 7)),
            ((32, 16, 12), (8, 8, 4), (12, 10, 6), (11, 9, 5)), ((48, 8, 70), (48, 8, 66), (2, 2, 2), (2, 2, 2)),
            ((16, 6, 9), (3, 2, 4), (0, 1, 0), (1, 0, 2)), ((256, 4, 3), (250, 4, 3), (3, 0, 0), (3, 0, 0))]
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        for sdims, brick, neg, pos in cfgs:
            v = cases.mk(sdims, fmt, 7, -1000.0, 3000.0)
            want = o.brick_decompose(v, brick, neg, pos)
            for env in ({}, {"VKTB_BRICK_KZ": "1"}, {"VKTB_BRICK_KZ": "5"}, {"VKTB_BRICK_NO_TMA": "1"}):
                for k in ("VKTB_BRICK_KZ", "VKTB_BRICK_NO_TMA"):
                    monkeypatch.delenv(k, raising=False)
                for k, val in env.items():
                    monkeypatch.setenv(k, val)
                got = gpu.brick_decompose(v, brick, neg, pos)
                assert len(want) == len(got)
                for i, (a, b) in enumerate(zip(want, got)):
                    assert a.shape == b.shape and a.tobytes() == b.tobytes(), (fmt, sdims, brick, neg, pos, env, i)


def test_transform_flat_and_rows(gpu):
    """Transform over contiguous ranges (flat stream kernel: whole volumes, whole planes, part of one plane) and over
    boxed ranges (row kernel); odd row lengths make vectors straddle rows and misalign the range start."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        for dims in ((67, 13, 9), (64, 32, 8), (5, 3, 2), (1, 1, 1), (1, 7, 300)):
            X, Y, Z = dims
            ranges = [((0, 0, 0), dims), ((0, 0, Z // 2), dims), ((0, Y // 3, Z - 1), (X, Y, Z)),
                      ((0, 0, 0), (X, Y, max(1, Z - 1))), ((X // 2, 0, 0), dims), ((0, 1 % Y, 0), (X, Y, Z))]
            for first, last in ranges:
                for op in range(4):
                    a, b = cases.mk(dims, fmt, 5), cases.mk(dims, fmt, 5)
                    o.transform_range1(a, first, last, op)
                    gpu.transform_range1(b, first, last, op)
                    assert same_bits(b.a, a.a), (fmt, dims, first, last, op)


def test_resample_shapes(gpu):
    o = po.oracle()
    for sfmt, dfmt in ((po.UINT16, po.UINT16), (po.UINT8, po.UINT16), (po.FLOAT32, po.FLOAT32), (po.UINT16, po.UINT8)):
        src = cases.mk((128, 96, 80), sfmt, 1)
        for ddims in ((64, 48, 40), (192, 144, 120), (100, 37, 53), (128, 96, 80)):
            for filt in (po.NEAREST, po.LINEAR, po.TRILINEAR):
                a, b = po.empty(ddims, dfmt, -1.0, 2.0), po.empty(ddims, dfmt, -1.0, 2.0)
                o.resample(a, src, filt)
                gpu.resample(b, src, filt)
                assert same_bits(b.a, a.a), (sfmt, dfmt, ddims, filt)


def test_resample_float_specials_linear(gpu):
    """Float32 sources keep the 8-tap path: NaN/Inf/-0 taps propagate exactly as in the reference."""
    o = po.oracle()
    src = cases.mk((32, 24, 20), po.FLOAT32, 1)
    flat = src.a.ravel()
    flat[::97] = np.nan
    flat[5::131] = np.inf
    flat[7::89] = -0.0
    for ddims in ((48, 36, 30), (16, 12, 10)):
        a, b = po.empty(ddims, po.FLOAT32), po.empty(ddims, po.FLOAT32)
        o.resample(a, src, po.LINEAR)
        gpu.resample(b, src, po.LINEAR)
        assert same_bits(b.a, a.a), ddims


def test_resample_halve_x_kernel(gpu):
    """8-tap modes with an exact 2:1 x map (coalesced four-row gather): every format, y / z ratios of any kind,
    Float32 specials, and a dst width whose x map is NOT exact (falls back to the other kernels)."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32):
        src = cases.mk((96, 40, 24), fmt, 1)
        if fmt == po.FLOAT32:
            flat = src.a.ravel()
            flat[::53] = np.nan
            flat[7::71] = np.inf
            flat[3::67] = -0.0
        for ddims in ((48, 20, 12), (48, 60, 36), (48, 33, 7), (48, 40, 24)):
            for filt in (po.LINEAR, po.TRILINEAR):
                a, b = po.empty(ddims, fmt, -1.0, 2.0), po.empty(ddims, fmt, -1.0, 2.0)
                o.resample(a, src, filt)
                gpu.resample(b, src, filt)
                assert same_bits(b.a, a.a), (fmt, ddims, filt)
    # 2:1 with a width for which (x / float(Dx)) * Sx dips below 2x somewhere: the host check must reject it
    for Dx in (16, 24, 40, 56, 80, 176):
        src = cases.mk((2 * Dx, 6, 5), po.FLOAT32, 2)
        a, b = po.empty((Dx, 3, 5), po.FLOAT32), po.empty((Dx, 3, 5), po.FLOAT32)
        o.resample(a, src, po.LINEAR)
        gpu.resample(b, src, po.LINEAR)
        assert same_bits(b.a, a.a), Dx


def test_resample_zslab_windows(gpu):
    """z-slab sharding on one device ("virtual ranks"): slabs computed one by one equal the whole."""
    from volkit_b200 import capi
    o = po.oracle()
    S, D = (40, 30, 32), (60, 45, 48)
    src = cases.mk(S, po.UINT16, 1)
    for filt in (po.NEAREST, po.LINEAR, po.TRILINEAR):
        whole = po.empty(D, po.UINT16)
        o.resample(whole, src, filt)
        for n in (2, 3, 4):
            for g in range(n):
                dz0, dz1 = g * D[2] // n, (g + 1) * D[2] // n
                a, b = ctypes.c_int32(), ctypes.c_int32()
                capi.check(gpu.lib.vktb_resample_window_src_range(filt, dz0, dz1, D[2], S[2], ctypes.byref(a), ctypes.byref(b)))
                win = po.Vol(src.a[a.value:b.value].copy(), po.UINT16)
                part = po.empty((D[0], D[1], dz1 - dz0), po.UINT16)
                gpu.resample_window(part, win, filt, dz0, D, a.value, S)
                assert part.a.tobytes() == whole.a[dz0:dz1].tobytes(), (filt, n, g)


def test_resample_row_pipeline(gpu):
    """One-tap Resample through the TMA row pipeline (rows 16-byte aligned): several x tiles, R > 1 row groups,
    long y/z runs (8x up-sampling), pure-move and converting variants, and z-slab windows."""
    from volkit_b200 import capi
    o = po.oracle()
    U8, U16, U32, F32 = po.UINT8, po.UINT16, po.UINT32, po.FLOAT32
    shapes = [
        ((5000, 6, 5), (8200, 9, 7), U16, U16, (0.0, 1.0), (0.0, 1.0)),     # 3 x tiles, move
        ((8200, 5, 4), (5000, 3, 3), U16, U16, (0.0, 1.0), (0.0, 1.0)),     # down, 2 x tiles
        ((2056, 4, 4), (3080, 6, 6), F32, F32, (0.0, 1.0), (0.0, 1.0)),     # f32 move, 2 x tiles
        ((16, 8, 8), (128, 64, 64), U16, U16, (0.0, 1.0), (0.0, 1.0)),      # 8x up: 64 stores per src row
        ((256, 64, 48), (64, 16, 12), F32, U16, (0.0, 1.0), (-1.0, 2.0)),   # 4x down, converting, R = 8
        ((128, 40, 24), (192, 60, 36), U16, U16, (-1000.0, 3000.0), (0.0, 1.0)),
        ((128, 40, 24), (192, 60, 36), U8, U32, (0.0, 1.0), (0.0, 1.0)),
        ((64, 40, 24), (96, 60, 36), U32, U8, (0.0, 1.0), (0.25, 0.75)),
        ((64, 33, 17), (32, 50, 40), U8, U8, (0.0, 1.0), (0.0, 1.0)),       # mixed up/down per axis
    ]
    for S, D, sfmt, dfmt, smap, dmap in shapes:
        src = cases.mk(S, sfmt, 1, *smap)
        for filt in (po.NEAREST, po.LINEAR):
            if filt == po.LINEAR and sfmt == F32:
                continue   # Float32 Linear keeps the 8-tap kernel (covered elsewhere)
            a, b = cases.mk(D, dfmt, 3, *dmap), cases.mk(D, dfmt, 3, *dmap)
            o.resample(a, src, filt)
            gpu.resample(b, src, filt)
            assert same_bits(b.a, a.a), (S, D, sfmt, dfmt, filt)
    # z-slab windows on aligned rows
    S, D = (64, 30, 32), (96, 45, 48)
    src = cases.mk(S, U16, 1)
    whole = po.empty(D, U16)
    o.resample(whole, src, po.NEAREST)
    for n in (2, 3, 5):
        for g in range(n):
            dz0, dz1 = g * D[2] // n, (g + 1) * D[2] // n
            a, b = ctypes.c_int32(), ctypes.c_int32()
            capi.check(gpu.lib.vktb_resample_window_src_range(po.NEAREST, dz0, dz1, D[2], S[2], ctypes.byref(a), ctypes.byref(b)))
            win = po.Vol(src.a[a.value:b.value].copy(), U16)
            part = po.empty((D[0], D[1], dz1 - dz0), U16)
            gpu.resample_window(part, win, po.NEAREST, dz0, D, a.value, S)
            assert part.a.tobytes() == whole.a[dz0:dz1].tobytes(), (n, g)


def test_resample_march_kernel(gpu):
    """8-tap Resample through the TMA z-march (16-byte aligned src rows): several tiles in x and y, multi-box src bricks
    (3:1 down-sampling), the reference's unclamped hi.x spill column, NaN / Inf / -0 taps, z-slab windows."""
    from volkit_b200 import capi
    o = po.oracle()
    U8, U16, U32, F32 = po.UINT8, po.UINT16, po.UINT32, po.FLOAT32
    shapes = [
        ((200, 70, 40), (300, 100, 64), F32, F32, (0.0, 1.0), (0.0, 1.0), (po.LINEAR, po.TRILINEAR)),
        ((256, 64, 48), (128, 32, 24), F32, F32, (0.0, 1.0), (0.0, 1.0), (po.LINEAR, po.TRILINEAR)),    # exact 2:1
        ((264, 50, 30), (88, 20, 12), U16, U16, (0.0, 1.0), (-1.0, 2.0), (po.TRILINEAR,)),              # 3:1, 2 boxes in x
        ((768, 20, 10), (256, 7, 5), U8, U8, (0.0, 1.0), (0.0, 1.0), (po.TRILINEAR,)),                   # 3:1 UInt8, 4 boxes
        ((64, 40, 24), (200, 130, 70), U16, U8, (-1000.0, 3000.0), (0.0, 1.0), (po.TRILINEAR,)),        # 3x up, converting
        ((64, 33, 17), (36, 50, 40), F32, U16, (0.0, 1.0), (0.0, 1.0), (po.LINEAR, po.TRILINEAR)),      # mixed per axis
        ((48, 20, 12), (64, 30, 18), U32, F32, (0.0, 1.0), (0.0, 1.0), (po.TRILINEAR,)),
    ]
    for S, D, sfmt, dfmt, smap, dmap, filters in shapes:
        src = cases.mk(S, sfmt, 1, *smap)
        if sfmt == F32:
            flat = src.a.ravel()
            flat[::211] = np.nan
            flat[3::173] = np.inf
            flat[5::97] = -0.0
            flat[-1] = np.nan          # the tap the unclamped hi.x reads at the very last voxel's row end
        for filt in filters:
            a, b = cases.mk(D, dfmt, 3, *dmap), cases.mk(D, dfmt, 3, *dmap)
            o.resample(a, src, filt)
            gpu.resample(b, src, filt)
            assert same_bits(b.a, a.a), (S, D, sfmt, dfmt, filt)
    # z-slab windows, Float32 Linear (8 taps) and true trilinear
    S, D = (64, 30, 32), (96, 45, 48)
    src = cases.mk(S, F32, 1)
    src.a.ravel()[::101] = np.inf
    for filt in (po.LINEAR, po.TRILINEAR):
        whole = po.empty(D, F32)
        o.resample(whole, src, filt)
        for n in (2, 3):
            for g in range(n):
                dz0, dz1 = g * D[2] // n, (g + 1) * D[2] // n
                a, b = ctypes.c_int32(), ctypes.c_int32()
                capi.check(gpu.lib.vktb_resample_window_src_range3(filt, dz0, dz1, capi.I3(*D), capi.I3(*S), ctypes.byref(a),
                                                                   ctypes.byref(b)))
                win = po.Vol(src.a[a.value:b.value].copy(), F32)
                part = po.empty((D[0], D[1], dz1 - dz0), F32)
                gpu.resample_window(part, win, filt, dz0, D, a.value, S)
                assert same_bits(part.a, whole.a[dz0:dz1]), (filt, n, g)


@pytest.fixture()
def hosted(monkeypatch):
    from gpu_adapter import HostedLib
    monkeypatch.setenv("VKTB_HOSTED_SLAB_KB", "24")     # a few planes per slab: every call runs many pipeline stages
    return HostedLib()


def test_hosted_arith_streams_slabs(gpu, hosted):
    """Host-resident volumes streamed through the device in z-slabs give the bytes of the whole-volume call."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        dims = (48, 40, 37)
        s1, s2 = cases.mk(dims, fmt, 1), cases.mk(dims, fmt, 2)
        for op in (po.SAFESUM, po.DIFF, po.SAFEPROD):
            a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
            o.arith_range(op, a, s1, s2, (0, 0, 0), dims)
            hosted.arith_range(op, b, s1, s2, (0, 0, 0), dims)
            assert same_bits(b.a, a.a), (fmt, op)
        # sub-box with a dst offset: the untouched dst voxels of every plane must survive the round trip
        a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
        o.arith_range(po.SAFESUM, a, s1, s2, (3, 5, 7), (40, 33, 30), (2, 1, 4))
        hosted.arith_range(po.SAFESUM, b, s1, s2, (3, 5, 7), (40, 33, 30), (2, 1, 4))
        assert same_bits(b.a, a.a), fmt
        # in place: dst is source 1
        a, b = s1.copy(), s1.copy()
        o.arith_range(po.SAFEDIFF, a, a, s2, (0, 0, 0), dims)
        hosted.arith_range(po.SAFEDIFF, b, b, s2, (0, 0, 0), dims)
        assert same_bits(b.a, a.a), fmt
    # mixed formats, non-default mappings
    d1, d2 = po.empty((32, 24, 20), po.UINT16, -1.0, 3.0), po.empty((32, 24, 20), po.UINT16, -1.0, 3.0)
    s1, s2 = cases.mk((32, 24, 20), po.UINT8, 1), cases.mk((32, 24, 20), po.FLOAT32, 2)
    o.arith_range(po.SAFEPROD, d1, s1, s2, (0, 0, 0), (32, 24, 20))
    hosted.arith_range(po.SAFEPROD, d2, s1, s2, (0, 0, 0), (32, 24, 20))
    assert same_bits(d2.a, d1.a)


def test_hosted_transform_and_resample(gpu, hosted):
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        dims = (40, 24, 50)
        for first, last in (((0, 0, 0), dims), ((3, 2, 5), (37, 20, 41))):
            for op in range(4):
                a, b = cases.mk(dims, fmt, 5), cases.mk(dims, fmt, 5)
                o.transform_range1(a, first, last, op)     # ops read absolute z: slabs must not renumber planes
                hosted.transform_range1(b, first, last, op)
                assert same_bits(b.a, a.a), (fmt, first, op)
    for sfmt, dfmt in ((po.UINT16, po.UINT16), (po.FLOAT32, po.FLOAT32), (po.UINT8, po.UINT16)):
        src = cases.mk((64, 48, 40), sfmt, 1)
        for ddims in ((32, 24, 20), (96, 72, 60), (50, 37, 53), (64, 48, 40)):
            for filt in (po.NEAREST, po.LINEAR, po.TRILINEAR):
                a, b = po.empty(ddims, dfmt, -1.0, 2.0), po.empty(ddims, dfmt, -1.0, 2.0)
                o.resample(a, src, filt)
                hosted.resample(b, src, filt)
                assert same_bits(b.a, a.a), (sfmt, dfmt, ddims, filt)


def test_hosted_rejects_bad_calls(gpu, hosted):
    from volkit_b200 import capi
    v = cases.mk((16, 16, 16), po.UINT8, 1)
    with pytest.raises(RuntimeError):
        hosted.arith_range(po.SUM, v, v, v, (0, 0, 0), (16, 16, 17))        # range outside the volumes
    with pytest.raises(RuntimeError):
        hosted.arith_range(po.SUM, v, v, v, (0, 0, 0), (16, 16, 8), (0, 0, 8))   # in place with an offset
    assert capi.load().vktb_last_error()


def test_flip_kernels(gpu):
    """Pair kernel (X axis, unaligned rows) and the 128-bit row-swap kernel (Y / Z with 16-byte aligned rows)."""
    o = po.oracle()
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        for dims, first, last, off, ddims in (((128, 40, 24), (0, 0, 0), (128, 40, 24), (0, 0, 0), (128, 40, 24)),
                                              ((128, 40, 24), (16, 3, 2), (112, 37, 21), (16, 2, 1), (160, 48, 30)),
                                              ((67, 21, 10), (1, 2, 3), (66, 20, 9), (3, 1, 2), (80, 30, 20))):
            for axis in range(3):
                src = cases.mk(dims, fmt, 1)
                a, b = cases.mk(ddims, fmt, 3), cases.mk(ddims, fmt, 3)
                o.flip_range(a, src, first, last, off, axis)
                gpu.flip_range(b, src, first, last, off, axis)
                assert a.a.tobytes() == b.a.tobytes(), (fmt, dims, axis)
                if dims == ddims:
                    a, b = src.copy(), src.copy()
                    o.flip_range(a, a, first, last, off, axis)
                    gpu.flip_range(b, b, first, last, off, axis)
                    assert a.a.tobytes() == b.a.tobytes(), ("in place", fmt, axis)
    with pytest.raises(RuntimeError):
        gpu.flip_range(cases.mk((8, 8, 8), po.UINT16, 1), cases.mk((8, 8, 8), po.UINT8, 1), (0, 0, 0), (8, 8, 8))


def test_histogram_kernels(gpu):
    """Shared-memory counters (bins fit in shared memory) and global atomics (they do not); dropped voxels."""
    o = po.oracle()
    for fmt, lo, hi in ((po.UINT8, 0.0, 1.0), (po.UINT16, 0.0, 1.0), (po.UINT16, -1000.0, 3000.0), (po.UINT32, 0.0, 1.0),
                        (po.FLOAT32, 0.0, 1.0)):
        v = cases.mk((200, 60, 33), fmt, 7, lo, hi)
        for first, last in (((0, 0, 0), (200, 60, 33)), ((3, 5, 7), (199, 58, 30))):
            for nb in (1, 64, 256, 4096, 24576, 100000):
                ref, rd = o.histogram_range(v, first, last, nb)
                got, gd = gpu.histogram_range(v, first, last, nb)
                assert rd == gd == 0 and got.tobytes() == ref.tobytes(), (fmt, lo, first, nb)
    # contiguous runs of >= 2^20 voxels (flat path); odd row length + plane offset misalign the run's start; a constant
    # volume sends every voxel to one bin
    for fmt, lo, hi in ((po.UINT8, 0.0, 1.0), (po.UINT16, 0.0, 1.0), (po.UINT16, -1000.0, 3000.0), (po.FLOAT32, 0.0, 1.0)):
        v = cases.mk((251, 130, 41), fmt, 9, lo, hi)
        for first, last in (((0, 0, 0), (251, 130, 41)), ((0, 0, 3), (251, 130, 40))):
            for nb in (1, 100, 256):
                ref, rd = o.histogram_range(v, first, last, nb)
                got, gd = gpu.histogram_range(v, first, last, nb)
                assert rd == gd == 0 and got.tobytes() == ref.tobytes(), ("flat", fmt, lo, first, nb)
    v = po.Vol(np.full((64, 128, 256), 40000, np.uint16), po.UINT16)
    ref, _ = o.histogram_range(v, (0, 0, 0), (256, 128, 64), 256)
    got, _ = gpu.histogram_range(v, (0, 0, 0), (256, 128, 64), 256)
    assert got.tobytes() == ref.tobytes() and int(got.max()) == 256 * 128 * 64
    # Float32 voxels outside the mapping, NaN and Inf: the reference indexes out of bounds, both sides drop and count
    v = cases.mk((64, 32, 16), po.FLOAT32, 3, 0.25, 0.75)
    flat = v.a.ravel()
    flat[::101] = np.nan
    flat[5::97] = np.inf
    flat[7::89] = -3.0
    ref, rd = o.histogram_range(v, (0, 0, 0), (64, 32, 16), 128)
    got, gd = gpu.histogram_range(v, (0, 0, 0), (64, 32, 16), 128)
    assert rd == gd and rd > 0 and got.tobytes() == ref.tobytes()
    assert int(got.sum()) + gd == 64 * 32 * 16


def test_affine_and_filter_kernels(gpu):
    """Scale / Rotate / ApplyFilter at sizes that engage the vector stores, the z-march chunks (ragged last chunk, z
    chunk borders) and the one-tap / eight-tap / generic kernels; Float32 sources with NaN / Inf / -0 taps."""
    o = po.oracle()
    for sfmt, dfmt in ((po.UINT16, po.UINT16), (po.UINT8, po.UINT8), (po.FLOAT32, po.FLOAT32), (po.UINT16, po.FLOAT32),
                       (po.UINT32, po.UINT16), (po.INT16, po.INT16)):
        src = cases.mk((96, 40, 37), sfmt, 2, -3.0, 5.0)
        if sfmt == po.FLOAT32:
            src.a[3, 5, 7], src.a[10, 20, 95], src.a[36, 39, 95], src.a[0, 0, 0] = np.nan, np.inf, -np.inf, -0.0
        for ddims, op in (((96, 40, 37), ("scale", (1.25, 0.8, 1.5), (40.0, 20.0, 18.0))),
                          ((128, 33, 40), ("rotate", (0.3, -1.0, 2.0), 0.9, (48.0, 20.0, 18.0))),
                          ((96, 40, 37), ("scale", (1.0, 1.0, 1.0), (0.0, 0.0, 0.0))),
                          # every dst voxel maps inside the source: the bulk-copy row pipeline (integer sources)
                          ((96, 40, 37), ("scale", (1.5, 2.0, 1.25), (48.0, 20.0, 18.0))),
                          ((192, 80, 50), ("scale", (2.0, 2.0, 1.4), (0.0, 0.0, 0.0))),
                          ((64, 24, 30), ("scale", (0.75, 0.625, 0.9), (0.0, 0.0, 0.0))),
                          ((2048, 9, 5), ("scale", (21.5, 1.0, 1.0), (7.25, 3.0, 2.0)))):
            a, b = cases.mk(ddims, dfmt, 5, 0.0, 1.0), cases.mk(ddims, dfmt, 5, 0.0, 1.0)
            for lib, d in ((gpu, a), (o, b)):
                if op[0] == "scale":
                    lib.scale(d, src, op[1], op[2])
                else:
                    lib.rotate(d, src, op[1], op[2], op[3])
            assert same_bits(a.a, b.a), (sfmt, dfmt, ddims, op)
        for fd, seed in (((3, 3, 3), 0), ((3, 3, 3), 3), ((2, 3, 1), 4)):
            w = cases.filter_weights(fd, seed)
            for ident in (True, False):
                s2 = po.Vol(src.a, sfmt, 0.0, 1.0) if ident else src
                a, b = cases.mk((96, 40, 37), dfmt, 5, 0.0 if ident else -1.0, 1.0), None
                b = a.copy()
                gpu.apply_filter(a, s2, w); o.apply_filter(b, s2, w)
                assert same_bits(a.a, b.a), (sfmt, dfmt, fd, seed, ident)
        if sfmt == dfmt:
            a, b = src.copy(), src.copy()
            w = cases.filter_weights((3, 3, 3), 9)
            gpu.apply_filter(a, a, w); o.apply_filter(b, b, w)
            assert same_bits(a.a, b.a), (sfmt, "in place")
    # thin volumes: everything is border
    for dims in ((40, 2, 9), (1, 1, 1), (2, 2, 2), (3, 3, 3), (40, 9, 1)):
        src = cases.mk(dims, po.UINT16, 2)
        a, b = cases.mk(dims, po.UINT16, 5), cases.mk(dims, po.UINT16, 5)
        w = cases.filter_weights((3, 3, 3), 2)
        gpu.apply_filter(a, src, w); o.apply_filter(b, src, w)
        assert a.a.tobytes() == b.a.tobytes(), dims
    with pytest.raises(Exception):
        gpu.apply_filter(a, src, np.zeros((5, 3, 3), np.float32))


def test_stream_in_out_many_chunks(gpu):
    """vktb_stream_in / vktb_stream_out (what InputStream::read / OutputStream::write use under the GPU policy): 70 MiB
    + 5 bytes travel through the 16 MiB staging slots in both directions; the callbacks' return values are summed."""
    from volkit_b200 import capi
    lib = gpu.lib
    n = 70 * (1 << 20) + 5
    src = np.random.default_rng(1).integers(0, 256, n, dtype=np.uint8)
    dev = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(dev), n))
    pos, calls = [0], [0]

    def produce(ctx, buf, length):
        ctypes.memmove(buf, src.ctypes.data + pos[0], length)
        pos[0] += length
        calls[0] += 1
        return length

    total = ctypes.c_size_t(0)
    capi.check(lib.vktb_stream_in(dev, n, capi.READ_FN(produce), None, ctypes.byref(total)))
    assert total.value == n and pos[0] == n and calls[0] == 5
    out = np.zeros(n, np.uint8)
    wpos = [0]

    def consume(ctx, buf, length):
        ctypes.memmove(out.ctypes.data + wpos[0], buf, length)
        wpos[0] += length
        return 1                                    # an item-counting sink, like RawFile::write

    capi.check(lib.vktb_stream_out(dev, n, capi.WRITE_FN(consume), None, ctypes.byref(total)))
    assert total.value == 5 and wpos[0] == n and out.tobytes() == src.tobytes()
    # and the device really holds the bytes: an ordinary D2H agrees
    back = np.zeros(n, np.uint8)
    capi.check(lib.vktb_memcpy(back.ctypes.data, dev, n, capi.D2H))
    assert back.tobytes() == src.tobytes()
    capi.check(lib.vktb_free(dev))


def test_aggregates(gpu):
    """ComputeAggregatesRange: min/max/argmin/argmax bit-exact (first occurrence, ties, NaN skipped, -0 / +0), the
    accumulated statistics inside the reference's own rounding bound (cases.check_aggregates); flat and row kernels,
    packed-integer and float min/max paths, vectors + scalar heads and tails."""
    o = po.oracle()
    rng = np.random.default_rng(11)
    for fmt in cases.FMT4 + [po.INT16]:
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
            for dims, first, last in cases.AGG_CASES + [((70, 33, 20), (0, 0, 0), (70, 33, 20)),
                                                        ((70, 33, 20), (3, 1, 2), (69, 30, 19)),
                                                        ((128, 16, 9), (0, 0, 2), (128, 16, 7))]:
                v = cases.mk(dims, fmt, 4, lo, hi)
                cases.check_aggregates(o, gpu.aggregates_range(v, first, last), o.aggregates_range(v, first, last),
                                       v, first, last, (fmt, lo, dims, first))
    # ties: few distinct values, so the first occurrence decides argmin / argmax
    for fmt, top in ((po.UINT8, 3), (po.UINT16, 5), (po.UINT32, 2), (po.FLOAT32, 4)):
        for lo, hi in ((0.0, 1.0), (-2.0, 5.0)):
            a = rng.integers(0, top, size=(9, 17, 67)).astype(po.DTYPES[fmt])
            v = po.Vol(a, fmt, lo, hi)
            for first, last in (((0, 0, 0), (67, 17, 9)), ((2, 1, 1), (66, 16, 8))):
                cases.check_aggregates(o, gpu.aggregates_range(v, first, last), o.aggregates_range(v, first, last),
                                       v, first, last, ("ties", fmt, lo, first))
    # constant volume: min == max everywhere, argmin = argmax = first voxel of the range; var exactly 0 for a whole volume
    v = po.Vol(np.full((6, 10, 48), 777, np.uint16), po.UINT16)
    g = gpu.aggregates_range(v, (0, 0, 0), (48, 10, 6))
    cases.check_aggregates(o, g, o.aggregates_range(v, (0, 0, 0), (48, 10, 6)), v, (0, 0, 0), (48, 10, 6), "constant")
    # Float32 specials: NaN never enters min/max, -0 / +0 compare equal (first one's bits are kept), +-inf do enter
    f = rng.random((5, 9, 40), dtype=np.float32) * 2 - 1
    f[0, 0, 3] = np.nan; f[2, 4, 7] = np.nan
    for mn, mx in (((-0.0, 0.0), None), ((0.0, -0.0), None), (None, (np.inf, np.inf)), ((-np.inf, -np.inf), (np.inf, np.inf))):
        w = f.copy()
        if mn:
            w = np.where(w < 0, -w, w); w[1, 2, 5], w[3, 3, 9] = mn
        if mx:
            w[1, 6, 30], w[4, 1, 2] = mx
        v = po.Vol(w, po.FLOAT32)
        g, r = gpu.aggregates_range(v, (0, 0, 0), (40, 9, 5)), o.aggregates_range(v, (0, 0, 0), (40, 9, 5))
        assert cases.agg_exact_words(g).tobytes() == cases.agg_exact_words(r).tobytes(), (mn, mx)
    # no voxel qualifies: every voxel NaN -> the initial values
    v = po.Vol(np.full((2, 3, 16), np.nan, np.float32), po.FLOAT32)
    g, r = gpu.aggregates_range(v, (0, 0, 0), (16, 3, 2)), o.aggregates_range(v, (0, 0, 0), (16, 3, 2))
    assert cases.agg_exact_words(g).tobytes() == cases.agg_exact_words(r).tobytes()
    # empty range: the reference's initial values; first > last is rejected
    v = cases.mk((16, 4, 4), po.UINT16, 1)
    g, r = gpu.aggregates_range(v, (3, 1, 1), (3, 4, 4)), o.aggregates_range(v, (3, 1, 1), (3, 4, 4))
    assert g.words() == r.words()
    with pytest.raises(Exception):
        gpu.aggregates_range(v, (5, 1, 1), (3, 4, 4))


def test_aggregates_partials_combine(gpu):
    """z-slab partials folded by vktb_aggregates_finish equal the one-call result (exact fields identical, moments to
    rounding): the path aggregates_sharded() takes."""
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    lib = gpu.lib
    for fmt in (po.UINT16, po.FLOAT32):
        vol = cases.mk((50, 21, 24), fmt, 9, -3.0, 7.0)
        dv = DeviceVolume.from_numpy(vol.a, fmt, -3.0, 7.0)
        whole = capi.VktbAggregates()
        z3, l3 = capi.I3(0, 0, 0), capi.I3(50, 21, 24)
        capi.check(lib.vktb_aggregates_range(ctypes.byref(dv.desc()), z3, l3, ctypes.byref(whole)))
        parts = (capi.VktbAggPartial * 3)()
        for k, (z0, z1) in enumerate(((0, 7), (7, 16), (16, 24))):
            capi.check(lib.vktb_aggregates_partial(ctypes.byref(dv.desc()), capi.I3(0, 0, z0), capi.I3(50, 21, z1),
                                                   ctypes.byref(parts[k])))
        folded = capi.VktbAggregates()
        capi.check(lib.vktb_aggregates_finish(parts, 3, 50 * 21 * 24, ctypes.byref(folded)))
        for name in ("min", "max"):
            assert np.float32(getattr(folded, name)).tobytes() == np.float32(getattr(whole, name)).tobytes()
        assert list(folded.argmin) == list(whole.argmin) and list(folded.argmax) == list(whole.argmax)
        for name in ("sum", "mean", "var", "stddev"):
            a, b = getattr(folded, name), getattr(whole, name)
            assert abs(a - b) <= 4e-7 * max(abs(a), abs(b)), (name, a, b)
        dv.free()


def test_cabi_from_many_threads(gpu):
    """The C-ABI may be called from any thread (SURVEY 8(b) threading): four host threads run different algorithms on
    their own volumes at the same time (ctypes releases the GIL during the calls); every result must equal the oracle's."""
    import threading
    o = po.oracle()
    dims = (96, 64, 48)
    errors = []

    def worker(k):
        try:
            for rep in range(6):
                fmt = (po.UINT8, po.UINT16, po.FLOAT32, po.UINT16)[k]
                s1, s2 = cases.mk(dims, fmt, 10 + k), cases.mk(dims, fmt, 20 + k)
                a, b = cases.mk(dims, fmt, 3), cases.mk(dims, fmt, 3)
                if k == 0:
                    o.arith_range(po.SAFESUM, a, s1, s2, (0, 0, 0), dims)
                    gpu.arith_range(po.SAFESUM, b, s1, s2, (0, 0, 0), dims)
                elif k == 1:
                    o.copy_range(a, s1, (-2, 1, 3), (90, 60, 40), (1, 2, 3))
                    gpu.copy_range(b, s1, (-2, 1, 3), (90, 60, 40), (1, 2, 3))
                elif k == 2:
                    a, b = po.empty((144, 96, 72), fmt), po.empty((144, 96, 72), fmt)
                    o.resample(a, s1, po.LINEAR)
                    gpu.resample(b, s1, po.LINEAR)
                else:
                    o.transform_range1(a, (1, 2, 3), (95, 60, 40), rep % 4)
                    gpu.transform_range1(b, (1, 2, 3), (95, 60, 40), rep % 4)
                if not same_bits(b.a, a.a):
                    errors.append((k, rep))
        except Exception as e:                      # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_empty_and_degenerate_ranges(gpu):
    v = cases.mk((9, 7, 5), po.UINT16, 3)
    before = v.a.copy()
    gpu.fill_range(v, (3, 3, 3), (3, 5, 5), 0.5)          # empty in x
    gpu.copy_range(v, cases.mk((4, 4, 4), po.UINT16, 1), (0, 0, 0), (0, 4, 4))
    assert v.a.tobytes() == before.tobytes()
    one = po.empty((1, 1, 1), po.UINT8)
    gpu.fill_range(one, (0, 0, 0), (1, 1, 1), 0.5)
    assert int(one.a[0, 0, 0]) == 127


def test_arith_mapped_proven_division(gpu):
    """map_dst divides by hi - lo once per voxel.  From 2^22 voxels on, the flat kernel uses the reciprocal-and-two-FMA
    form for divisors it has PROVEN against __fdiv_rn on all 2^32 dividends (arith.cu proveDivisor).  Bit-exact against
    the oracle for several ranges (4000, 3, 0.7, 1e-3 .. 1e3, a negative one) and for ops whose results span zero,
    tiny, huge, Inf and NaN before the division (Quot by zero, Prod of large values)."""
    from concurrent.futures import ThreadPoolExecutor
    o = po.oracle()
    dims = (256, 256, 70)
    pool = ThreadPoolExecutor(min(16, os.cpu_count() or 1))
    slabs = [(z, min(z + 5, dims[2])) for z in range(0, dims[2], 5)]
    for fmt in (po.UINT16, po.FLOAT32):
        s1, s2 = po.synth(dims, fmt, 1), po.synth(dims, fmt, 2)
        if fmt == po.FLOAT32:
            s1 = (s1 * np.float32(4000.0) - np.float32(1000.0)).astype(np.float32)
            s2[::7, ::3, ::5] = 0.0                     # Quot by zero -> Inf / NaN in front of the division
        for (lo, hi), ops in (((-1000.0, 3000.0), (po.SAFEPROD, po.QUOT, po.SUM)), ((0.0, 3.0), (po.SAFEQUOT, po.PROD)),
                              ((0.1, 0.8), (po.DIFF,)), ((1e-3, 1e3), (po.QUOT,)), ((2.0, -5.0), (po.SAFESUM,))):
            a, b = po.Vol(s1, fmt, lo, hi), po.Vol(s2, fmt)
            for op in ops:
                want, got = po.empty(dims, fmt, lo, hi), po.empty(dims, fmt, lo, hi)

                def one(zz, op=op, want=want, a=a, b=b):
                    z0, z1 = zz
                    o.arith_range(op, po.Vol(want.a[z0:z1], fmt, lo, hi), po.Vol(a.a[z0:z1], fmt, lo, hi), po.Vol(b.a[z0:z1], fmt),
                                  (0, 0, 0), (dims[0], dims[1], z1 - z0))
                list(pool.map(one, slabs))
                gpu.arith_range(op, got, a, b, (0, 0, 0), dims)
                assert same_bits(got.a, want.a), (fmt, lo, hi, op)
    pool.shutdown()


def test_histogram_integer_bins_and_byte_values(gpu):
    """histograms of a few million voxels through the shortcuts of histogramKernel: identity-mapped UInt16 with up to
    256 bins takes (raw * numBins) >> 16, UInt8 counts byte values and bins each value's total once, with the exact float
    expression.  Exact against the oracle: random data, a constant volume (a whole warp hits one bin), non-identity
    mappings whose bins fall partly outside [0, numBins) (the reference writes out of bounds there: dropped and counted),
    odd bin counts, more than 256 bins, a range that starts off a 16-byte boundary (whole planes from plane 1 of a volume
    with an odd plane size)."""
    o = po.oracle()
    for fmt, dims in ((po.UINT8, (255, 257, 70)), (po.UINT16, (256, 256, 70)), (po.UINT32, (256, 128, 130)),
                      (po.FLOAT32, (128, 256, 130))):
        data = po.synth(dims, fmt, 9)
        const = np.full_like(data, 77)
        for arr, lo, hi, nb in ((data, 0.0, 1.0, 256), (data, 0.0, 1.0, 37), (const, 0.0, 1.0, 256),
                                (data, 0.25, 0.75, 200), (data, -1.0, 2.0, 255), (data, 0.0, 1.0, 257), (data, 0.0, 1.0, 1000)):
            vol = po.Vol(arr, fmt, lo, hi)
            for first in ((0, 0, 0), (0, 0, 1)):
                want, wdrop = o.histogram_range(vol, first, dims, nb)
                got, gdrop = gpu.histogram_range(vol, first, dims, nb)
                assert got.tobytes() == want.tobytes() and gdrop == wdrop, (fmt, lo, hi, nb, first)
                assert int(got.sum()) + gdrop == dims[0] * dims[1] * (dims[2] - first[2])


def test_resample_f32_linear_gate(gpu):
    """Float32 under reference-faithful Linear: clean data takes scan -> one-tap rows -> gated (skipped) march; ONE
    Inf, NaN or -0 anywhere in the source flips the gate and the exact 8-tap march recomputes everything.  Bit-exact
    against the oracle either way (the poisoned voxels sit where they are the +1 taps of other voxels: 0 * Inf = NaN
    must spread exactly as in the reference, and -0 + +0 = +0 must lose its sign)."""
    o = po.oracle()
    S, D = (64, 48, 40), (96, 72, 60)          # 2:3 up-sampling, 16-byte aligned rows
    base = po.synth(S, po.FLOAT32, 7)
    base[0, 0, 0] = 0.0                          # a +0 voxel is clean
    poisons = [None,
               ((5, 7, 9), np.float32(np.inf)), ((39, 47, 63), np.float32(np.nan)), ((20, 3, 1), np.float32(-0.0)),
               ((0, 0, 1), np.float32(-np.inf))]
    for dfmt in (po.FLOAT32, po.UINT16):
        for poison in poisons:
            a = base.copy()
            if poison is not None:
                a[poison[0]] = poison[1]
            src = po.Vol(a, po.FLOAT32)
            want, got = po.empty(D, dfmt), po.empty(D, dfmt)
            o.resample(want, src, po.LINEAR)
            gpu.resample(got, src, po.LINEAR)
            assert same_bits(got.a, want.a), (dfmt, poison)
    # a whole plane of NaN
    a = base.copy()
    a[21, :, :] = np.float32(np.nan)
    src = po.Vol(a, po.FLOAT32)
    want, got = po.empty(D, po.FLOAT32), po.empty(D, po.FLOAT32)
    o.resample(want, src, po.LINEAR)
    gpu.resample(got, src, po.LINEAR)
    assert same_bits(got.a, want.a)
    # z-window form (sharded / striped volumes): dst planes [12, 30) read src planes [8, 21); the poison sits ONLY in
    # src plane 20 = the +1 tap plane of the last dst plane, which no dst plane of the window maps to as its lo plane
    # (the row pipeline never stages it: the small separate scan has to see it)
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    lib = gpu.lib
    dz0, dz1 = 12, 30
    s0, s1 = ctypes.c_int32(), ctypes.c_int32()
    capi.check(lib.vktb_resample_window_src_range3(capi.LINEAR, dz0, dz1, capi.I3(*D), capi.I3(*S), ctypes.byref(s0), ctypes.byref(s1)))
    assert int(29 / 60 * 40) == 19 and s0.value <= 8 and s1.value >= 21
    for poison_plane, value in ((20, np.float32(np.inf)), (None, None), (s0.value, np.float32(-0.0))):
        a = base.copy()
        if poison_plane is not None:
            a[poison_plane, 11, 13] = value
        win = np.ascontiguousarray(a[s0.value:s1.value])
        want = po.empty((D[0], D[1], dz1 - dz0), po.FLOAT32)
        o.resample_window(want, po.Vol(win, po.FLOAT32), po.LINEAR, dz0, D, s0.value, S)
        dsrc = DeviceVolume.from_numpy(win, capi.FLOAT32)
        ddst = DeviceVolume((D[0], D[1], dz1 - dz0), capi.FLOAT32)
        capi.check(lib.vktb_resample_window(ctypes.byref(ddst.desc()), ctypes.byref(dsrc.desc()), capi.LINEAR, dz0, capi.I3(*D),
                                            s0.value, capi.I3(*S)), "resample window")
        assert same_bits(ddst.download(), want.a), (poison_plane,)
        dsrc.free(); ddst.free()


def test_pageable_round_trip_through_the_copy_pool(gpu):
    """pageable host memory travels through the pinned staging ring; chunks of 2 MiB and more are cut into parts copied
    by several threads (memory.cu CopyPool): a 70 MB array of odd length comes back byte for byte, several times"""
    from volkit_b200 import capi
    lib = gpu.lib
    n = 70 * 1000 * 1000 + 13
    for seed in (1, 2, 3):
        a = po.synth_numpy(n, po.UINT8, seed)
        p = ctypes.c_void_p()
        capi.check(lib.vktb_alloc(ctypes.byref(p), n))
        capi.check(lib.vktb_memcpy(p, a.ctypes.data, n, capi.H2D))
        a_copy = a.copy()
        a[:] = 0                                            # the source may be overwritten on return
        back = np.empty(n, np.uint8)
        capi.check(lib.vktb_memcpy(back.ctypes.data, p, n, capi.D2H))
        assert back.tobytes() == a_copy.tobytes()
        capi.check(lib.vktb_free(p))


def test_memset_and_pools(gpu):
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    d = DeviceVolume((1000, 3, 1), capi.UINT8)
    pat = (ctypes.c_uint8 * 3)(1, 2, 3)
    capi.check(gpu.lib.vktb_memset_range(d.ptr, pat, 3000, 3))
    assert d.download().ravel().tolist() == [1, 2, 3] * 1000
    # pinned pool round trip
    p = ctypes.c_void_p()
    capi.check(gpu.lib.vktb_host_alloc(ctypes.byref(p), 1 << 20))
    ctypes.memset(p.value, 7, 1 << 20)
    e = DeviceVolume((1 << 20, 1, 1), capi.UINT8)
    capi.check(gpu.lib.vktb_memcpy(e.ptr, p.value, 1 << 20, capi.H2D))
    capi.check(gpu.lib.vktb_host_free(p.value))      # freed right after the call, as migrate() does
    assert (e.download() == 7).all()
