"""Pins the C restatement (oracle/vkt_oracle.c) against the UNMODIFIED reference serial path
(oracle/_ref/libvkt_ref.so, built from /root/reference by oracle/Makefile): byte equality on
identical seeded inputs for every hot-path function, all four in-scope formats (+Int16)."""
import itertools

import numpy as np
import pytest

import cases

from oracle import pyoracle as po

pytestmark = pytest.mark.needs_ref

FMTS = [po.UINT8, po.UINT16, po.UINT32, po.FLOAT32, po.INT16]
MAPPINGS = [(0.0, 1.0), (-1000.0, 3000.0)]


def mk(dims, fmt, seed, lo=0.0, hi=1.0):
    return po.Vol(po.synth(dims, fmt, seed), fmt, lo, hi)


def same(a: po.Vol, b: po.Vol):
    return a.a.tobytes() == b.a.tobytes()


def test_scalar_codec_sweep():
    o, r = po.oracle(), po.reference()
    for lo, hi in MAPPINGS + [(0.25, 0.75)]:
        for fmt, n in ((po.UINT8, 256), (po.UINT16, 65536), (po.INT16, 65536)):
            for i in range(0, n, 1 if n == 256 else 17):
                raw = int(i).to_bytes(2, "little")[: 1 if n == 256 else 2]
                vo, vr = o.unmap_voxel(raw, fmt, lo, hi), r.unmap_voxel(raw, fmt, lo, hi)
                assert np.float32(vo).tobytes() == np.float32(vr).tobytes()
                assert o.map_voxel(vo, fmt, lo, hi) == r.map_voxel(vr, fmt, lo, hi)


def test_map_edge_values():
    o, r = po.oracle(), po.reference()
    edge = [0.0, -0.0, 0.1, 0.5, 0.999, 1.0, 1.5, -0.25, 2.0, -1.0, 1e-30, -1e-30, 1e10, -1e10, 3e38, -3e38,
            float("inf"), float("-inf"), float("nan"), 0.99999994, 32767.5 / 65536, 255.0 / 255.999]
    for fmt in FMTS:
        for lo, hi in MAPPINGS + [(1.0, 1.0), (1.0, 0.0)]:
            for v in edge:
                assert o.map_voxel(v, fmt, lo, hi) == r.map_voxel(v, fmt, lo, hi), (fmt, lo, hi, v)


@pytest.mark.parametrize("fmt", FMTS)
def test_fill(fmt):
    o, r = po.oracle(), po.reference()
    for lo, hi in MAPPINGS:
        for value in (0.1, 1.0, 1.5, -0.25, 2500.0):
            a, b = mk((9, 7, 5), fmt, 3, lo, hi), mk((9, 7, 5), fmt, 3, lo, hi)
            o.fill_range(a, (1, 2, 0), (8, 6, 4), value)
            r.fill_range(b, (1, 2, 0), (8, 6, 4), value)
            assert same(a, b)


@pytest.mark.parametrize("dfmt,sfmt", list(itertools.product(FMTS[:4], FMTS[:4])))
def test_copy(dfmt, sfmt):
    o, r = po.oracle(), po.reference()
    for (dlo, dhi), (slo, shi) in (((0, 1), (0, 1)), ((0, 1), (-1000, 3000))):
        src = mk((8, 6, 5), sfmt, 1, slo, shi)
        for first, last, off in (((0, 0, 0), (8, 6, 5), (0, 0, 0)), ((1, 2, 1), (7, 5, 4), (3, 2, 1)),
                                 ((-2, -1, -3), (10, 8, 7), (1, 1, 1))):
            a, b = mk((16, 14, 12), dfmt, 3, dlo, dhi), mk((16, 14, 12), dfmt, 3, dlo, dhi)
            o.copy_range(a, src, first, last, off)
            r.copy_range(b, src, first, last, off)
            assert same(a, b)


@pytest.mark.parametrize("op", range(10))
def test_arith(op):
    o, r = po.oracle(), po.reference()
    for dfmt, f1, f2 in ((po.UINT16,) * 3, (po.UINT8,) * 3, (po.UINT32,) * 3, (po.FLOAT32,) * 3,
                         (po.UINT8, po.UINT16, po.FLOAT32), (po.FLOAT32, po.UINT8, po.UINT32), (po.INT16,) * 3):
        for (dlo, dhi) in MAPPINGS:
            s1, s2 = mk((11, 6, 5), f1, 1), mk((11, 6, 5), f2, 2, *MAPPINGS[op % 2])
            a, b = mk((16, 9, 8), dfmt, 3, dlo, dhi), mk((16, 9, 8), dfmt, 3, dlo, dhi)
            o.arith_range(op, a, s1, s2, (1, 0, 2), (10, 6, 5), (4, 3, 2))
            r.arith_range(op, b, s1, s2, (1, 0, 2), (10, 6, 5), (4, 3, 2))
            assert same(a, b), (op, dfmt, f1, f2)


def test_arith_exhaustive_u8_pairs():
    """all 65536 UInt8 x UInt8 pairs for each of the ten ops."""
    o, r = po.oracle(), po.reference()
    g = np.arange(256, dtype=np.uint8)
    A = np.repeat(g, 256).reshape(1, 256, 256).copy()
    B = np.tile(g, 256).reshape(1, 256, 256).copy()
    for op in range(10):
        for lo, hi in MAPPINGS:
            a, b = po.empty((256, 256, 1), po.UINT8, lo, hi), po.empty((256, 256, 1), po.UINT8, lo, hi)
            o.arith_range(op, a, po.Vol(A, po.UINT8), po.Vol(B, po.UINT8, lo, hi), (0, 0, 0), (256, 256, 1))
            r.arith_range(op, b, po.Vol(A, po.UINT8), po.Vol(B, po.UINT8, lo, hi), (0, 0, 0), (256, 256, 1))
            assert same(a, b), op


@pytest.mark.parametrize("fmt", FMTS[:4])
def test_transform(fmt):
    o, r = po.oracle(), po.reference()
    for op in range(4):
        a, b = mk((20, 19, 18), fmt, 1, -1000, 3000), mk((20, 19, 18), fmt, 1, -1000, 3000)
        o.transform_range1(a, (1, 2, 3), (19, 18, 17), op)
        r.transform_range1(b, (1, 2, 3), (19, 18, 17), op)
        assert same(a, b), op
    for op in range(2):
        a1, a2 = mk((9, 8, 7), fmt, 1), mk((9, 8, 7), fmt, 2)
        b1, b2 = a1.copy(), a2.copy()
        o.transform_range2(a1, a2, (1, 0, 1), (9, 7, 6), op)
        r.transform_range2(b1, b2, (1, 0, 1), (9, 7, 6), op)
        assert same(a1, b1) and same(a2, b2)


@pytest.mark.parametrize("sfmt,dfmt", list(itertools.product(FMTS[:4], FMTS[:4])))
def test_resample(sfmt, dfmt):
    o, r = po.oracle(), po.reference()
    src = mk((12, 10, 8), sfmt, 1, 0.0, 1.0)
    for ddims in ((6, 5, 4), (18, 15, 12), (12, 10, 8), (7, 13, 5), (24, 3, 9)):
        for filt in (po.NEAREST, po.LINEAR):
            a, b = po.empty(ddims, dfmt, -1.0, 2.0), po.empty(ddims, dfmt, -1.0, 2.0)
            o.resample(a, src, filt)
            r.resample(b, src, filt)
            assert same(a, b), (ddims, filt)


@pytest.mark.parametrize("fmt", FMTS[:4])
def test_flip(fmt):
    """the restatement keeps the reference's quirks: `first` ignored, z mirrored on Axis::Y, nothing mirrored on Axis::Z"""
    o, r = po.oracle(), po.reference()
    for dims, ddims, first, last, off in (((9, 7, 6), (9, 7, 6), (0, 0, 0), (9, 7, 6), (0, 0, 0)),
                                          ((12, 10, 8), (14, 13, 10), (2, 0, 1), (9, 7, 6), (1, 2, 1)),
                                          ((8, 8, 8), (8, 8, 8), (1, 1, 1), (8, 8, 8), (0, 0, 0))):
        for axis in range(3):
            src = mk(dims, fmt, 1)
            a, b = mk(ddims, fmt, 3), mk(ddims, fmt, 3)
            o.flip_range(a, src, first, last, off, axis)
            r.flip_range(b, src, first, last, off, axis)
            assert a.a.tobytes() == b.a.tobytes(), (dims, axis)
            if off == (0, 0, 0) and dims == ddims:
                a, b = src.copy(), src.copy()
                o.flip_range(a, a, first, last, off, axis)
                r.flip_range(b, b, first, last, off, axis)
                assert a.a.tobytes() == b.a.tobytes(), ("in place", dims, axis)
    # the documented quirks, literally
    v = po.Vol(np.arange(4 * 3 * 2, dtype=np.uint8).reshape(2, 3, 4), po.UINT8)
    d = po.empty((4, 3, 2), po.UINT8)
    r.flip_range(d, v, (0, 0, 0), (4, 3, 2), (0, 0, 0), 2)               # Axis::Z: lower half copied, nothing mirrored
    assert d.a[0].tobytes() == v.a[0].tobytes() and not d.a[1].any()
    d = po.empty((4, 3, 2), po.UINT8)
    r.flip_range(d, v, (0, 0, 0), (4, 3, 2), (0, 0, 0), 1)               # Axis::Y: y and z mirrored together
    assert d.a[1, 2, 0] == v.a[0, 0, 0] and d.a[0, 0, 0] == v.a[1, 2, 0]


@pytest.mark.parametrize("fmt", FMTS[:4])
def test_histogram(fmt):
    o, r = po.oracle(), po.reference()
    for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
        if fmt == po.FLOAT32 and lo != 0.0:
            continue                                   # Float32 voxels are not mapped on read: they would leave the bins
        v = mk((25, 12, 10), fmt, 2, lo, hi)
        for nb in (1, 16, 256, 4099):
            ha, dropped = o.histogram_range(v, (1, 2, 3), (24, 11, 9), nb)
            hb, _ = r.histogram_range(v, (1, 2, 3), (24, 11, 9), nb)
            assert dropped == 0 and ha.tobytes() == hb.tobytes(), (lo, nb)
            assert int(ha.sum()) == 23 * 9 * 6


@pytest.mark.parametrize("fmt", FMTS)
def test_scale_rotate_filter(fmt):
    """Scale / Rotate / ApplyFilter: the restatements reproduce the reference build byte for byte; an identity scale is
    a plain (converting) copy and a z rotation by a quarter turn about the centre of a square plane permutes voxels."""
    o, r = po.oracle(), po.reference()
    for sdims, ddims, op in cases.AFFINE_CASES:
        for dfmt in (fmt, po.FLOAT32):
            src = mk(sdims, fmt, 2, -1000.0, 3000.0)
            a, b = mk(ddims, dfmt, 5, -1.0, 2.0), mk(ddims, dfmt, 5, -1.0, 2.0)
            for lib, d in ((o, a), (r, b)):
                if op[0] == "scale":
                    lib.scale(d, src, op[1], op[2])
                else:
                    lib.rotate(d, src, op[1], op[2], op[3])
            assert a.a.tobytes() == b.a.tobytes(), (sdims, ddims, op)
            if op[0] == "scale" and op[1] == (1.0, 1.0, 1.0):
                want = mk(ddims, dfmt, 5, -1.0, 2.0)
                o.copy_range(want, src, (0, 0, 0), sdims, (0, 0, 0))
                assert a.a.tobytes() == want.a.tobytes()
    rng = np.random.default_rng(5)
    for fd in cases.FILTER_DIMS:
        for seed in (0, 7):
            w = cases.filter_weights(fd, seed)
            src = mk((26, 13, 11), fmt, 2, -1.0, 2.0)
            a, b = mk((29, 13, 12), fmt, 5), mk((29, 13, 12), fmt, 5)
            o.apply_filter(a, src, w); r.apply_filter(b, src, w)
            assert a.a.tobytes() == b.a.tobytes(), (fd, seed)
            c, d = src.copy(), src.copy()
            o.apply_filter(c, c, w); r.apply_filter(d, d, w)
            assert c.a.tobytes() == d.a.tobytes(), (fd, seed, "in place")
            # the in-place call filters a copy: same voxels as the out-of-place call into a same-shaped volume
            e = po.empty((26, 13, 11), fmt, -1.0, 2.0)
            o.apply_filter(e, src, w)
            assert e.a.tobytes() == c.a.tobytes()
    del rng


def test_stream_io(tmp_path):
    """OutputStream::writeRange / InputStream::readRange / RawFile: the byte-buffer restatement against the reference
    run on real files, and against a numpy statement (lines of last.x - first.x voxels from the START of each row)."""
    o, r = po.oracle(), po.reference()
    for fmt in (po.UINT8, po.UINT16, po.FLOAT32):
        vol = mk((21, 9, 8), fmt, 2)
        for first, last in (((0, 0, 0), (21, 9, 8)), ((3, 1, 2), (17, 8, 7)), ((5, 0, 0), (6, 9, 1)), ((2, 2, 2), (2, 5, 5))):
            path = str(tmp_path / f"range_{fmt}_{first[0]}_{last[0]}.raw")
            err = r.write_range(vol, first, last, path)
            want = o.rows_gather(vol, first, last)
            assert open(path, "rb").read() == want, (fmt, first, last)
            nx = last[0] - first[0]
            assert want == b"".join(vol.a[z, y, :nx].tobytes() for z in range(first[2], last[2]) for y in range(first[1], last[1]))
            # RawFile::write returns fwrite's item count, so the reference reports WriteError (3) unless one item == the total
            items = (last[1] - first[1]) * (last[2] - first[2]) if nx > 0 else 0     # fwrite of 0 bytes returns 0
            assert err == (0 if items == len(want) else 3)
            a, b = mk((21, 9, 8), fmt, 5), mk((21, 9, 8), fmt, 5)
            r.read_range(a, first, last, path)
            o.rows_scatter(b, first, last, want)
            assert a.a.tobytes() == b.a.tobytes(), (fmt, first, last)
        path = str(tmp_path / f"whole_{fmt}.raw")
        assert r.write(vol, path) == 3 and open(path, "rb").read() == vol.a.tobytes()    # bytes arrive, the code says WriteError
        back = mk((21, 9, 8), fmt, 5)
        assert r.read(back, path) == 2 and back.a.tobytes() == vol.a.tobytes()           # ... and ReadError
    for name in ("vol_48x40x36_uint16.raw", "a_b_512x256x128_int8_x.raw", "foo.raw", "x_10x20x30.raw", "s_uint32_7x8x9_uint64.raw",
                 "trailing_4x4x4_uint16_"):
        full = tmp_path / name
        full.write_bytes(b"")
        assert o.rawfile_parse(str(full)) == r.rawfile_parse(str(full)), name
    assert o.rawfile_parse("vol_48x40x36_uint16.raw") == ((48, 40, 36), po.UINT16)


@pytest.mark.parametrize("fmt", FMTS)
def test_aggregates(fmt):
    """ComputeAggregatesRange: the restatement reproduces all 13 words of the reference's vkt::Aggregates; min / max /
    argmin / argmax also equal an independent numpy statement (first occurrence in x-fastest order)."""
    o, r = po.oracle(), po.reference()
    for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
        for dims, first, last in cases.AGG_CASES + [((21, 6, 5), (2, 2, 2), (2, 6, 5))]:      # + an empty range
            v = mk(dims, fmt, 4, lo, hi)
            a, b = o.aggregates_range(v, first, last), r.aggregates_range(v, first, last)
            assert a.words() == b.words(), (dims, first, last, lo)
            box = cases.unmapped(o, v)[first[2]:last[2], first[1]:last[1], first[0]:last[0]]
            if box.size == 0:
                assert a.min == np.finfo(np.float32).max and a.prod == 1.0 and a.sum == 0.0
                continue
            nx, ny = last[0] - first[0], last[1] - first[1]
            for val, arg, pos in ((a.min, a.argmin, int(np.argmin(box))), (a.max, a.argmax, int(np.argmax(box)))):
                assert np.float32(val) == box.ravel()[pos]
                assert list(arg) == [first[0] + pos % nx, first[1] + (pos // nx) % ny, first[2] + pos // (nx * ny)]
            cases.check_aggregates(o, a, b, v, first, last, parallel=False)   # the bound holds trivially for ref == ref


@pytest.mark.parametrize("fmt", FMTS[:4])
def test_scan(fmt):
    """ScanRange: oracle (reference loops restated) == reference == numpy cumsum along x, y, z modulo 2^(8 bpv)."""
    o, r = po.oracle(), po.reference()
    for ddims, sdims, first, last, off, in_place in cases.SCAN_CASES:
        src = mk(sdims, fmt, 2, -1000.0, 3000.0)
        if in_place:
            a, b = src.copy(), src.copy()
            o.scan_range(a, a, first, last, off)
            r.scan_range(b, b, first, last, off)
        else:
            a, b = mk(ddims, fmt, 5, -1000.0, 3000.0), mk(ddims, fmt, 5, -1000.0, 3000.0)
            o.scan_range(a, src, first, last, off)
            r.scan_range(b, src, first, last, off)
        assert a.a.tobytes() == b.a.tobytes(), (ddims, first, last, off, in_place)
        # independent statement: three cumulative sums of the raw integers, walked in the range's directions
        base = src if in_place else mk(ddims, fmt, 5, -1000.0, 3000.0)
        want = base.a.copy()
        if not in_place:
            o.copy_range(po.Vol(want, fmt, -1000.0, 3000.0), src, first, last, off)
        sl = []
        for ax in (2, 1, 0):                                   # numpy axes (z, y, x)
            f, l = first[ax] + off[ax], last[ax] + off[ax]
            sl.append(slice(f, l) if f < l else slice(f, l if l >= 0 else None, -1))
        box = want[sl[0], sl[1], sl[2]].view({1: np.uint8, 2: np.uint16, 4: np.uint32}[want.dtype.itemsize])
        acc = box.astype(np.uint64)
        for ax in range(3):
            acc = np.cumsum(acc, axis=ax, dtype=np.uint64)
        box[...] = acc.astype(box.dtype)
        assert want.tobytes() == a.a.tobytes(), (ddims, first, last, off, in_place)


@pytest.mark.parametrize("fmt", FMTS[:4])
def test_brick_decompose(fmt):
    """BrickDecomposeResize's brick grid and BrickDecompose's contents (halo voxels clamped at the volume border)."""
    o, r = po.oracle(), po.reference()
    for sdims, brick, neg, pos in cases.BRICK_CASES:
        assert o.brick_layout(sdims, brick, neg, pos) == r.brick_layout(sdims, brick, neg, pos), (sdims, brick)
        v = mk(sdims, fmt, 3, -1000.0, 3000.0)
        a, b = o.brick_decompose(v, brick, neg, pos), r.brick_decompose(v, brick, neg, pos)
        assert len(a) == len(b)
        for i, (x, y) in enumerate(zip(a, b)):
            assert x.shape == y.shape and x.tobytes() == y.tobytes(), (sdims, brick, neg, pos, i)
    # without halos the bricks tile the volume exactly
    v = mk((16, 12, 10), fmt, 3)
    bricks = o.brick_decompose(v, (8, 4, 5))
    assert np.array_equal(bricks[0], v.a[:5, :4, :8]) and np.array_equal(bricks[-1], v.a[5:, 8:, 8:])


def test_resample_linear_equals_nearest_in_reference():
    """the reference's Linear mode is numerically nearest-floor."""
    r = po.reference()
    src = mk((16, 16, 16), po.UINT16, 1)
    a, b = po.empty((24, 24, 24), po.UINT16), po.empty((24, 24, 24), po.UINT16)
    r.resample(a, src, po.NEAREST)
    r.resample(b, src, po.LINEAR)
    assert same(a, b)


def test_resample_window_matches_whole():
    o = po.oracle()
    src = mk((10, 9, 16), po.UINT16, 1)
    for ddims in ((15, 7, 24), (5, 9, 8)):
        for filt in (po.NEAREST, po.LINEAR, po.TRILINEAR):
            whole = po.empty(ddims, po.UINT16)
            o.resample(whole, src, filt)
            for dz0, dz1 in ((0, ddims[2] // 2), (ddims[2] // 2, ddims[2])):
                zs = [int(np.float32(np.float32(z) / np.float32(ddims[2])) * np.float32(16)) for z in (dz0, dz1 - 1)]
                sz0, sz1 = zs[0], min(zs[1] + 2, 16)
                win = po.Vol(src.a[sz0:sz1].copy(), po.UINT16)
                part = po.empty((ddims[0], ddims[1], dz1 - dz0), po.UINT16)
                o.resample_window(part, win, filt, dz0, ddims, sz0, (10, 9, 16))
                assert part.a.tobytes() == whole.a[dz0:dz1].tobytes()
