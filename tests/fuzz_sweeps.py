"""Random case generators shared by tests/test_oracle_fuzz.py (restatement against the compiled reference, CPU) and
tests/test_gpu_parity.py::test_random_sweep_against_the_oracle (CUDA path through the C-ABI against the restatement).
Every sweep takes two backends with the pyoracle interface, the number of cases, and `reference`: True when the second
backend is the compiled reference, whose undefined corner (out-of-bounds histogram bins) must be avoided -- the other one,
the tap one voxel past the source buffer, is made deterministic by oracle/ref_shim.cpp (it zeroes the malloc slack); the restatement and the CUDA path define those corners and are compared there too."""
import numpy as np

import cases
from oracle import pyoracle as po

FMTS = [po.UINT8, po.UINT16, po.UINT32, po.FLOAT32]
MAPS = [(0.0, 1.0), (-1000.0, 3000.0), (0.25, 0.75), (1.0, 0.0), (-3.5, -1.25), (0.0, 65535.0)]


def rdims(rng, lo=1, hi=14):
    return tuple(int(v) for v in rng.randint(lo, hi + 1, size=3))


def rrange(rng, dims):
    """a non-empty box [first, last) inside dims"""
    first = tuple(int(rng.randint(0, d)) for d in dims)
    last = tuple(int(rng.randint(f + 1, d + 1)) for f, d in zip(first, dims))
    return first, last


def rmap(rng):
    return MAPS[rng.randint(len(MAPS))]


def same(a, b):
    """equal bytes; Float32: equal bytes or NaN in the same places (payloads may differ between builds)"""
    if a.tobytes() == b.tobytes():
        return True
    if a.dtype != np.float32:
        return False
    an, bn = np.isnan(a), np.isnan(b)
    return bool((an == bn).all()) and a[~an].tobytes() == b[~bn].tobytes()


def sweep_fill_transform(o, r, n, reference):
    rng = np.random.RandomState(101)
    values = [0.0, 1.0, -1.0, 0.5, 0.1, 3.0e9, -2.5e4, 65535.0, 1e-30, float("inf"), float("-inf"), float("nan")]
    for i in range(n):
        fmt = FMTS[rng.randint(4)]
        dims = rdims(rng)
        lo, hi = rmap(rng)
        first, last = rrange(rng, dims)
        a, b = cases.mk(dims, fmt, i, lo, hi), cases.mk(dims, fmt, i, lo, hi)
        if rng.randint(2):
            v = values[rng.randint(len(values))] if rng.randint(3) else float(rng.uniform(lo - 1, hi + 1))
            o.fill_range(a, first, last, v)
            r.fill_range(b, first, last, v)
            what = ("fill", fmt, dims, lo, hi, first, last, v)
        else:
            op = int(rng.randint(4))
            if r.transform_range1(b, first, last, op) is False:
                continue                                    # an op this backend has no host function for
            o.transform_range1(a, first, last, op)
            what = ("transform", fmt, dims, first, last, op)
        assert same(a.a, b.a), what


def sweep_copy_range(o, r, n, reference):
    rng = np.random.RandomState(202)
    for i in range(n):
        sfmt, dfmt = FMTS[rng.randint(4)], FMTS[rng.randint(4)]
        sdims, ddims = rdims(rng), rdims(rng)
        # the copied box fits into dst at `off`; its source coordinates may start before / end after the source volume
        ext = tuple(int(rng.randint(1, d + 1)) for d in ddims)
        off = tuple(int(rng.randint(0, d - e + 1)) for d, e in zip(ddims, ext))
        first = tuple(int(rng.randint(-3, s + 2)) for s in sdims)
        last = tuple(f + e for f, e in zip(first, ext))
        src = cases.mk(sdims, sfmt, i, *rmap(rng))
        dlo, dhi = rmap(rng)
        a, b = cases.mk(ddims, dfmt, 1000 + i, dlo, dhi), cases.mk(ddims, dfmt, 1000 + i, dlo, dhi)
        o.copy_range(a, src, first, last, off)
        r.copy_range(b, src, first, last, off)
        assert same(a.a, b.a), (sfmt, dfmt, sdims, ddims, first, last, off)


def sweep_arithmetic(o, r, n, reference):
    rng = np.random.RandomState(303)
    for i in range(n):
        op = int(rng.randint(10))
        f1, f2, fd = (FMTS[rng.randint(4)] for _ in range(3))
        sdims = rdims(rng, 2, 12)
        first, last = rrange(rng, sdims)
        ext = tuple(l - f for f, l in zip(first, last))
        ddims = tuple(int(e + rng.randint(0, 4)) for e in ext)
        ddims = tuple(max(d, f + e) if rng.randint(2) else d for d, f, e in zip(ddims, first, ext))
        # dst receives the box at first + off: keep it inside
        off = tuple(int(rng.randint(-f, d - (f + e) + 1)) if d >= e and d - (f + e) >= -f else -f
                    for d, f, e in zip(ddims, first, ext))
        if any(f + o_ < 0 or f + o_ + e > d for f, o_, e, d in zip(first, off, ext, ddims)):
            continue
        s1, s2 = cases.mk(sdims, f1, i, *rmap(rng)), cases.mk(sdims, f2, 500 + i, *rmap(rng))
        dlo, dhi = rmap(rng)
        a, b = cases.mk(ddims, fd, 900 + i, dlo, dhi), cases.mk(ddims, fd, 900 + i, dlo, dhi)
        o.arith_range(op, a, s1, s2, first, last, off)
        r.arith_range(op, b, s1, s2, first, last, off)
        assert same(a.a, b.a), (op, f1, f2, fd, sdims, ddims, first, last, off)


def sweep_resample(o, r, n, reference):
    rng = np.random.RandomState(404)
    for i in range(n):
        sfmt, dfmt = FMTS[rng.randint(4)], FMTS[rng.randint(4)]
        sdims, ddims = rdims(rng, 1, 11), rdims(rng, 1, 13)
        if rng.randint(6) == 0:
            ddims = sdims                                  # the equal-dims branch (a converting copy)
        filt = (po.NEAREST, po.LINEAR)[rng.randint(2)]
        src = cases.mk(sdims, sfmt, i, *rmap(rng))
        if sfmt == po.FLOAT32 and rng.randint(3) == 0:     # special values travel through the 8-tap chain
            flat = src.a.reshape(-1)
            flat[rng.randint(flat.size)] = [np.inf, -np.inf, np.nan, -0.0][rng.randint(4)]
        dlo, dhi = rmap(rng)
        a, b = cases.mk(ddims, dfmt, 700 + i, dlo, dhi), cases.mk(ddims, dfmt, 700 + i, dlo, dhi)
        o.resample(a, src, filt)
        r.resample(b, src, filt)
        assert same(a.a, b.a), (sfmt, dfmt, sdims, ddims, filt)


def sweep_resample_aligned(o, r, n, reference):
    """rows of whole 16-byte vectors, 16 .. 96 voxels per axis: the shapes that reach the bulk-copy row pipeline and the TMA
    z-march instead of the direct-gather kernel"""
    rng = np.random.RandomState(707)
    for i in range(n):
        sfmt, dfmt = FMTS[rng.randint(4)], FMTS[rng.randint(4)]
        sdims, ddims = tuple(8 * v for v in rdims(rng, 2, 10)), tuple(8 * v for v in rdims(rng, 2, 12))
        if rng.randint(6) == 0:
            ddims = sdims                                  # the equal-dims branch (a converting copy)
        filt = (po.NEAREST, po.LINEAR)[rng.randint(2)]
        src = cases.mk(sdims, sfmt, i, *rmap(rng))
        if sfmt == po.FLOAT32 and rng.randint(3) == 0:     # special values travel through the 8-tap chain
            flat = src.a.reshape(-1)
            flat[rng.randint(flat.size)] = [np.inf, -np.inf, np.nan, -0.0][rng.randint(4)]
        dlo, dhi = rmap(rng)
        a, b = cases.mk(ddims, dfmt, 700 + i, dlo, dhi), cases.mk(ddims, dfmt, 700 + i, dlo, dhi)
        o.resample(a, src, filt)
        r.resample(b, src, filt)
        assert same(a.a, b.a), (sfmt, dfmt, sdims, ddims, filt)


def sweep_flip_scan_histogram_bricks(o, r, n, reference):
    rng = np.random.RandomState(505)
    for i in range(n):
        fmt = FMTS[rng.randint(4)]
        dims = rdims(rng, 1, 12)
        lo, hi = rmap(rng)
        v = cases.mk(dims, fmt, i, lo, hi)
        kind = rng.randint(4)
        if kind == 0:                                       # in-place flip of a sub-range
            first, last = rrange(rng, dims)
            axis = int(rng.randint(3))
            a, b = v.copy(), v.copy()
            o.flip_range(a, a, first, last, (0, 0, 0), axis)
            r.flip_range(b, b, first, last, (0, 0, 0), axis)
            assert same(a.a, b.a), ("flip", fmt, dims, first, last, axis)
        elif kind == 1:                                     # in-place scan, any direction per axis
            first, last = rrange(rng, dims)
            f, l = list(first), list(last)
            for ax in range(3):
                if rng.randint(3) == 0:
                    f[ax], l[ax] = last[ax] - 1, first[ax] - 1      # right to left
            a, b = v.copy(), v.copy()
            o.scan_range(a, a, tuple(f), tuple(l))
            r.scan_range(b, b, tuple(f), tuple(l))
            assert same(a.a, b.a), ("scan", fmt, dims, f, l)
        elif kind == 2:
            if hi == lo:
                continue
            first, last = rrange(rng, dims)
            nb = int((1, 2, 7, 64, 256, 1000)[rng.randint(6)])
            if reference and (lo > hi or (lo, hi) != (0.0, 1.0) and fmt == po.FLOAT32):
                continue                                    # bins outside [0, nb): the reference writes out of bounds
            if not hasattr(r, "histogram_range"):
                continue
            ha, da = o.histogram_range(v, first, last, nb)
            hb, db = r.histogram_range(v, first, last, nb)
            if not reference:
                assert da == db and ha.tobytes() == hb.tobytes(), ("hist", fmt, dims, lo, hi, first, last, nb, da, db)
            elif da == 0 and db == 0:
                assert ha.tobytes() == hb.tobytes(), ("hist", fmt, dims, lo, hi, first, last, nb)
        else:
            brick = tuple(int(rng.randint(1, d + 2)) for d in dims)
            neg = tuple(int(rng.randint(0, 3)) for _ in range(3))
            pos = tuple(int(rng.randint(0, 3)) for _ in range(3))
            if not hasattr(r, "brick_decompose"):
                continue
            wa = o.brick_decompose(v, brick, neg, pos)
            wb = r.brick_decompose(v, brick, neg, pos)
            assert len(wa) == len(wb) and all(same(x, y) for x, y in zip(wa, wb)), ("bricks", fmt, dims, brick, neg, pos)


def sweep_scale_rotate_filter(o, r, n, reference):
    """random scale factors / rotation axes and angles / centres between volumes of different dims and formats, and random
    filter weights of every size 1..3 per axis: the libm calls (sinf / cosf in the quaternion) and every float operation
    of the coordinate transform run in both builds"""
    rng = np.random.RandomState(606)
    for i in range(n):
        sfmt = FMTS[rng.randint(4)]
        dfmt = sfmt if rng.randint(2) else FMTS[rng.randint(4)]
        sdims, ddims = rdims(rng, 2, 12), rdims(rng, 2, 12)
        src = cases.mk(sdims, sfmt, i, *rmap(rng))
        dlo, dhi = rmap(rng)
        a, b = cases.mk(ddims, dfmt, 300 + i, dlo, dhi), cases.mk(ddims, dfmt, 300 + i, dlo, dhi)
        center = tuple(float(rng.uniform(-1, d)) for d in sdims)
        kind = rng.randint(3)
        if kind == 0:
            factor = tuple(float(rng.choice([0.5, 0.75, 1.0, 1.5, 2.0, 3.0, float(rng.uniform(0.3, 3))])) for _ in range(3))
            o.scale(a, src, factor, center)
            r.scale(b, src, factor, center)
            what = ("scale", factor)
        elif kind == 1:
            axis = tuple(float(v) for v in rng.uniform(-2, 2, size=3))
            if rng.randint(3) == 0:
                axis = [(1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0)][rng.randint(3)]
            angle = float(rng.uniform(-3.2, 3.2))
            o.rotate(a, src, axis, angle, center)
            r.rotate(b, src, axis, angle, center)
            what = ("rotate", axis, angle)
        else:
            fd = tuple(int(rng.randint(1, 4)) for _ in range(3))
            w = rng.standard_normal(fd[::-1]).astype(np.float32)
            if not hasattr(r, "apply_filter"):
                continue
            a, b = cases.mk(sdims, dfmt, 300 + i, dlo, dhi), cases.mk(sdims, dfmt, 300 + i, dlo, dhi)
            o.apply_filter(a, src, w)
            r.apply_filter(b, src, w)
            what = ("filter", fd)
        assert same(a.a, b.a), (what, sfmt, dfmt, sdims, ddims, center)


def sweep_aggregates(o, r, n, reference):
    """ComputeAggregatesRange word for word (min, max, mean, stddev, var, sum, prod, argmin, argmax): only for backends that
    accumulate sequentially in binary32 like the reference (the restatement, the compiled reference, the host mirror's CPU
    policy); the CUDA path has its own tolerance contract (tests/cases.py check_aggregates)"""
    rng = np.random.RandomState(808)
    for i in range(n):
        fmt = FMTS[rng.randint(4)]
        dims = rdims(rng, 1, 10)
        lo, hi = rmap(rng)
        v = cases.mk(dims, fmt, i, lo, hi)
        if fmt == po.FLOAT32 and rng.randint(3) == 0:
            flat = v.a.reshape(-1)
            flat[rng.randint(flat.size)] = [np.inf, -np.inf, np.nan, -0.0, 0.0][rng.randint(5)]
        first, last = rrange(rng, dims)
        if rng.randint(3) == 0:
            first, last = (0, 0, 0), dims
        wa, wb = o.aggregates_range(v, first, last), r.aggregates_range(v, first, last)
        a, b = np.frombuffer(wa.words(), np.float32, 7), np.frombuffer(wb.words(), np.float32, 7)
        assert same(a, b) and wa.words()[28:] == wb.words()[28:], (fmt, dims, lo, hi, first, last, a, b)


def sweep_transform2(o, r, n, reference):
    """TransformRange with a binary op (both volumes are modified), same and mixed formats"""
    rng = np.random.RandomState(909)
    for i in range(n):
        f1, f2 = FMTS[rng.randint(4)], FMTS[rng.randint(4)]
        dims = rdims(rng, 1, 12)
        first, last = rrange(rng, dims)
        op = int(rng.randint(2))
        m1, m2 = rmap(rng), rmap(rng)
        a1, a2 = cases.mk(dims, f1, i, *m1), cases.mk(dims, f2, 400 + i, *m2)
        b1, b2 = a1.copy(), a2.copy()
        if not hasattr(r, "transform_range2"):
            continue
        o.transform_range2(a1, a2, first, last, op)
        r.transform_range2(b1, b2, first, last, op)
        assert same(a1.a, b1.a) and same(a2.a, b2.a), (f1, f2, dims, first, last, op)
