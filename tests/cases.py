"""Shared case catalogue for the golden fixtures.

Every case is (name, kind, params); inputs are regenerated from seeds (src1 = 1,
src2 = 2, dst pre-fill = 3) so only the reference's OUTPUT bytes are stored in tests/golden/golden.npz.
`run_case(lib, case)` runs a case on any implementation exposing the pyoracle._Lib method set (the C
restatement, the compiled reference, or the CUDA product adapter in tests/gpu_adapter.py).
"""
from __future__ import annotations

import numpy as np

from oracle import pyoracle as po

U8, U16, U32, F32 = po.UINT8, po.UINT16, po.UINT32, po.FLOAT32
FMT4 = [U8, U16, U32, F32]


# (source dims, brick size, halo neg, halo pos) for BrickDecompose
BRICK_CASES = [((16, 12, 10), (8, 4, 5), (0, 0, 0), (0, 0, 0)), ((17, 13, 9), (8, 8, 8), (0, 0, 0), (0, 0, 0)),
               ((17, 13, 9), (5, 6, 4), (1, 1, 1), (1, 1, 1)), ((20, 7, 6), (7, 7, 7), (2, 0, 1), (0, 3, 1)),
               ((9, 9, 9), (16, 16, 16), (1, 2, 3), (3, 2, 1)), ((12, 1, 5), (4, 1, 2), (0, 0, 0), (1, 0, 0))]


# ScanRange: (dst dims, src dims, first, last, dstOffset, in place)
SCAN_CASES = [((12, 9, 7), (12, 9, 7), (0, 0, 0), (12, 9, 7), (0, 0, 0), False),
              ((12, 9, 7), (12, 9, 7), (0, 0, 0), (12, 9, 7), (0, 0, 0), True),
              ((20, 14, 11), (12, 9, 7), (1, 2, 1), (11, 8, 6), (5, 3, 4), False),
              ((17, 9, 8), (17, 9, 8), (3, 1, 2), (16, 8, 7), (0, 0, 0), True),
              ((17, 9, 8), (17, 9, 8), (15, 1, 2), (2, 8, 7), (0, 0, 0), True),      # x right to left
              ((17, 9, 8), (17, 9, 8), (3, 7, 6), (16, 0, 1), (0, 0, 0), True),      # y, z downwards
              ((16, 8, 8), (16, 8, 8), (15, 7, 7), (-1, -1, -1), (0, 0, 0), True),   # every axis reversed, whole volume
              ((40, 3, 2), (40, 3, 2), (0, 0, 0), (40, 1, 1), (0, 1, 1), False),     # a single row
              ((5, 6, 40), (5, 6, 40), (2, 0, 0), (3, 6, 40), (0, 0, 0), True)]      # one voxel wide


# Scale / Rotate: (source dims, dest dims, ("scale", factor, center) | ("rotate", axis, angle, center))
AFFINE_CASES = [((21, 13, 11), (21, 13, 11), ("scale", (1.5, 0.75, 2.0), (10.0, 6.0, 5.0))),
                ((21, 13, 11), (30, 9, 14), ("scale", (0.5, 1.0, 3.0), (0.0, 0.0, 0.0))),
                ((16, 16, 16), (16, 16, 16), ("scale", (1.0, 1.0, 1.0), (3.5, 2.25, 8.0))),     # identity: a plain copy
                ((21, 13, 11), (21, 13, 11), ("rotate", (1.0, 2.0, 0.5), 0.7, (10.0, 6.0, 5.0))),
                ((21, 13, 11), (30, 9, 14), ("rotate", (0.0, 0.0, 1.0), 1.5707964, (10.0, 6.0, 5.0))),
                ((24, 24, 8), (24, 24, 8), ("rotate", (0.0, 0.0, -3.0), 3.1415927, (11.5, 11.5, 4.0)))]
# ApplyFilter: filter dims (x, y, z), each 1..3
FILTER_DIMS = [(3, 3, 3), (1, 2, 3), (2, 2, 2), (3, 1, 1), (1, 1, 1)]


def filter_weights(fdims, seed=0):
    """(fz, fy, fx) float32 weights; seed 0 = the 3x3x3 box blur the reference's users pass, else random"""
    n = fdims[0] * fdims[1] * fdims[2]
    if seed == 0:
        return np.full(fdims[::-1], np.float32(1.0) / np.float32(n), np.float32)
    return np.random.default_rng(seed).standard_normal(fdims[::-1]).astype(np.float32)


# ComputeAggregatesRange: (dims, first, last); whole volume, a range (numElems stays the whole volume's), a few voxels
# (prod stays inside the binary32 range), one row, one voxel
AGG_CASES = [((33, 9, 8), (0, 0, 0), (33, 9, 8)), ((33, 9, 8), (1, 2, 1), (30, 9, 7)), ((33, 9, 8), (5, 3, 2), (8, 5, 4)),
             ((40, 3, 2), (0, 1, 1), (40, 2, 2)), ((7, 5, 3), (6, 4, 2), (7, 5, 3))]


def agg_exact_words(a) -> np.ndarray:
    """the order-independent fields of vkt::Aggregates: min, max (bit patterns), argmin, argmax"""
    w = np.frombuffer(a.words(), np.uint32)
    return np.concatenate([w[0:2], w[7:13]]).copy()


def agg_accumulated_words(a) -> np.ndarray:
    """mean, stddev, var, sum, prod: sequential binary32 accumulations in the reference"""
    return np.frombuffer(a.words(), np.uint32)[2:7].copy()


def unmapped(o, vol) -> np.ndarray:
    """every voxel's getValue() as float32 (Z, Y, X): the oracle's equal-dims Resample into Float32 with the (0, 1)
    mapping is map(unmap(v)) = unmap(v)"""
    dims = vol.a.shape[::-1]
    f = po.empty(dims, po.FLOAT32)
    o.resample(f, vol, po.NEAREST)
    return f.a


def check_aggregates(o, got, ref, vol, first, last, what="", parallel=True):
    """The contract of vktb_aggregates_range (include/vktb/backend.h): min/max/argmin/argmax identical to the
    reference; sum, mean, var, stddev, prod within the reference's own a-priori rounding bound of the real-number
    value, and the parallel result itself within 1e-6 of that value (float64 numpy)."""
    assert agg_exact_words(got).tobytes() == agg_exact_words(ref).tobytes(), (what, "min/max/argmin/argmax")
    v = unmapped(o, vol)[first[2]:last[2], first[1]:last[1], first[0]:last[0]].astype(np.float64).ravel()
    n, eps = v.size, 2.0 ** -24
    num_elems = float(vol.a.size)
    if not np.isfinite(v).all():
        return
    sabs, ssum = np.abs(v).sum(), v.sum()
    tiny = 1e-30
    # sum: sequential binary32 summation error <= (n - 1) eps sum|v| (+ the final rounding of either side)
    bsum = n * eps * sabs + 2 * eps * abs(ssum) + tiny
    assert abs(float(got.sum) - float(ref.sum)) <= bsum + 1e-6 * sabs, (what, "sum", got.sum, ref.sum, bsum)
    if parallel:
        assert abs(float(got.sum) - ssum) <= 1e-6 * sabs + tiny, (what, "sum vs float64", got.sum, ssum)
    bmean = bsum / num_elems + eps * abs(float(ref.mean))
    assert abs(float(got.mean) - float(ref.mean)) <= bmean + 1e-6 * sabs / num_elems, (what, "mean", got.mean, ref.mean)
    # var: terms (v - mean)^2 carry 3 roundings each, the running sum n more; the two means differ by <= dm
    m = float(ref.mean)
    dev = v - m
    sq = (dev * dev).sum()
    dm = abs(float(got.mean) - m)
    bvar = ((n + 4) * eps * sq + 2 * abs(dev.sum()) * dm + n * dm * dm) / num_elems + 2 * eps * float(ref.var) + tiny
    assert abs(float(got.var) - float(ref.var)) <= bvar + 1e-6 * sq / num_elems, (what, "var", got.var, ref.var, bvar)
    if parallel:
        assert abs(float(got.var) - ((v - float(got.mean)) ** 2).sum() / num_elems) <= 2e-6 * sq / num_elems + tiny, \
            (what, "var vs float64")
    assert np.float32(got.stddev) == np.sqrt(np.float32(got.var)), (what, "stddev = sqrtf(var)")
    # prod: relative bound while the running product stays a normal binary32 number; 0 / inf classes otherwise
    rp, gp = float(ref.prod), float(got.prod)
    if np.isnan(rp):
        return                                   # inf * 0 inside the reference's running product
    if np.isinf(rp) or np.isinf(gp):
        with np.errstate(all="ignore"):
            mags = np.cumsum(np.log2(np.maximum(np.abs(v), 1e-300)))
        if np.isinf(rp) and np.isinf(gp):
            assert rp == gp, (what, "prod sign", gp, rp)
        else:                                    # the running product left the binary32 range on the way
            assert mags.max() > 126, (what, "prod", gp, rp)
        return
    assert abs(gp - rp) <= 2 * n * eps * abs(rp) + 1.2e-38, (what, "prod", gp, rp)


def mk(dims, fmt, seed, lo=0.0, hi=1.0):
    return po.Vol(po.synth(dims, fmt, seed), fmt, lo, hi)


def cases():
    out = []
    # --- Fill: SURVEY.md F6 table + ranges
    for fmt in FMT4:
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
            for value in (0.0, 0.1, 0.5, 0.999, 1.0, 1.5, -0.25, 2999.0):
                out.append((f"fill/{fmt}/{lo}/{hi}/{value}", "fill",
                            dict(dims=(13, 7, 5), fmt=fmt, lo=lo, hi=hi, first=(1, 0, 2), last=(12, 7, 4), value=value)))
    # --- Copy: same-format (bytewise) and converting, dstOffset, clamped src
    for dfmt in FMT4:
        for sfmt in FMT4:
            for smap in ((0.0, 1.0), (-1000.0, 3000.0)):
                for first, last, off in (((0, 0, 0), (12, 9, 7), (0, 0, 0)), ((1, 2, 1), (11, 8, 6), (8, 8, 8)),
                                         ((-3, -2, -1), (15, 11, 9), (1, 2, 3))):
                    out.append((f"copy/{dfmt}/{sfmt}/{smap[0]}/{first}/{off}", "copy",
                                dict(ddims=(32, 24, 20), sdims=(12, 9, 7), dfmt=dfmt, sfmt=sfmt, smap=smap,
                                     first=first, last=last, off=off)))
    # --- Arithmetic: 10 ops x formats x mappings, whole + range/offset
    combos = [(U16, U16, U16), (U8, U8, U8), (U32, U32, U32), (F32, F32, F32), (U8, U16, F32), (F32, U8, U32),
              (U16, F32, U8)]
    for op in range(10):
        for dfmt, f1, f2 in combos:
            for dmap in ((0.0, 1.0), (-1000.0, 3000.0)):
                out.append((f"arith/{op}/{dfmt}/{f1}/{f2}/{dmap[0]}", "arith",
                            dict(op=op, ddims=(40, 12, 10), sdims=(33, 9, 7), dfmt=dfmt, f1=f1, f2=f2, dmap=dmap,
                                 s2map=((0.0, 1.0), (-1000.0, 3000.0))[op % 2],
                                 first=(3, 1, 2), last=(32, 9, 6), off=(8, 3, 4))))
        out.append((f"arith/{op}/whole", "arith",
                    dict(op=op, ddims=(37, 11, 6), sdims=(37, 11, 6), dfmt=U16, f1=U16, f2=U16, dmap=(0.0, 1.0),
                         s2map=(0.0, 1.0), first=(0, 0, 0), last=(37, 11, 6), off=(0, 0, 0))))
    # --- Transform
    for fmt in FMT4:
        for op in range(4):
            out.append((f"transform1/{fmt}/{op}", "transform1",
                        dict(dims=(21, 19, 18), fmt=fmt, lo=-1000.0, hi=3000.0, first=(1, 2, 3), last=(20, 18, 17), op=op)))
        for op in range(2):
            out.append((f"transform2/{fmt}/{op}", "transform2",
                        dict(dims=(21, 9, 8), fmt=fmt, first=(1, 0, 1), last=(21, 8, 7), op=op)))
    # --- Resample: nearest + linear(reference) over down/up/equal/anisotropic dims
    for sfmt in FMT4:
        for dfmt in FMT4:
            for ddims in ((8, 6, 5), (24, 18, 15), (16, 12, 10), (9, 17, 7), (33, 5, 11)):
                for filt in (po.NEAREST, po.LINEAR):
                    out.append((f"resample/{sfmt}/{dfmt}/{ddims}/{filt}", "resample",
                                dict(sdims=(16, 12, 10), ddims=ddims, sfmt=sfmt, dfmt=dfmt, filt=filt,
                                     smap=(0.0, 1.0), dmap=(-1.0, 2.0))))
    # --- Flip: all three axes (Y mirrors y and z, Z mirrors nothing -- as the reference is
    # written), whole volume / range that ignores `first` / dstOffset, and in place
    for fmt in (U8, U16, F32):
        for axis in range(3):
            out.append((f"flip/{fmt}/{axis}/whole", "flip",
                        dict(sdims=(22, 9, 8), ddims=(22, 9, 8), fmt=fmt, first=(0, 0, 0), last=(22, 9, 8), off=(0, 0, 0),
                             axis=axis, inplace=False)))
            out.append((f"flip/{fmt}/{axis}/range", "flip",
                        dict(sdims=(22, 9, 8), ddims=(25, 12, 10), fmt=fmt, first=(3, 1, 2), last=(20, 8, 7), off=(2, 3, 1),
                             axis=axis, inplace=False)))
            out.append((f"flip/{fmt}/{axis}/inplace", "flip",
                        dict(sdims=(16, 9, 8), ddims=(16, 9, 8), fmt=fmt, first=(0, 0, 0), last=(16, 9, 8), off=(0, 0, 0),
                             axis=axis, inplace=True)))
    # --- Histogram: bins as uint64
    for fmt, lo, hi in ((U8, 0.0, 1.0), (U16, 0.0, 1.0), (U16, -1000.0, 3000.0), (U32, 0.0, 1.0), (F32, 0.0, 1.0)):
        for nb in (7, 256, 5000):
            out.append((f"hist/{fmt}/{lo}/{nb}", "hist",
                        dict(dims=(31, 14, 9), fmt=fmt, lo=lo, hi=hi, first=(1, 0, 2), last=(30, 13, 9), bins=nb)))
    # --- BrickDecompose and Scan (rank 4)
    for fmt in (U8, U16, F32):
        for i, (sdims, brick, neg, pos) in enumerate(BRICK_CASES):
            out.append((f"bricks/{fmt}/{i}", "bricks", dict(sdims=sdims, fmt=fmt, brick=brick, neg=neg, pos=pos)))
        for i, (ddims, sdims, first, last, off, inplace) in enumerate(SCAN_CASES):
            out.append((f"scan/{fmt}/{i}", "scan",
                        dict(ddims=ddims, sdims=sdims, fmt=fmt, first=first, last=last, off=off, inplace=inplace)))
    # --- Scale / Rotate / ApplyFilter
    for sfmt, dfmt in ((U8, U8), (U16, U16), (U32, U32), (F32, F32), (U16, F32), (F32, U8)):
        for i in range(len(AFFINE_CASES)):
            out.append((f"affine/{sfmt}/{dfmt}/{i}", "affine", dict(sfmt=sfmt, dfmt=dfmt, i=i)))
        for i, fd in enumerate(FILTER_DIMS):
            for inplace in ((False, True) if sfmt == dfmt else (False,)):
                out.append((f"filter/{sfmt}/{dfmt}/{i}/{int(inplace)}", "filter",
                            dict(sfmt=sfmt, dfmt=dfmt, fdims=fd, seed=i, inplace=inplace)))
    # --- Aggregates: [exact fields, accumulated fields] of the reference
    for fmt in FMT4:
        for lo, hi in ((0.0, 1.0), (-1000.0, 3000.0)):
            for i, (dims, first, last) in enumerate(AGG_CASES):
                out.append((f"agg/{fmt}/{lo}/{i}", "agg", dict(dims=dims, fmt=fmt, lo=lo, hi=hi, first=first, last=last)))
    # --- BASELINE config 1 (CLI pipeline on a 64^3 uint8 volume -> 32^3 uint16)
    out.append(("cli/config1", "cli1", dict()))
    return out


def run_case(lib, case) -> list[np.ndarray]:
    """Returns the list of output arrays of the case."""
    _, kind, p = case
    if kind == "fill":
        v = mk(p["dims"], p["fmt"], 3, p["lo"], p["hi"])
        lib.fill_range(v, p["first"], p["last"], p["value"])
        return [v.a]
    if kind == "copy":
        d = mk(p["ddims"], p["dfmt"], 3)
        s = mk(p["sdims"], p["sfmt"], 1, *p["smap"])
        lib.copy_range(d, s, p["first"], p["last"], p["off"])
        return [d.a]
    if kind == "arith":
        d = mk(p["ddims"], p["dfmt"], 3, *p["dmap"])
        s1 = mk(p["sdims"], p["f1"], 1)
        s2 = mk(p["sdims"], p["f2"], 2, *p["s2map"])
        lib.arith_range(p["op"], d, s1, s2, p["first"], p["last"], p["off"])
        return [d.a]
    if kind == "transform1":
        v = mk(p["dims"], p["fmt"], 1, p["lo"], p["hi"])
        lib.transform_range1(v, p["first"], p["last"], p["op"])
        return [v.a]
    if kind == "transform2":
        a, b = mk(p["dims"], p["fmt"], 1), mk(p["dims"], p["fmt"], 2)
        lib.transform_range2(a, b, p["first"], p["last"], p["op"])
        return [a.a, b.a]
    if kind == "resample":
        s = mk(p["sdims"], p["sfmt"], 1, *p["smap"])
        d = mk(p["ddims"], p["dfmt"], 3, *p["dmap"])
        lib.resample(d, s, p["filt"])
        return [d.a]
    if kind == "flip":
        s = mk(p["sdims"], p["fmt"], 1)
        if p["inplace"]:
            lib.flip_range(s, s, p["first"], p["last"], p["off"], p["axis"])
            return [s.a]
        d = mk(p["ddims"], p["fmt"], 3)
        lib.flip_range(d, s, p["first"], p["last"], p["off"], p["axis"])
        return [d.a]
    if kind == "hist":
        v = mk(p["dims"], p["fmt"], 1, p["lo"], p["hi"])
        bins, dropped = lib.histogram_range(v, p["first"], p["last"], p["bins"])
        assert dropped == 0, "case must stay inside the reference's defined behaviour"
        return [bins]
    if kind == "affine":
        sdims, ddims, op = AFFINE_CASES[p["i"]]
        src = mk(sdims, p["sfmt"], 2, -1000.0, 3000.0)
        d = mk(ddims, p["dfmt"], 5, -1.0, 2.0)
        if op[0] == "scale":
            lib.scale(d, src, op[1], op[2])
        else:
            lib.rotate(d, src, op[1], op[2], op[3])
        return [d.a]
    if kind == "filter":
        src = mk((26, 13, 11), p["sfmt"], 2, -1.0, 2.0)
        w = filter_weights(p["fdims"], p["seed"])
        if p["inplace"]:
            lib.apply_filter(src, src, w)
            return [src.a]
        d = mk((29, 13, 12), p["dfmt"], 5, 0.0, 1.0)
        lib.apply_filter(d, src, w)
        return [d.a]
    if kind == "agg":
        v = mk(p["dims"], p["fmt"], 4, p["lo"], p["hi"])
        a = lib.aggregates_range(v, p["first"], p["last"])
        return [agg_exact_words(a), agg_accumulated_words(a)]
    if kind == "bricks":
        s = mk(p["sdims"], p["fmt"], 3, -1000.0, 3000.0)
        return list(lib.brick_decompose(s, p["brick"], p["neg"], p["pos"]))
    if kind == "scan":
        s = mk(p["sdims"], p["fmt"], 2)
        if p["inplace"]:
            lib.scan_range(s, s, p["first"], p["last"], p["off"])
            return [s.a]
        d = mk(p["ddims"], p["fmt"], 5)
        lib.scan_range(d, s, p["first"], p["last"], p["off"])
        return [d.a]
    if kind == "cli1":
        # declare-sv 64^3 uint8 | fill 0.1 | fill-range 1 [0,8)^3 | resample --dims 32^3 uint16
        # (reference src/cli/main.cpp:402-442, 521-574, 701-740; the CLI resamples with Nearest, :734)
        v = po.empty((64, 64, 64), U8)
        lib.fill_range(v, (0, 0, 0), (64, 64, 64), 0.1)
        lib.fill_range(v, (0, 0, 0), (8, 8, 8), 1.0)
        d = po.empty((32, 32, 32), U16)
        lib.resample(d, v, po.NEAREST)
        return [v.a, d.a]
    raise ValueError(kind)
