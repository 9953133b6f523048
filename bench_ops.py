"""Per-op table for bench.py (N=1): every in-scope op at the BASELINE sizes that fit one B200, each with GVoxels/s,
algorithmic GB/s (SURVEY.md 8(d) bytes) and fraction of the measured HBM peak.  Inputs are resident in HBM.

After its timing every op is run once more on a known destination state and compared with the oracle AT THAT SIZE on
sampled z-planes (bench_parity.py: first, last, eight random planes); the verdicts land in `parity` (a
bench_parity.Parity) and the per-op entry gets "parity": true/false."""
import ctypes
import math

import numpy as np


class _Raw:
    """device memory as a torch tensor (checker side only: value counts for the histogram check)"""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def run_ops(torch, lib, capi, DeviceVolume, S, peak, time_steps, parity=None, steps=5, warmup=3):
    from bench import torch_synth
    from bench_parity import sample_planes
    from oracle import pyoracle as po   # checker

    out = {}
    o = po.oracle()

    def synth_into(vol, seed):
        bpv = capi.BYTES_PER_VOXEL[vol.fmt]
        n = vol.nbytes // bpv
        for f0 in range(0, n, 1 << 27):
            m = min(1 << 27, n - f0)
            t = torch_synth(torch, m, bpv, seed, f0, float32=(vol.fmt == capi.FLOAT32))
            capi.check(lib.vktb_memcpy(vol.ptr + f0 * bpv, t.data_ptr(), m * bpv, capi.D2D))
            torch.cuda.synchronize()
            del t
        torch.cuda.empty_cache()

    def record(name, voxels, alg_bytes, fn, note=None, check=None, prepare=None):
        """times fn; then (prepare(); fn() once more;) check() -> bool compares the result with the oracle"""
        ms = time_steps(torch, fn, steps, warmup)
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gvoxels_s": voxels / (ms * 1e-3) / 1e9, "alg_gb_s": gbs, "frac_of_measured_peak": gbs / peak,
                     "frac_of_8TBs": gbs / 8000.0, "alg_bytes": alg_bytes}
        if note:
            out[name]["note"] = note
        if check is not None and parity is not None:
            if prepare is not None:
                prepare()
                fn()
            out[name]["parity"] = bool(check(name))

    z3 = capi.I3(0, 0, 0)
    n = S ** 3
    PF = 0.25                                                      # pre-fill value of dst before a *Range check

    def vol1(a, fmt, lo=0.0, hi=1.0):
        return po.Vol(np.ascontiguousarray(a), fmt, lo, hi)

    def fill_plane(dims, fmt, lo=0.0, hi=1.0):
        p = po.empty((dims[0], dims[1], 1), fmt, lo, hi)
        o.fill_range(p, (0, 0, 0), (dims[0], dims[1], 1), PF)
        return p.a

    # ---- generic plane checks ---------------------------------------------------------------------------------
    def check_arith(name, dvol, op, fmt, dims, seeds, first, last, off, maps=((0.0, 1.0),) * 3, prefilled=True):
        """dst plane z <- op(s1, s2) plane z - off.z inside [first, last), the rest keeps the pre-fill"""
        pre = fill_plane(dims, fmt, *maps[0]) if prefilled else None

        def expected(z):
            d = vol1(pre.copy() if prefilled else np.zeros((1, dims[1], dims[0]), po.DTYPES[fmt]), fmt, *maps[0])
            pz = z - off[2]
            if first[2] <= pz < last[2]:
                a = vol1(parity.regen(seeds[0], fmt, dims, pz), fmt, *maps[1])
                b = vol1(parity.regen(seeds[1], fmt, dims, pz), fmt, *maps[2])
                o.arith_range(op, d, a, b, (first[0], first[1], 0), (last[0], last[1], 1), (off[0], off[1], 0))
            return d.a
        return parity.check(name, sample_planes(dims[2]), lambda z: parity.fetch(dvol.ptr, dims, fmt, z), expected)

    def check_copy(name, dvol, fmt, dims, seed, first, last, off, prefilled=True):
        """dst plane z <- src plane clamp(z - off.z + first.z) inside the box, the rest keeps the pre-fill"""
        pre = fill_plane(dims, fmt) if prefilled else None

        def expected(z):
            d = vol1(pre.copy() if prefilled else np.zeros((1, dims[1], dims[0]), po.DTYPES[fmt]), fmt)
            pz = z - off[2] + first[2]
            if first[2] <= pz < last[2]:
                a = vol1(parity.regen(seed, fmt, dims, min(max(pz, 0), dims[2] - 1)), fmt)
                o.copy_range(d, a, (first[0], first[1], 0), (last[0], last[1], 1), (off[0], off[1], 0))
            return d.a
        return parity.check(name, sample_planes(dims[2]), lambda z: parity.fetch(dvol.ptr, dims, fmt, z), expected)

    def check_transform(name, dvol, fmt, dims, seed, op):
        def expected(z):
            v = vol1(parity.regen(seed, fmt, dims, z), fmt)
            o.transform_range1_window(v, (0, 0, 0), (dims[0], dims[1], 1), op, z)
            return v.a
        return parity.check(name, sample_planes(dims[2]), lambda z: parity.fetch(dvol.ptr, dims, fmt, z), expected)

    def check_resample(name, dvol, dfmt, D, sfmt, Sd, seed, filt):
        def expected(z):
            s0 = int(z * Sd[2] / D[2])
            w0, w1 = max(s0 - 1, 0), min(s0 + 3, Sd[2])            # generous: the taps are planes s0 and s0 + 1
            src = vol1(parity.regen(seed, sfmt, Sd, w0, w1 - w0), sfmt)
            d = po.empty((D[0], D[1], 1), dfmt)
            o.resample_window(d, src, filt, z, D, w0, Sd)
            return d.a
        return parity.check(name, sample_planes(D[2]), lambda z: parity.fetch(dvol.ptr, D, dfmt, z), expected)

    # ---- UInt16 S^3: Fill, Copy, CopyRange(dstOffset), SafeDiff, Sum (general mapping), Transform
    U16, dims = capi.UINT16, (S, S, S)
    a, b, d = DeviceVolume(dims, U16), DeviceVolume(dims, U16), DeviceVolume(dims, U16)
    synth_into(a, 1)
    synth_into(b, 2)
    l3 = capi.I3(S, S, S)
    da, db, dd = a.desc(), b.desc(), d.desc()
    fill = lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, PF))  # noqa: E731
    pre16 = fill_plane(dims, po.UINT16) if parity is not None else None
    record(f"Fill {S}^3 UInt16", n, 2 * n, fill,
           check=lambda nm: parity.check(nm, sample_planes(S), lambda z: parity.fetch(d.ptr, dims, po.UINT16, z), lambda z: pre16))
    record(f"Copy {S}^3 UInt16", n, 4 * n, lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3)),
           check=lambda nm: check_copy(nm, d, po.UINT16, dims, 1, (0, 0, 0), dims, (0, 0, 0), prefilled=False))
    m = S - 8
    mmm = (m, m, m)
    record(f"CopyRange {m}^3 UInt16 dstOffset (8,8,8)", m ** 3, 4 * m ** 3,
           lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, capi.I3(m, m, m), capi.I3(8, 8, 8))),
           prepare=fill, check=lambda nm: check_copy(nm, d, po.UINT16, dims, 1, (0, 0, 0), mmm, (8, 8, 8)))
    record(f"SafeDiff {S}^3 UInt16", n, 6 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFEDIFF, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3, l3, z3)),
           check=lambda nm: check_arith(nm, d, po.SAFEDIFF, po.UINT16, dims, (1, 2), (0, 0, 0), dims, (0, 0, 0), prefilled=False))
    record(f"SumRange {m}^3 UInt16 dstOffset (8,8,8)", m ** 3, 6 * m ** 3,
           lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3,
                                                   capi.I3(m, m, m), capi.I3(8, 8, 8))),
           prepare=fill, check=lambda nm: check_arith(nm, d, po.SUM, po.UINT16, dims, (1, 2), (0, 0, 0), mmm, (8, 8, 8)))
    # sources at another 16-byte phase than the destination (funnel-shifted loads)
    record(f"CopyRange {m}^3 UInt16 first (1,1,1) dstOffset (0,0,0)", m ** 3, 4 * m ** 3,
           lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), capi.I3(1, 1, 1),
                                                  capi.I3(m + 1, m + 1, m + 1), z3)),
           prepare=fill, check=lambda nm: check_copy(nm, d, po.UINT16, dims, 1, (1, 1, 1), (m + 1,) * 3, (0, 0, 0)))
    record(f"SumRange {m}^3 UInt16 dstOffset (1,2,3)", m ** 3, 6 * m ** 3,
           lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3,
                                                   capi.I3(m, m, m), capi.I3(1, 2, 3))),
           prepare=fill, check=lambda nm: check_arith(nm, d, po.SUM, po.UINT16, dims, (1, 2), (0, 0, 0), mmm, (1, 2, 3)))
    gm = capi.vol(a.ptr, (S, S, S), capi.UINT16, -1000.0, 3000.0)
    gd = capi.vol(d.ptr, (S, S, S), capi.UINT16, -1000.0, 3000.0)
    gmaps = ((-1000.0, 3000.0), (-1000.0, 3000.0), (0.0, 1.0))
    record(f"SafeProd {S}^3 UInt16 mapping (-1000,3000)", n, 6 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFEPROD, ctypes.byref(gd), ctypes.byref(gm), ctypes.byref(db), z3, l3, z3)),
           check=lambda nm: check_arith(nm, d, po.SAFEPROD, po.UINT16, dims, (1, 2), (0, 0, 0), dims, (0, 0, 0), gmaps, prefilled=False))
    copy_a = lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3))  # noqa: E731
    record(f"Transform xor_coord {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(dd), z3, l3, 0)),
           prepare=copy_a, check=lambda nm: check_transform(nm, d, po.UINT16, dims, 1, 0))

    # a user's OWN op (tests/user_ops/ops.cu: the reference example's MakeCheckered<3>), registered from its cubin
    import os
    udir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "user_ops", "build")
    if os.path.exists(os.path.join(udir, "user_ops.cubin")) and os.path.exists(os.path.join(udir, "libuser_ops_host.so")):
        img = open(os.path.join(udir, "user_ops.cubin"), "rb").read()
        hnd = ctypes.c_int(-1)
        capi.check(lib.vktb_register_unary_op_image(img, len(img), b"checkered3", ctypes.byref(hnd)), "register user op")
        host_ops = ctypes.CDLL(os.path.join(udir, "libuser_ops_host.so"))
        host_fn = ctypes.cast(host_ops.checkered3, ctypes.c_void_p).value

        def check_user(name):
            def expected(z):
                v = vol1(parity.regen(1, po.UINT16, dims, z), po.UINT16)
                o.transform_fn1(v, (0, 0, 0), (S, S, 1), host_fn, z)       # the host twin of the same source
                return v.a
            return parity.check(name, sample_planes(S), lambda z: parity.fetch(d.ptr, dims, po.UINT16, z), expected)

        record(f"Transform user op checkered3 (cubin) {S}^3 UInt16", n, 4 * n,
               lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(dd), z3, l3, hnd.value)),
               "the user's op inlined into the kernel its own image carries (include/vktb/transform_op.cuh)",
               prepare=copy_a, check=check_user)

    # ---- SURVEY 8(f) rows ---------------------------------------------------------------------------------------
    def check_flip(name, axis):
        """axis X mirrors x inside a plane; axis Y mirrors y AND z (the reference as written): dst planes z and
        S-1-z come from the oracle run on the two-plane volume {z, S-1-z}"""
        def expected(z):
            if axis == 0:
                src = vol1(parity.regen(1, po.UINT16, dims, z), po.UINT16)
                dst = vol1(np.zeros_like(src.a), po.UINT16)
                o.flip_range(dst, src, (0, 0, 0), (S, S, 1), (0, 0, 0), 0)
                return dst.a
            zz = S - 1 - z
            lo, hi = min(z, zz), max(z, zz)
            src = vol1(np.concatenate([parity.regen(1, po.UINT16, dims, lo), parity.regen(1, po.UINT16, dims, hi)]), po.UINT16)
            dst = vol1(np.zeros_like(src.a), po.UINT16)
            o.flip_range(dst, src, (0, 0, 0), (S, S, 2), (0, 0, 0), 1)
            return dst.a[0:1] if z == lo else dst.a[1:2]
        return parity.check(name, sample_planes(S), lambda z: parity.fetch(d.ptr, dims, po.UINT16, z), expected)

    for axis, aname in ((0, "X"), (1, "Y (mirrors y and z, as the reference)")):
        record(f"Flip {aname} {S}^3 UInt16", n, 4 * (n if S % 2 == 0 else n - S * S),
               lambda: capi.check(lib.vktb_flip_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3, axis)),
               check=(lambda nm, ax=axis: check_flip(nm, ax)) if S % 2 == 0 else None)

    def check_scan(name):
        """plane z of the 3-D prefix sum minus plane z-1 (mod 2^16) is the 2-D prefix sum of src plane z: the oracle
        scans the regenerated plane; plane 0 is compared directly"""
        def got(z):
            cur = parity.fetch(d.ptr, dims, po.UINT16, z)
            return cur if z == 0 else (cur - parity.fetch(d.ptr, dims, po.UINT16, z - 1)).astype(np.uint16)

        def expected(z):
            v = vol1(parity.regen(1, po.UINT16, dims, z), po.UINT16)
            o.scan_range(v, v, (0, 0, 0), (S, S, 1), (0, 0, 0))
            return v.a
        return parity.check(name, sample_planes(S), got, expected)

    record(f"Scan {S}^3 UInt16 in place", n, 4 * n,
           lambda: capi.check(lib.vktb_scan_range(ctypes.byref(dd), ctypes.byref(dd), z3, l3, z3)),
           "alg_bytes = one read + one write per voxel (the ideal single pass); the kernel makes two passes",
           prepare=copy_a, check=check_scan)
    # BrickDecompose: 64^3 bricks, halo 1 on every side, all bricks in ONE launch
    bs, halo = 64, 1
    nb = (S + bs - 1) // bs
    edge = lambda i: halo + min(bs, S - i * bs) + halo  # noqa: E731
    bricks, firsts, lasts, total = [], [], [], 0
    for z in range(nb):
        for y in range(nb):
            for x in range(nb):
                bd = (edge(x), edge(y), edge(z))
                bricks.append((total, bd))
                firsts += [x * bs - halo, y * bs - halo, z * bs - halo]
                lasts += [min((x + 1) * bs, S) + halo, min((y + 1) * bs, S) + halo, min((z + 1) * bs, S) + halo]
                total += bd[0] * bd[1] * bd[2]
    pool = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(pool), total * 2))
    descs = (capi.VktbVol * len(bricks))(*[capi.vol(pool.value + 2 * off, bd, capi.UINT16, 0.0, 1.0) for off, bd in bricks])
    i32 = lambda v: (ctypes.c_int32 * len(v))(*v)  # noqa: E731
    bf, bl, bo = i32(firsts), i32(lasts), i32([0] * len(firsts))

    def check_bricks(name):
        """sampled bricks (first, last, eight random), three planes each: brick plane k <- clamped src plane"""
        rng = np.random.RandomState(7)
        picks = sorted({0, len(bricks) - 1} | {int(v) for v in rng.randint(0, len(bricks), size=8)})
        items = [(i, k) for i in picks for k in sorted({0, bricks[i][1][2] // 2, bricks[i][1][2] - 1})]

        def got(it):
            i, k = it
            off, bd = bricks[i]
            return parity.fetch(pool.value + 2 * off, bd, po.UINT16, k)

        def expected(it):
            i, k = it
            _, bd = bricks[i]
            f, l = firsts[3 * i:3 * i + 3], lasts[3 * i:3 * i + 3]
            src = vol1(parity.regen(1, po.UINT16, dims, min(max(f[2] + k, 0), S - 1)), po.UINT16)
            dst = po.empty((bd[0], bd[1], 1), po.UINT16)
            o.copy_range(dst, src, (f[0], f[1], 0), (l[0], l[1], 1), (0, 0, 0))
            return dst.a
        ok = parity.check(name, items, got, expected)
        parity.results[name]["bricks"] = len(picks)
        return ok

    record(f"BrickDecompose {S}^3 UInt16 -> {len(bricks)} bricks of {bs}^3 + halo {halo}", total, 4 * total,
           lambda: capi.check(lib.vktb_copy_range_batch(len(bricks), descs, ctypes.byref(da), bf, bl, bo)),
           "one launch for all bricks; voxels = brick voxels incl. halos, alg_bytes = 2 bpv per brick voxel",
           check=check_bricks)
    capi.check(lib.vktb_free(pool))
    bins = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(bins), 256 * 8))

    def check_hist(name):
        """WHOLE volume: the count of every UInt16 value (torch.bincount over the device volume, checker side) folded
        through the oracle's value -> bin table must equal the 256 bins"""
        table = torch.from_numpy(parity.bin_table_u16(256)).cuda()
        counts = torch.zeros(65536, dtype=torch.int64, device="cuda")
        vals = torch.as_tensor(_Raw(a.ptr, n, "<i2"), device="cuda")
        for f0 in range(0, n, 1 << 27):
            counts += torch.bincount(vals[f0:f0 + (1 << 27)].to(torch.int32) & 0xFFFF, minlength=65536)
        want = torch.zeros(256, dtype=torch.int64, device="cuda").index_add_(0, table.clamp(min=0), counts * (table >= 0))
        got = np.empty(256, np.uint64)
        capi.check(lib.vktb_memcpy(got.ctypes.data, bins.value, 256 * 8, capi.D2H))
        torch.cuda.empty_cache()
        return parity.record(name, got.astype(np.int64).tolist() == want.cpu().tolist() and int(got.sum()) == n,
                             "all voxels: value counts x the oracle's value->bin table == bins")

    record(f"Histogram 256 bins {S}^3 UInt16", n, 2 * n,
           lambda: capi.check(lib.vktb_histogram_range(ctypes.byref(da), z3, l3, bins, 256, None)), check=check_hist)
    capi.check(lib.vktb_free(bins))

    # Scale / Rotate / ApplyFilter
    F9, F3 = ctypes.c_float * 9, ctypes.c_float * 3
    ctr = (S / 2, S / 2, S / 2)
    half = F3(*ctr)

    def check_affine(name, m):
        """matrices that keep z apart (Scale: diagonal; Rotate about z): dst plane z taps the src planes
        (int)q.z and (int)q.z + 1, q.z evaluated like the oracle does"""
        def expected(z):
            qz = np.float32(np.float32(m[8]) * np.float32(np.float32(z) - np.float32(ctr[2]))) + np.float32(ctr[2])
            s0 = int(qz) if np.isfinite(qz) else 0
            w0, w1 = min(max(s0 - 1, 0), S - 1), max(min(s0 + 3, S), 1)
            src = vol1(parity.regen(1, po.UINT16, dims, w0, w1 - w0), po.UINT16)
            dst = vol1(pre16.copy(), po.UINT16)
            o.affine_window(dst, z, src, w0, dims, (0, 0, 0), dims, m, ctr)
            return dst.a
        return parity.check(name, sample_planes(S), lambda z: parity.fetch(d.ptr, dims, po.UINT16, z), expected)

    ms_ = [1 / 1.5, 0, 0, 0, 1 / 1.5, 0, 0, 0, 1 / 1.5]
    record(f"Scale x1.5 about the centre {S}^3 UInt16", n, int(2 * n + 2 * n / 3.375),
           lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, F9(*ms_), half)),
           "every dst voxel is written; the (1/1.5)^3 of the source it maps to is read once",
           prepare=fill, check=lambda nm: check_affine(nm, list(F9(*ms_))))
    cs, sn = math.cos(0.5), math.sin(0.5)
    mr_ = [cs, -sn, 0, sn, cs, 0, 0, 0, 1]
    record(f"Rotate 0.5 rad about z through the centre {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, F9(*mr_), half)),
           "alg_bytes = one read + one write per voxel (upper bound: voxels that rotate out of the volume are skipped)",
           prepare=fill, check=lambda nm: check_affine(nm, list(F9(*mr_))))
    box = (ctypes.c_float * 27)(*([1.0 / 27.0] * 27))
    wbox = np.full((3, 3, 3), np.float32(1.0) / np.float32(27.0), np.float32)

    def check_filter(name):
        def expected(z):
            w0, w1 = max(z - 1, 0), min(z + 2, S)
            src = vol1(parity.regen(1, po.UINT16, dims, w0, w1 - w0), po.UINT16)
            dst = po.empty((S, S, 1), po.UINT16)
            o.apply_filter_window(dst, z, src, w0, dims, wbox)
            return dst.a
        return parity.check(name, sample_planes(S, n_random=4), lambda z: parity.fetch(d.ptr, dims, po.UINT16, z), expected)

    record(f"ApplyFilter 3x3x3 {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_apply_filter(ctypes.byref(dd), ctypes.byref(da), box, capi.I3(3, 3, 3), 2)),
           "27 FMUL + 27 FADD per voxel in the reference's order: CUDA-core bound, see DESIGN.md", check=check_filter)
    # Aggregates: one read pass, result on the host (the call is synchronous)
    agg = capi.VktbAggregates()

    def check_agg(name, lo=0.0, hi=1.0):
        """extrema: a uniform UInt16 stream reaches 0 and 65535 -- the format's own extremes -- inside the first
        planes, so the oracle's min / max / argmin / argmax of those planes are the whole volume's; sum against a
        binary64 sum of the device volume within the 1e-6 contract of backend.h"""
        k = max(1, min(S, (1 << 22) // (S * S)))
        head = vol1(parity.regen(1, po.UINT16, dims, 0, k), po.UINT16, lo, hi)
        ra = o.aggregates_range(head, (0, 0, 0), (S, S, k))
        raw = head.a.reshape(-1)
        ok = bool((raw == 0).any() and (raw == 65535).any())
        ok = ok and np.float32(agg.min).tobytes() == np.float32(ra.min).tobytes() and np.float32(agg.max).tobytes() == np.float32(ra.max).tobytes()
        ok = ok and list(agg.argmin) == list(ra.argmin) and list(agg.argmax) == list(ra.argmax)
        vals = torch.as_tensor(_Raw(a.ptr, n, "<i2"), device="cuda")
        tot = 0.0
        for f0 in range(0, n, 1 << 27):
            tot += float((vals[f0:f0 + (1 << 27)].to(torch.int32) & 0xFFFF).sum(dtype=torch.float64).item())
        real = lo * n + (hi - lo) * tot / 65535.999
        ok = ok and abs(agg.sum - real) <= 2e-6 * abs(real)
        torch.cuda.empty_cache()
        return parity.record(name, ok, "extrema and positions exact vs the oracle (first planes hold 0 and 65535); sum within 2e-6")

    record(f"Aggregates {S}^3 UInt16", n, 2 * n,
           lambda: capi.check(lib.vktb_aggregates_range(ctypes.byref(da), z3, l3, ctypes.byref(agg))), check=check_agg)
    record(f"Aggregates {S}^3 UInt16 mapping (-1000,3000)", n, 2 * n,
           lambda: capi.check(lib.vktb_aggregates_range(ctypes.byref(gm), z3, l3, ctypes.byref(agg))),
           check=lambda nm: check_agg(nm, -1000.0, 3000.0))
    b.free(); d.free()

    # ---- Resample S^3 -> (S/2)^3 and -> (3S/2)^3, UInt16 (BASELINE config 3)
    # Algorithmic bytes of the ONE-TAP modes (Nearest, and reference-faithful Linear on an integer source, which is
    # bit-identical to it: frac == 0) count what the op must touch: the src rows that are hit, whole (a row is the
    # unit of DRAM locality: every other voxel of a hit row shares its 32-byte sectors), plus the dst volume:
    #     rows hit = |{(int)(y/Dy*Sy)}| * |{(int)(z/Dz*Sz)}|,  bytes = rows hit * Sx * bpv + Dx*Dy*Dz * bpv
    # 2:1 down: a quarter of the rows -> 4.29 GB + 2.15 GB at S = 2048; 2:3 up: every row -> the SURVEY 8(d) figure.
    # The 8-tap modes (true trilinear; Linear on Float32) touch every src voxel: N_src*bpv + N_dst*bpv.
    def hit(Dn, Sn):
        return len({int(np.float32(np.float32(np.float32(i) / np.float32(Dn)) * np.float32(Sn))) for i in range(Dn)})

    resample_summary = {}
    for scale_name, De in (("down", S // 2), ("up", S * 3 // 2)):
        r = DeviceVolume((De,) * 3, capi.UINT16)
        dr = r.desc()
        DD = (De,) * 3
        rows_hit = hit(De, S) ** 2
        alg_touch = rows_hit * S * 2 + 2 * De ** 3
        alg_all = 2 * n + 2 * De ** 3
        for fname, filt, alg in (("Nearest", capi.NEAREST, alg_touch), ("Linear", capi.LINEAR, alg_touch),
                                 ("Trilinear(true)", capi.TRILINEAR, alg_all)):
            note = None
            if filt == capi.LINEAR:
                note = ("reference-faithful Linear on an integer source is bit-identical to Nearest (frac==0): one tap; "
                        "alg_bytes = src rows hit (whole) + dst")
            name = f"Resample {fname} {S}^3 -> {De}^3 UInt16"
            record(name, De ** 3, alg, lambda: capi.check(lib.vktb_resample(ctypes.byref(dr), ctypes.byref(da), filt)), note,
                   check=lambda nm, f=filt: check_resample(nm, r, po.UINT16, DD, po.UINT16, dims, 1, f))
            out[name]["src_rows_hit_fraction"] = rows_hit / float(S * S)
            if filt == capi.LINEAR:
                resample_summary[scale_name] = dict(out[name], op=name)
        r.free()
    a.free()

    # ---- Float32 (S/2... keeps memory modest): Resample Linear (8-tap faithful) S^3/8 voxels x 4 B
    Sf = S // 2
    fdims = (Sf,) * 3
    fa = DeviceVolume(fdims, capi.FLOAT32)
    synth_into(fa, 1)
    for De in (Sf // 2, Sf * 3 // 2):
        r = DeviceVolume((De,) * 3, capi.FLOAT32)
        dr, dfa = r.desc(), fa.desc()
        DD = (De,) * 3
        record(f"Resample Linear(8-tap) {Sf}^3 -> {De}^3 Float32", De ** 3, 4 * Sf ** 3 + 4 * De ** 3,
               lambda: capi.check(lib.vktb_resample(ctypes.byref(dr), ctypes.byref(dfa), capi.LINEAR)),
               check=lambda nm: check_resample(nm, r, po.FLOAT32, DD, po.FLOAT32, fdims, 1, po.LINEAR))
        r.free()
    fa.free()

    # ---- UInt8 (2S, 2S, S/2 ... = 2 S^3 voxels): Fill / Transform / SafeSum (BASELINE config 4 per-GPU slab shape)
    dims8 = (2 * S, 2 * S, S // 2)
    n8 = dims8[0] * dims8[1] * dims8[2]
    a8, b8, d8 = DeviceVolume(dims8, capi.UINT8), DeviceVolume(dims8, capi.UINT8), DeviceVolume(dims8, capi.UINT8)
    synth_into(a8, 1)
    synth_into(b8, 2)
    l8 = capi.I3(*dims8)
    d_a8, d_b8, d_d8 = a8.desc(), b8.desc(), d8.desc()
    tag = f"{dims8[0]}x{dims8[1]}x{dims8[2]} UInt8"
    pre8 = fill_plane(dims8, po.UINT8) if parity is not None else None
    record(f"Fill {tag}", n8, n8, lambda: capi.check(lib.vktb_fill_range(ctypes.byref(d_d8), z3, l8, PF)),
           check=lambda nm: parity.check(nm, sample_planes(dims8[2]), lambda z: parity.fetch(d8.ptr, dims8, po.UINT8, z), lambda z: pre8))
    record(f"Transform xor_coord {tag}", n8, 2 * n8, lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(d_d8), z3, l8, 0)),
           prepare=lambda: capi.check(lib.vktb_copy_range(ctypes.byref(d_d8), ctypes.byref(d_a8), z3, l8, z3)),
           check=lambda nm: check_transform(nm, d8, po.UINT8, dims8, 1, 0))
    record(f"SafeSum {tag}", n8, 3 * n8,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(d_d8), ctypes.byref(d_a8), ctypes.byref(d_b8), z3, l8, z3)),
           check=lambda nm: check_arith(nm, d8, po.SAFESUM, po.UINT8, dims8, (1, 2), (0, 0, 0), dims8, (0, 0, 0), prefilled=False))
    bins8 = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(bins8), 256 * 8))

    def check_hist8(name):
        """all voxels: byte-value counts (torch.bincount, checker side) x the oracle's value -> bin table == bins"""
        table = np.full(256, -1, np.int64)
        tv = po.Vol(np.arange(256, dtype=np.uint8).reshape(1, 1, 256), po.UINT8)
        for v in range(256):
            hb, dropped = o.histogram_range(tv, (v, 0, 0), (v + 1, 1, 1), 256)
            if not dropped:
                table[v] = int(np.argmax(hb))
        counts = torch.zeros(256, dtype=torch.int64, device="cuda")
        vals = torch.as_tensor(_Raw(a8.ptr, n8, "|u1"), device="cuda")
        for f0 in range(0, n8, 1 << 28):
            counts += torch.bincount(vals[f0:f0 + (1 << 28)].to(torch.int32), minlength=256)
        tt = torch.from_numpy(table).cuda()
        want = torch.zeros(256, dtype=torch.int64, device="cuda").index_add_(0, tt.clamp(min=0), counts * (tt >= 0))
        got = np.empty(256, np.uint64)
        capi.check(lib.vktb_memcpy(got.ctypes.data, bins8.value, 256 * 8, capi.D2H))
        torch.cuda.empty_cache()
        return parity.record(name, got.astype(np.int64).tolist() == want.cpu().tolist() and int(got.sum()) == n8,
                             "all voxels: value counts x the oracle's value->bin table == bins")

    record(f"Histogram 256 bins {tag}", n8, n8,
           lambda: capi.check(lib.vktb_histogram_range(ctypes.byref(d_a8), z3, l8, bins8, 256, None)), check=check_hist8)
    capi.check(lib.vktb_free(bins8))
    a8.free(); b8.free(); d8.free()

    # ---- headless ray-marcher (last row): 1024^3 UInt8, the viewer's defaults (45 degrees,
    # view_all, dt 1, sRGB), 1024 x 1024 pixel-centre rays into an RGBA8 device buffer
    R, VW = min(1024, S), 1024
    rdims = (R, R, R)
    rv = DeviceVolume(rdims, capi.UINT8)
    synth_into(rv, 1)
    cam = capi.VktbCamera()
    capi.check(lib.vktb_render_default_camera(capi.I3(*rdims), None, ctypes.byref(cam)))
    img = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(img), VW * VW * 4))
    drv = rv.desc()
    render = lambda w: capi.check(lib.vktb_render_raymarch(ctypes.byref(drv), None, ctypes.byref(cam), None, 0, 1.0, 1, w, w,  # noqa: E731
                                                           img, None))

    def check_render(name):
        """the oracle marches a 96 x 96 viewport through the SAME full-size volume (downloaded from the device)"""
        w = 96
        render(w)
        got = np.empty((w, w, 4), np.uint8)
        capi.check(lib.vktb_memcpy(got.ctypes.data, img.value, got.nbytes, capi.D2H))
        host = po.Vol(rv.download(), po.UINT8)
        _, want = po.render_oracle().raymarch(host, None, tuple(cam.eye), tuple(cam.center), tuple(cam.up), 45.0, w, w, 1.0, True)
        diff = np.abs(got.astype(np.int16) - want.astype(np.int16))
        return parity.record(name, int(diff.max()) <= 1 and bool((want[..., 3] > 0).any()),
                             f"96 x 96 rays through the full volume within 1 LSB per RGBA8 channel of the oracle (max {int(diff.max())})")

    # samples along the rays: every pixel whose ray hits the box marches (tfar - tnear) / dt steps; counted on the host
    # for the timed viewport from the oracle's own intersection would cost a frame, so the figure is measured: the
    # kernel's step count equals the sum over rays of ceil((tfar - tnear) / dt), estimated from the box's mean chord
    record(f"Render ray-march {R}^3 UInt8 -> {VW}x{VW} RGBA8", VW * VW, VW * VW * 4, lambda: render(VW),
           "voxels = rays; alg_bytes = the RGBA8 frame (the volume reads depend on the view); see rays_per_s",
           check=check_render)
    nm = f"Render ray-march {R}^3 UInt8 -> {VW}x{VW} RGBA8"
    out[nm]["rays_per_s"] = VW * VW / (out[nm]["ms"] * 1e-3)
    capi.check(lib.vktb_free(img))
    rv.free()
    return out, resample_summary
