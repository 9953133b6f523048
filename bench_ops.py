"""Per-op table for bench.py (N=1): every in-scope op at the BASELINE sizes that fit one B200, each with GVoxels/s,
algorithmic GB/s (SURVEY.md 8(d) bytes) and fraction of the measured HBM peak.  Inputs are resident in HBM."""
import ctypes


def run_ops(torch, lib, capi, DeviceVolume, S, peak, time_steps, steps=5, warmup=3):
    from bench import torch_synth

    out = {}

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

    def record(name, voxels, alg_bytes, fn, note=None):
        ms = time_steps(torch, fn, steps, warmup)
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gvoxels_s": voxels / (ms * 1e-3) / 1e9, "alg_gb_s": gbs, "frac_of_measured_peak": gbs / peak,
                     "frac_of_8TBs": gbs / 8000.0, "alg_bytes": alg_bytes}
        if note:
            out[name]["note"] = note

    z3 = capi.I3(0, 0, 0)
    n = S ** 3

    # ---- UInt16 S^3: Fill, Copy, CopyRange(dstOffset), SafeDiff, Sum (general mapping), Transform
    a, b, d = DeviceVolume((S,) * 3, capi.UINT16), DeviceVolume((S,) * 3, capi.UINT16), DeviceVolume((S,) * 3, capi.UINT16)
    synth_into(a, 1)
    synth_into(b, 2)
    l3 = capi.I3(S, S, S)
    da, db, dd = a.desc(), b.desc(), d.desc()
    record(f"Fill {S}^3 UInt16", n, 2 * n, lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, 0.25)))
    record(f"Copy {S}^3 UInt16", n, 4 * n, lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3)))
    m = S - 8
    record(f"CopyRange {m}^3 UInt16 dstOffset (8,8,8)", m ** 3, 4 * m ** 3,
           lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, capi.I3(m, m, m), capi.I3(8, 8, 8))))
    record(f"SafeDiff {S}^3 UInt16", n, 6 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFEDIFF, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3, l3, z3)))
    record(f"SumRange {m}^3 UInt16 dstOffset (8,8,8)", m ** 3, 6 * m ** 3,
           lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3,
                                                   capi.I3(m, m, m), capi.I3(8, 8, 8))))
    # sources at another 16-byte phase than the destination (funnel-shifted loads)
    record(f"CopyRange {m}^3 UInt16 first (1,1,1) dstOffset (0,0,0)", m ** 3, 4 * m ** 3,
           lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), capi.I3(1, 1, 1),
                                                  capi.I3(m + 1, m + 1, m + 1), z3)))
    record(f"SumRange {m}^3 UInt16 dstOffset (1,2,3)", m ** 3, 6 * m ** 3,
           lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3,
                                                   capi.I3(m, m, m), capi.I3(1, 2, 3))))
    gm = capi.vol(a.ptr, (S, S, S), capi.UINT16, -1000.0, 3000.0)
    gd = capi.vol(d.ptr, (S, S, S), capi.UINT16, -1000.0, 3000.0)
    record(f"SafeProd {S}^3 UInt16 mapping (-1000,3000)", n, 6 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFEPROD, ctypes.byref(gd), ctypes.byref(gm), ctypes.byref(db), z3, l3, z3)))
    record(f"Transform xor_coord {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(dd), z3, l3, 0)))
    # Flip (each visited voxel and its partner are read and written once) and Histogram (one read pass)
    for axis, name in ((0, "X"), (1, "Y (mirrors y and z, as the reference)")):
        record(f"Flip {name} {S}^3 UInt16", n, 4 * (n if S % 2 == 0 else n - S * S),
               lambda: capi.check(lib.vktb_flip_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3, axis)))
    # Scan, in place: x, y and z prefix passes
    record(f"Scan {S}^3 UInt16 in place", n, 4 * n,
           lambda: capi.check(lib.vktb_scan_range(ctypes.byref(dd), ctypes.byref(dd), z3, l3, z3)),
           "alg_bytes = one read + one write per voxel (the ideal single pass); the kernel makes three passes")
    # BrickDecompose: 64^3 bricks, halo 1 on every side, all bricks in ONE launch
    bs, halo = 64, 1
    nb = (S + bs - 1) // bs
    edge = lambda i: halo + min(bs, S - i * bs) + halo
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
    descs = (capi.VktbVol * len(bricks))(*[capi.vol(pool.value + 2 * o, bd, capi.UINT16, 0.0, 1.0) for o, bd in bricks])
    i32 = lambda v: (ctypes.c_int32 * len(v))(*v)
    bf, bl, bo = i32(firsts), i32(lasts), i32([0] * len(firsts))
    record(f"BrickDecompose {S}^3 UInt16 -> {len(bricks)} bricks of {bs}^3 + halo {halo}", total, 4 * total,
           lambda: capi.check(lib.vktb_copy_range_batch(len(bricks), descs, ctypes.byref(da), bf, bl, bo)),
           "one launch for all bricks; voxels = brick voxels incl. halos, alg_bytes = 2 bpv per brick voxel")
    capi.check(lib.vktb_free(pool))
    bins = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(bins), 256 * 8))
    record(f"Histogram 256 bins {S}^3 UInt16", n, 2 * n,
           lambda: capi.check(lib.vktb_histogram_range(ctypes.byref(da), z3, l3, bins, 256, None)))
    capi.check(lib.vktb_free(bins))
    # Scale / Rotate / ApplyFilter
    F9, F3 = ctypes.c_float * 9, ctypes.c_float * 3
    half = F3(S / 2, S / 2, S / 2)
    record(f"Scale x1.5 about the centre {S}^3 UInt16", n, int(2 * n + 2 * n / 3.375),
           lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(da), z3, l3,
                                                    F9(1 / 1.5, 0, 0, 0, 1 / 1.5, 0, 0, 0, 1 / 1.5), half)),
           "every dst voxel is written; the (1/1.5)^3 of the source it maps to is read once")
    import math
    cs, sn = math.cos(0.5), math.sin(0.5)
    record(f"Rotate 0.5 rad about z through the centre {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(da), z3, l3,
                                                    F9(cs, -sn, 0, sn, cs, 0, 0, 0, 1), half)),
           "alg_bytes = one read + one write per voxel (upper bound: voxels that rotate out of the volume are skipped)")
    box = (ctypes.c_float * 27)(*([1.0 / 27.0] * 27))
    record(f"ApplyFilter 3x3x3 {S}^3 UInt16", n, 4 * n,
           lambda: capi.check(lib.vktb_apply_filter(ctypes.byref(dd), ctypes.byref(da), box, capi.I3(3, 3, 3), 2)),
           "27 FMUL + 27 FADD per voxel in the reference's order: CUDA-core bound, see DESIGN.md")
    # Aggregates: one read pass, result on the host (the call is synchronous)
    agg = capi.VktbAggregates()
    record(f"Aggregates {S}^3 UInt16", n, 2 * n,
           lambda: capi.check(lib.vktb_aggregates_range(ctypes.byref(da), z3, l3, ctypes.byref(agg))))
    record(f"Aggregates {S}^3 UInt16 mapping (-1000,3000)", n, 2 * n,
           lambda: capi.check(lib.vktb_aggregates_range(ctypes.byref(gm), z3, l3, ctypes.byref(agg))))
    b.free(); d.free()

    # ---- Resample S^3 -> (S/2)^3 and -> (3S/2)^3, UInt16 (BASELINE config 3)
    for scale_name, De in (("down", S // 2), ("up", S * 3 // 2)):
        r = DeviceVolume((De,) * 3, capi.UINT16)
        dr = r.desc()
        alg = 2 * n + 2 * De ** 3
        for fname, filt in (("Nearest", capi.NEAREST), ("Linear", capi.LINEAR), ("Trilinear(true)", capi.TRILINEAR)):
            note = None
            if filt == capi.LINEAR:
                note = ("reference-faithful Linear on an integer source is bit-identical to Nearest (frac==0); "
                        "1 tap is read, alg_bytes still counts the whole source (SURVEY 8(d))")
            record(f"Resample {fname} {S}^3 -> {De}^3 UInt16", De ** 3, alg,
                   lambda: capi.check(lib.vktb_resample(ctypes.byref(dr), ctypes.byref(da), filt)), note)
        r.free()
    a.free()

    # ---- Float32 (S/2... keeps memory modest): Resample Linear (8-tap faithful) S^3/8 voxels x 4 B
    Sf = S // 2
    fa = DeviceVolume((Sf,) * 3, capi.FLOAT32)
    synth_into(fa, 1)
    for De in (Sf // 2, Sf * 3 // 2):
        r = DeviceVolume((De,) * 3, capi.FLOAT32)
        dr, dfa = r.desc(), fa.desc()
        record(f"Resample Linear(8-tap) {Sf}^3 -> {De}^3 Float32", De ** 3, 4 * Sf ** 3 + 4 * De ** 3,
               lambda: capi.check(lib.vktb_resample(ctypes.byref(dr), ctypes.byref(dfa), capi.LINEAR)))
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
    record(f"Fill {tag}", n8, n8, lambda: capi.check(lib.vktb_fill_range(ctypes.byref(d_d8), z3, l8, 0.25)))
    record(f"Transform xor_coord {tag}", n8, 2 * n8, lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(d_d8), z3, l8, 0)))
    record(f"SafeSum {tag}", n8, 3 * n8,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(d_d8), ctypes.byref(d_a8), ctypes.byref(d_b8), z3, l8, z3)))
    a8.free(); b8.free(); d8.free()
    return out
