"""Sharded-volume table for bench.py at N > 1 (one process per GPU): BASELINE configs 4 and 5 and the up-sampling
case that really needs a halo.  Every op is timed as the MAX over ranks of CUDA-event time on the launching stream,
halo exchange (NCCL point-to-point over NVLink) included in the Resample step and also reported on its own."""
import ctypes


def run_sharded_ops(torch, dist, lib, capi, ShardedVolume, sharded, rank, world, peak, time_steps, edge=4096, steps=3,
                    warmup=3, parity=None):
    """parity (bench_parity.Parity or None): after its timing every op is compared with the oracle on sampled planes of
    THIS rank's slab -- first, last (= the slab-boundary planes), the planes a halo feeds and four random ones -- and
    the verdicts are AND-ed over the ranks."""
    import numpy as np
    from bench import torch_synth
    from bench_parity import sample_planes
    from oracle import pyoracle as po   # checker
    out = {}
    o = po.oracle()

    def slab_fetch(vol, z, fmt):
        """global plane z of a ShardedVolume's owned slab as a (1, Y, X) array"""
        t = vol.planes(z, z + 1).cpu().numpy()
        return t.view(po.DTYPES[fmt]).reshape(1, vol.dims[1], vol.dims[0])

    def agree(name, ok):
        """AND over ranks; records the verdict"""
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        good = bool(t.item())
        out[name]["parity"] = good
        if parity is not None:
            parity.results[name]["ok"] = good
        return good

    def check_planes(name, vol, fmt, expected, extra=()):
        if parity is None:
            return
        planes = sorted(set(sample_planes(vol.z1 - vol.z0, n_random=4, lo=vol.z0, seed=99 + rank)) |
                        {z for z in extra if vol.z0 <= z < vol.z1})
        ok = parity.check(name, planes, lambda z: slab_fetch(vol, z, fmt), expected)
        agree(name, ok)

    def synth_into(t, bpv, seed, first_voxel, f32=False):
        n = t.numel() // bpv
        for f0 in range(0, n, 1 << 27):
            m = min(1 << 27, n - f0)
            v = torch_synth(torch, m, bpv, seed, first_voxel + f0, float32=f32)
            t[f0 * bpv:(f0 + m) * bpv].copy_(v.view(torch.uint8).reshape(-1))
            del v
        torch.cuda.synchronize()
        torch.cuda.empty_cache()

    def record(name, voxels_total, alg_bytes_total, fn, extra=None, check=None, prepare=None):
        ms = time_steps(torch, fn, steps, warmup, dist)
        gbs = alg_bytes_total / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gvoxels_s": voxels_total / (ms * 1e-3) / 1e9, "alg_gb_s_aggregate": gbs,
                     "frac_of_measured_peak_x_n": gbs / (peak * world), "alg_bytes": alg_bytes_total, "n_gpus": world}
        if extra:
            out[name].update(extra)
        if check is not None and parity is not None:
            if prepare is not None:
                prepare()
                fn()
            torch.cuda.synchronize()
            check(name)

    z3 = capi.I3(0, 0, 0)

    # ---- config 4: edge^3 UInt8, Fill / Transform / SafeSum, z-slab sharded, no communication -------------------
    E = edge
    dims = (E, E, E)
    ZOFF = 8                                         # z offset of the sharded SumRange line below
    a = ShardedVolume(torch, dims, capi.UINT8, rank, world, halo=(ZOFF, 0))   # room for the planes that line pulls in
    b = ShardedVolume(torch, dims, capi.UINT8, rank, world, halo=(ZOFF, 0))
    d = ShardedVolume(torch, dims, capi.UINT8, rank, world)
    first = a.z0 * E * E
    synth_into(a.data, 1, 1, first)
    synth_into(b.data, 1, 2, first)
    l3 = capi.I3(E, E, a.z1 - a.z0)
    da, db, dd = a.desc(), b.desc(), d.desc()
    n = E ** 3
    tag = f"{E}^3 UInt8 over {world} GPUs"
    U8 = po.UINT8
    pre = None
    if parity is not None:
        p_ = po.empty((E, E, 1), U8)
        o.fill_range(p_, (0, 0, 0), (E, E, 1), 0.25)
        pre = p_.a

    def arith_plane(op, z, off_z=0, last=None):
        """dst plane z of op(a, b) over [0, last) written at + (0, 0, off_z); the rest keeps the pre-fill"""
        dpl = po.Vol(pre.copy(), U8)
        pz = z - off_z
        if 0 <= pz < (last or dims)[2]:
            av, bv = po.Vol(parity.regen(1, U8, dims, pz), U8), po.Vol(parity.regen(2, U8, dims, pz), U8)
            o.arith_range(op, dpl, av, bv, (0, 0, 0), (E, E, 1), (0, 0, 0))
        return dpl.a

    def xor_plane(z):
        v = po.Vol(parity.regen(1, U8, dims, z), U8)
        o.transform_range1_window(v, (0, 0, 0), (E, E, 1), 0, z)      # the op sees the ABSOLUTE z
        return v.a

    fill = lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, 0.25))  # noqa: E731
    record(f"Fill {tag}", n, n, fill, check=lambda nm: check_planes(nm, d, U8, lambda z: pre))
    record(f"Transform xor_coord {tag}", n, 2 * n, lambda: sharded.transform_sharded(d, (0, 0, 0), dims, 0),
           prepare=lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3)),
           check=lambda nm: check_planes(nm, d, U8, xor_plane))
    record(f"SafeSum {tag}", n, 3 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3, l3, z3)),
           check=lambda nm: check_planes(nm, d, U8, lambda z: arith_plane(po.SAFESUM, z)))
    # *Range with a z offset: dst plane z reads src plane z - 8 -> 8 planes per source cross every slab boundary
    nr = E * E * (E - ZOFF)
    got = [0]

    def sum_range():
        got[0] = sharded.arith_range_sharded(capi.SUM, d, a, b, (0, 0, 0), (E, E, E - ZOFF), (0, 0, ZOFF), dist)

    # the dst planes [z0, z0 + ZOFF) of every slab but the first are computed from the neighbour's planes: check them all
    record(f"SumRange dstOffset (0,0,{ZOFF}) {tag}", nr, 3 * nr, sum_range, prepare=fill,
           check=lambda nm: check_planes(nm, d, U8, lambda z: arith_plane(po.SUM, z, ZOFF, (E, E, E - ZOFF)),
                                         extra=range(d.z0, d.z0 + ZOFF + 1)))
    t = torch.tensor([got[0]], dtype=torch.int64, device="cuda")
    dist.all_reduce(t) if dist is not None else None
    out[f"SumRange dstOffset (0,0,{ZOFF}) {tag}"]["halo_bytes_all_ranks"] = int(t.item())
    # the two ops with a real exchange step: per-rank partial results, one small collective
    last_bins, last_agg = [None], [None]

    def hist():
        last_bins[0] = sharded.histogram_sharded(a, (0, 0, 0), (E, E, E), 256, dist)

    def check_hist(name):
        """all voxels: every rank counts its slab's byte values (torch.bincount, checker side), the counts are summed
        over the ranks and folded through the oracle's value -> bin table"""
        table = np.full(256, -1, np.int64)
        vol = po.Vol(np.arange(256, dtype=np.uint8).reshape(1, 1, 256), U8)
        for v in range(256):
            bins, dropped = o.histogram_range(vol, (v, 0, 0), (v + 1, 1, 1), 256)
            if not dropped:
                table[v] = int(np.argmax(bins))
        counts = torch.zeros(256, dtype=torch.int64, device="cuda")
        flat = a.data
        for f0 in range(0, flat.numel(), 1 << 28):
            counts += torch.bincount(flat[f0:f0 + (1 << 28)].to(torch.int32), minlength=256)
        if dist is not None:
            dist.all_reduce(counts)
        tt = torch.from_numpy(table).cuda()
        want = torch.zeros(256, dtype=torch.int64, device="cuda").index_add_(0, tt.clamp(min=0), counts * (tt >= 0))
        ok = bool((last_bins[0] == want).all().item()) and int(last_bins[0].sum().item()) == n
        parity.record(name, ok, "all voxels: value counts of every slab, summed, x the oracle's value->bin table == bins")
        agree(name, ok)
        torch.cuda.empty_cache()

    def agg():
        last_agg[0] = sharded.aggregates_sharded(a, (0, 0, 0), (E, E, E), dist)

    def check_agg(name):
        """extrema: the uniform byte stream reaches 0 and 255 inside plane 0, so the oracle's min / max / argmin /
        argmax of plane 0 are the whole volume's -- every rank must hold them after the all-gather + fold"""
        head = po.Vol(parity.regen(1, U8, dims, 0), U8)
        ra = o.aggregates_range(head, (0, 0, 0), (E, E, 1))
        g = last_agg[0]
        ok = bool((head.a == 0).any() and (head.a == 255).any())
        ok = ok and np.float32(g.min).tobytes() == np.float32(ra.min).tobytes() and np.float32(g.max).tobytes() == np.float32(ra.max).tobytes()
        ok = ok and list(g.argmin) == list(ra.argmin) and list(g.argmax) == list(ra.argmax)
        tot = torch.zeros(1, dtype=torch.float64, device="cuda")
        flat = a.data
        for f0 in range(0, flat.numel(), 1 << 28):
            tot += flat[f0:f0 + (1 << 28)].sum(dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(tot)
        real = float(tot.item()) / 255.999
        ok = ok and abs(g.sum - real) <= 2e-6 * abs(real)
        parity.record(name, ok, "extrema and positions exact vs the oracle on plane 0; sum within 2e-6 of the binary64 sum")
        agree(name, ok)

    # the two ops with a real exchange step: per-rank partial results, one small collective
    record(f"Histogram 256 bins {tag} (per-rank bins + SUM all-reduce)", n, n, hist, check=check_hist)
    record(f"Aggregates {tag} (per-rank partial + 88-byte all-gather)", n, n, agg, check=check_agg)
    del a, b, d
    torch.cuda.empty_cache()

    # ---- Resample with halo exchange ---------------------------------------------------------------------------
    def resample_case(name, S, D, fmt, filt, f32):
        bpv = capi.BYTES_PER_VOXEL[fmt]
        plan = sharded.halo_plan(filt, D, S, world, lib)
        src = ShardedVolume(torch, S, fmt, rank, world, halo=sharded.halo_for(plan, S, world, rank))
        dst = ShardedVolume(torch, D, fmt, rank, world)
        synth_into(src.data, bpv, 1, src.z0 * S[0] * S[1], f32)
        halo_bytes = [0]

        def step():
            halo_bytes[0] = sharded.resample_sharded(dst, src, filt, dist, planner_lib=lib, plan=plan)

        def exchange_only():
            sharded.gather_window(src, plan[0][rank], plan[1], dist)

        nd = D[0] * D[1] * D[2]
        ns = S[0] * S[1] * S[2]
        ms_x = time_steps(torch, exchange_only, steps, warmup, dist)
        t = torch.tensor([halo_bytes[0]], dtype=torch.int64, device="cuda")
        pfmt = fmt                                       # capi and pyoracle share the DataFormat numbers

        def expected(z):
            s0 = int(z * S[2] / D[2])
            w0, w1 = max(s0 - 1, 0), min(s0 + 3, S[2])  # generous: the taps are planes s0 and s0 + 1
            sv = po.Vol(parity.regen(1, pfmt, S, w0, w1 - w0), pfmt)
            dv = po.empty((D[0], D[1], 1), pfmt)
            o.resample_window(dv, sv, filt, z, D, w0, S)
            return dv.a

        # the last dst planes of a slab are the ones that tap the next slab's first src plane (the halo): check the
        # last four and the first two of every slab besides the sampled ones
        record(name, nd, (ns + nd) * bpv, step,
               check=lambda nm: check_planes(nm, dst, pfmt, expected,
                                             extra=list(range(dst.z1 - 4, dst.z1)) + [dst.z0, dst.z0 + 1]))
        t[0] = halo_bytes[0]
        dist.all_reduce(t)
        out[name].update({"halo_exchange_ms": ms_x, "halo_bytes_all_ranks": int(t.item()),
                          "halo_transfers": len(plan[1])})
        del src, dst
        torch.cuda.empty_cache()

    H = E // 2
    # config 5: edge^3 Float32 -> (edge/2)^3, reference-faithful Linear (8 taps); exact 2:1 -> planner proves no halo
    resample_case(f"Resample Linear(8-tap) {E}^3 -> {H}^3 Float32 over {world} GPUs", (E, E, E), (H, H, H), capi.FLOAT32,
                  capi.LINEAR, True)
    # the case that needs the halo: 2:3 up-sampling, UInt16 (config 3 sharded)
    U = H * 3 // 2
    resample_case(f"Resample Linear {H}^3 -> {U}^3 UInt16 over {world} GPUs", (H, H, H), (U, U, U), capi.UINT16, capi.LINEAR,
                  False)
    resample_case(f"Resample Trilinear(true) {H}^3 -> {U}^3 UInt16 over {world} GPUs", (H, H, H), (U, U, U), capi.UINT16,
                  capi.TRILINEAR, False)
    if parity is not None:
        out["_parity"] = parity.summary()
    return out
