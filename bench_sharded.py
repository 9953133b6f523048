"""Sharded-volume table for bench.py at N > 1 (one process per GPU): BASELINE configs 4 and 5 and the up-sampling
case that really needs a halo.  Every op is timed as the MAX over ranks of CUDA-event time on the launching stream,
halo exchange (NCCL point-to-point over NVLink) included in the Resample step and also reported on its own."""
import ctypes


def run_sharded_ops(torch, dist, lib, capi, ShardedVolume, sharded, rank, world, peak, time_steps, edge=4096, steps=3,
                    warmup=3):
    from bench import torch_synth
    out = {}

    def synth_into(t, bpv, seed, first_voxel, f32=False):
        n = t.numel() // bpv
        for f0 in range(0, n, 1 << 27):
            m = min(1 << 27, n - f0)
            v = torch_synth(torch, m, bpv, seed, first_voxel + f0, float32=f32)
            t[f0 * bpv:(f0 + m) * bpv].copy_(v.view(torch.uint8).reshape(-1))
            del v
        torch.cuda.synchronize()
        torch.cuda.empty_cache()

    def record(name, voxels_total, alg_bytes_total, fn, extra=None):
        ms = time_steps(torch, fn, steps, warmup, dist)
        gbs = alg_bytes_total / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gvoxels_s": voxels_total / (ms * 1e-3) / 1e9, "alg_gb_s_aggregate": gbs,
                     "frac_of_measured_peak_x_n": gbs / (peak * world), "alg_bytes": alg_bytes_total, "n_gpus": world}
        if extra:
            out[name].update(extra)

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
    record(f"Fill {tag}", n, n, lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, 0.25)))
    record(f"Transform xor_coord {tag}", n, 2 * n, lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(dd), z3, l3, 0)))
    record(f"SafeSum {tag}", n, 3 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3, l3, z3)))
    # *Range with a z offset: dst plane z reads src plane z - 8 -> 8 planes per source cross every slab boundary
    nr = E * E * (E - ZOFF)
    got = [0]

    def sum_range():
        got[0] = sharded.arith_range_sharded(capi.SUM, d, a, b, (0, 0, 0), (E, E, E - ZOFF), (0, 0, ZOFF), dist)

    record(f"SumRange dstOffset (0,0,{ZOFF}) {tag}", nr, 3 * nr, sum_range)
    t = torch.tensor([got[0]], dtype=torch.int64, device="cuda")
    dist.all_reduce(t) if dist is not None else None
    out[f"SumRange dstOffset (0,0,{ZOFF}) {tag}"]["halo_bytes_all_ranks"] = int(t.item())
    # the two ops with a real exchange step: per-rank partial results, one small collective
    record(f"Histogram 256 bins {tag} (per-rank bins + SUM all-reduce)", n, n,
           lambda: sharded.histogram_sharded(a, (0, 0, 0), (E, E, E), 256, dist))
    record(f"Aggregates {tag} (per-rank partial + 88-byte all-gather)", n, n,
           lambda: sharded.aggregates_sharded(a, (0, 0, 0), (E, E, E), dist))
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
        record(name, nd, (ns + nd) * bpv, step)
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
    return out
