"""Single-process multi-GPU table: the path a `vkt::` caller gets after vktb_set_devices -- ONE process, volumes striped
over the GPUs of the box, every algorithm fanned out per stripe, migrate() of a striped volume on all copy engines.

Used two ways: rank 0 of `bench.py --gpus N` (N > 1) runs it after the multi-process table while the other ranks
wait (the driver's scaling run therefore measures this path too); stand-alone: `python bench_multi.py --gpus N`.

BASELINE configs: 4 (edge^3 UInt8 Fill / Transform / SafeSum), 3 strong-scaled (2:3 up-sampling, the case whose taps
cross slab boundaries -- here peer loads, no exchange step at all), 5 (edge^3 Float32 -> (edge/2)^3 + migrate round
trip: pinned host -> striped volume and back, all copy engines at once).  Every result is compared with the oracle on
sampled planes including the planes either side of every stripe boundary."""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def run_multi(torch, lib, capi, n_gpus, peak, time_steps, edge=4096, steps=3, warmup=3, parity=None):
    import numpy as np
    from bench import torch_synth, mem_available_bytes
    from bench_parity import sample_planes
    from oracle import pyoracle as po   # checker
    from volkit_b200.device import DeviceVolume

    out = {}
    o = po.oracle()
    ids = (ctypes.c_int * n_gpus)(*range(n_gpus))
    capi.check(lib.vktb_set_devices(ids, n_gpus), "vktb_set_devices")
    z3 = capi.I3(0, 0, 0)

    def synth_into(vol, seed, f32=False):
        bpv = capi.BYTES_PER_VOXEL[vol.fmt]
        n = vol.nbytes // bpv
        for f0 in range(0, n, 1 << 27):
            m = min(1 << 27, n - f0)
            t = torch_synth(torch, m, bpv, seed, f0, float32=f32)
            capi.check(lib.vktb_memcpy(vol.ptr + f0 * bpv, t.data_ptr(), m * bpv, capi.D2D))
            torch.cuda.synchronize()
            del t
        torch.cuda.empty_cache()

    def stripe_boundaries(vol):
        """planes that contain a stripe boundary (equal contiguous parts of the allocation, 2 MiB granularity)"""
        plane = vol.dims[0] * vol.dims[1] * capi.BYTES_PER_VOXEL[vol.fmt]
        gran = 2 << 20
        total = (vol.nbytes + gran - 1) // gran * gran
        chunk = ((total // gran + n_gpus - 1) // n_gpus) * gran
        return [min(vol.dims[2] - 1, (g * chunk) // plane) for g in range(1, n_gpus) if g * chunk < vol.nbytes]

    def record(name, voxels, alg_bytes, fn, check=None, prepare=None, extra=None):
        ms = time_steps(torch, fn, steps, warmup)
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gvoxels_s": voxels / (ms * 1e-3) / 1e9, "alg_gb_s_aggregate": gbs,
                     "frac_of_measured_peak_x_n": gbs / (peak * n_gpus), "alg_bytes": alg_bytes, "n_gpus": n_gpus}
        if extra:
            out[name].update(extra)
        if check is not None and parity is not None:
            if prepare is not None:
                prepare()
                fn()
            out[name]["parity"] = bool(check(name))

    def planes_of(vol):
        b = stripe_boundaries(vol)
        return sample_planes(vol.dims[2], boundaries=b + [z + 1 for z in b], n_random=4)

    # ---- config 4: edge^3 UInt8 ----------------------------------------------------------------------------------
    E = edge
    dims = (E, E, E)
    U8 = po.UINT8
    a, b, d = DeviceVolume(dims, capi.UINT8), DeviceVolume(dims, capi.UINT8), DeviceVolume(dims, capi.UINT8)
    st = ctypes.c_int()
    assert lib.vktb_is_striped(a.ptr, ctypes.byref(st)) == 1, "the bench volumes must be striped"
    synth_into(a, 1)
    synth_into(b, 2)
    l3 = capi.I3(*dims)
    da, db, dd = a.desc(), b.desc(), d.desc()
    n = E ** 3
    tag = f"{E}^3 UInt8, one process, {n_gpus} GPUs ({st.value} stripes)"
    pre = None
    if parity is not None:
        p_ = po.empty((E, E, 1), U8)
        o.fill_range(p_, (0, 0, 0), (E, E, 1), 0.25)
        pre = p_.a

    def arith_plane(z):
        dpl = po.empty((E, E, 1), U8)
        o.arith_range(po.SAFESUM, dpl, po.Vol(parity.regen(1, U8, dims, z), U8), po.Vol(parity.regen(2, U8, dims, z), U8),
                      (0, 0, 0), (E, E, 1))
        return dpl.a

    def xor_plane(z):
        v = po.Vol(parity.regen(1, U8, dims, z), U8)
        o.transform_range1_window(v, (0, 0, 0), (E, E, 1), 0, z)
        return v.a

    fetch = lambda vol, fmt: (lambda z: parity.fetch(vol.ptr, vol.dims, fmt, z))  # noqa: E731
    record(f"Fill {tag}", n, n, lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, 0.25)),
           check=lambda nm: parity.check(nm, planes_of(d), fetch(d, U8), lambda z: pre))
    record(f"Transform xor_coord {tag}", n, 2 * n, lambda: capi.check(lib.vktb_transform_range1(ctypes.byref(dd), z3, l3, 0)),
           prepare=lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(da), z3, l3, z3)),
           check=lambda nm: parity.check(nm, planes_of(d), fetch(d, U8), xor_plane))
    record(f"SafeSum {tag}", n, 3 * n,
           lambda: capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3, l3, z3)),
           check=lambda nm: parity.check(nm, planes_of(d), fetch(d, U8), arith_plane))
    # *Range with a z offset: dst plane z reads source plane z - 8 -- across a stripe boundary that is a peer load
    # inside the kernel (the multi-process form moves 8 planes per source and boundary with NCCL first)
    ZOFF = 8
    nr = E * E * (E - ZOFF)

    def sumrange_plane(z):
        dpl = po.Vol(pre.copy(), U8)
        pz = z - ZOFF
        if 0 <= pz < E - ZOFF:
            o.arith_range(po.SUM, dpl, po.Vol(parity.regen(1, U8, dims, pz), U8), po.Vol(parity.regen(2, U8, dims, pz), U8),
                          (0, 0, 0), (E, E, 1))
        return dpl.a

    def planes_zoff(vol):
        bnd = stripe_boundaries(vol)
        # dst planes [b, b + ZOFF) of the device behind boundary b read source planes of the device in front of it
        extra = [z + k for z in bnd for k in (-1, 0, 1, ZOFF - 1, ZOFF, ZOFF + 1)]
        return sorted(set(planes_of(vol)) | {z for z in extra if 0 <= z < vol.dims[2]})

    record(f"SumRange dstOffset (0,0,{ZOFF}) {tag}", nr, 3 * nr,
           lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(da), ctypes.byref(db), z3,
                                                   capi.I3(E, E, E - ZOFF), capi.I3(0, 0, ZOFF))),
           prepare=lambda: capi.check(lib.vktb_fill_range(ctypes.byref(dd), z3, l3, 0.25)),
           check=lambda nm: parity.check(nm, planes_zoff(d), fetch(d, U8), sumrange_plane))
    a.free(); b.free(); d.free()

    # ---- Resample ---------------------------------------------------------------------------------------------------
    def resample_case(name, S, D, fmt, pfmt, filt, f32, migrate=False):
        bpv = capi.BYTES_PER_VOXEL[fmt]
        src, dst = DeviceVolume(S, fmt), DeviceVolume(D, fmt)
        synth_into(src, 1, f32)
        ds, dd_ = src.desc(), dst.desc()

        def expected(z):
            s0 = int(z * S[2] / D[2])
            w0, w1 = max(s0 - 1, 0), min(s0 + 3, S[2])
            sv = po.Vol(parity.regen(1, pfmt, S, w0, w1 - w0), pfmt)
            dv = po.empty((D[0], D[1], 1), pfmt)
            o.resample_window(dv, sv, filt, z, D, w0, S)
            return dv.a

        nd, ns = D[0] * D[1] * D[2], S[0] * S[1] * S[2]
        extra = None
        if migrate:
            # migrate round trip of the dst-sized volume: pinned host -> striped device range and back
            nbytes = dst.nbytes
            if mem_available_bytes() > nbytes + (16 << 30):
                host = ctypes.c_void_p()
                capi.check(lib.vktb_host_alloc(ctypes.byref(host), nbytes))
                ctypes.memset(host, 1, nbytes)
                t_in = time_steps(torch, lambda: capi.check(lib.vktb_memcpy(dst.ptr, host, nbytes, capi.H2D)), 2, 1)
                t_out = time_steps(torch, lambda: capi.check(lib.vktb_memcpy(host, dst.ptr, nbytes, capi.D2H)), 2, 1)
                capi.check(lib.vktb_host_free(host))
                capi.check(lib.vktb_trim_pools())
                extra = {"migrate_bytes": nbytes, "migrate_in_ms": t_in, "migrate_out_ms": t_out,
                         "migrate_in_gb_s": nbytes / t_in / 1e6, "migrate_out_gb_s": nbytes / t_out / 1e6,
                         "migrate_note": "vktb_memcpy of the striped dst-sized volume from / to one pinned host buffer: "
                                         "every device's copy engine moves its own byte range"}
        record(name, nd, (ns + nd) * bpv, lambda: capi.check(lib.vktb_resample(ctypes.byref(dd_), ctypes.byref(ds), filt)),
               check=lambda nm: parity.check(nm, planes_of(dst), fetch(dst, pfmt), expected), extra=extra)
        out[name]["halo"] = "none as a step: taps in a neighbour's stripe are peer loads inside the kernel"
        src.free(); dst.free()

    H = E // 2
    U = H * 3 // 2
    resample_case(f"Resample Linear(8-tap) {E}^3 -> {H}^3 Float32 + migrate round trip, one process, {n_gpus} GPUs",
                  (E, E, E), (H, H, H), capi.FLOAT32, po.FLOAT32, capi.LINEAR, True, migrate=True)
    resample_case(f"Resample Linear {H}^3 -> {U}^3 UInt16, one process, {n_gpus} GPUs", (H, H, H), (U, U, U), capi.UINT16,
                  po.UINT16, capi.LINEAR, False)
    one = (ctypes.c_int * 1)(0)
    capi.check(lib.vktb_set_devices(one, 1))
    if parity is not None:
        out["_parity"] = parity.summary()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--edge", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import torch
    from bench import measured_peak, time_steps
    from bench_parity import Parity
    from volkit_b200 import capi
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    capi.check(lib.vktb_set_stream(stream.cuda_stream))
    peak, _ = measured_peak()
    par = Parity(lib, capi)
    out = run_multi(torch, lib, capi, args.gpus, peak, time_steps, edge=args.edge, steps=args.steps, parity=par)
    print(json.dumps({"single_process_multi_gpu": out, "n_gpus": args.gpus, "edge": args.edge}))


if __name__ == "__main__":
    main()
