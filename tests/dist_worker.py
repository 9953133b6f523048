"""One rank of the multi-rank sharding tests.  Spawned by tests/test_sharded_gloo.py (CPU tensors, gloo backend, the
oracle plugged in as the voxel compute so that the HOST-SIDE logic -- slab ranges, halo plan, point-to-point plane
exchange, window offsets -- is checked without a GPU) and by tests/test_sharded_gpu.py / bench (CUDA, NCCL, the
real C-ABI kernels).  Every rank checks its own dst slab against the whole-volume oracle result."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import pyoracle as po  # noqa: E402  (checker)
import cases  # noqa: E402  (tests/cases.py: tolerance contract of the aggregates)
from volkit_b200 import capi, sharded  # noqa: E402

CASES = [
    # (S, D, src fmt, dst fmt, filter)
    ((20, 12, 17), (30, 18, 23), po.UINT16, po.UINT16, po.NEAREST),
    ((20, 12, 17), (30, 18, 23), po.FLOAT32, po.FLOAT32, po.LINEAR),
    ((20, 12, 17), (30, 18, 23), po.UINT16, po.UINT8, po.TRILINEAR),
    ((16, 8, 32), (8, 4, 16), po.FLOAT32, po.FLOAT32, po.LINEAR),       # exact 2:1 -> planner proves "no halo"
    ((16, 8, 9), (24, 12, 31), po.UINT8, po.UINT16, po.LINEAR),         # thin src slabs: planes from non-neighbours
    ((12, 10, 14), (12, 10, 14), po.UINT16, po.FLOAT32, po.NEAREST),    # equal dims: conversion only
    ((64, 16, 24), (96, 24, 36), po.UINT16, po.UINT16, po.LINEAR),      # 16-byte aligned rows: TMA row pipeline on GPU
]


def np_view(desc, fmt):
    n = desc.dims[0] * desc.dims[1] * desc.dims[2]
    dt = np.dtype(po.DTYPES[fmt])
    buf = (ctypes.c_uint8 * (n * dt.itemsize)).from_address(desc.data)
    return np.frombuffer(buf, dt).reshape(desc.dims[2], desc.dims[1], desc.dims[0])


def oracle_compute(sfmt, dfmt, lo_hi_s, lo_hi_d):
    o = po.oracle()

    def run(dst_desc, src_desc, filt, dz0, D, sz0, S):
        d = po.Vol(np_view(dst_desc, dfmt), dfmt, *lo_hi_d)
        s = po.Vol(np_view(src_desc, sfmt), sfmt, *lo_hi_s)
        o.resample_window(d, s, filt, dz0, D, sz0, S)
    return run


def main():
    import torch
    import torch.distributed as dist
    backend = sys.argv[1]                 # gloo | nccl | gloo-cuda (several ranks on ONE GPU, planes staged by gloo)
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    gpu = backend != "gloo"
    if gpu:
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
            dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
        else:
            torch.cuda.set_device(0)
            dist.init_process_group("gloo")
        lib = capi.load()
        capi.check(lib.vktb_set_device(torch.cuda.current_device()))
        # a side stream as torch's current stream; the backend is NOT told: sharded.bound_stream must do the ordering
        stream = torch.cuda.Stream()
        torch.cuda.set_stream(stream)
    else:
        dist.init_process_group("gloo")
        lib = None
    device = "cuda" if gpu else "cpu"
    o = po.oracle()
    total_halo = 0
    for case_no, (S, D, sfmt, dfmt, filt) in enumerate(CASES):
        smap, dmap = (0.0, 1.0), (-1.0, 2.0)
        whole_src = po.Vol(po.synth(S, sfmt, 1), sfmt, *smap)
        ref = po.empty(D, dfmt, *dmap)
        o.resample(ref, whole_src, filt)

        # odd cases assemble the window in a fresh buffer, even ones in place (slab allocated with halo room)
        plan = sharded.halo_plan(filt, D, S, world, lib)
        halo = sharded.halo_for(plan, S, world, rank) if case_no % 2 == 0 else (0, 0)
        src = sharded.ShardedVolume(torch, S, sfmt, rank, world, *smap, device=device, halo=halo)
        dst = sharded.ShardedVolume(torch, D, dfmt, rank, world, *dmap, device=device)
        mine = np.ascontiguousarray(whole_src.a[src.z0:src.z1]).view(np.uint8).ravel()
        src.data.copy_(torch.from_numpy(mine.copy()))
        dst.data.zero_()

        compute = None if gpu else oracle_compute(sfmt, dfmt, smap, dmap)
        halo = sharded.resample_sharded(dst, src, filt, dist, compute=compute, planner_lib=lib, plan=plan)
        if gpu:
            torch.cuda.synchronize()
        total_halo += halo
        got = dst.data.cpu().numpy().view(po.DTYPES[dfmt]).reshape(dst.z1 - dst.z0, D[1], D[0])
        want = ref.a[dst.z0:dst.z1]
        same = got.tobytes() == want.tobytes()
        if not same and dfmt == po.FLOAT32:
            gn, wn = np.isnan(got), np.isnan(want)
            same = (gn == wn).all() and got[~gn].tobytes() == want[~wn].tobytes()
        if not same:
            print(f"rank {rank}: MISMATCH in case {S}->{D} fmt {sfmt}->{dfmt} filter {filt}", flush=True)
            sys.exit(1)
        # the planner's "no halo" claim for exact 2:1 down-sampling (SURVEY.md 8(e) caveat)
        if S == (16, 8, 32) and world == 2 and halo != 0:
            print(f"rank {rank}: expected no halo for exact 2:1, got {halo} bytes", flush=True)
            sys.exit(1)

    # *Range calls with a z offset / clamped Copy: dst slab <- window of src planes (same plane transfers)
    def fill_slab(vol, whole):
        part = np.ascontiguousarray(whole.a[vol.z0:vol.z1]).view(np.uint8).ravel()
        vol.data.copy_(torch.from_numpy(part.copy()))

    def check_slab(vol, ref, what):
        got = vol.data.cpu().numpy().view(po.DTYPES[vol.fmt]).reshape(vol.z1 - vol.z0, vol.dims[1], vol.dims[0])
        if got.tobytes() != ref.a[vol.z0:vol.z1].tobytes():
            print(f"rank {rank}: MISMATCH in {what}", flush=True)
            sys.exit(1)

    def as_vol(desc, fmt, lo_hi=(0.0, 1.0)):
        return po.Vol(np_view(desc, fmt), fmt, *lo_hi)

    fmt = po.UINT16
    Sd, Dd = (24, 10, 19), (28, 14, 26)
    whole_a, whole_b = po.Vol(po.synth(Sd, fmt, 1), fmt), po.Vol(po.synth(Sd, fmt, 2), fmt)
    for first, last, off in (((-3, -2, -4), (22, 9, 20), (1, 2, 2)),     # clamped on both z ends (24 planes land in dst planes 2..25)
                             ((2, 1, 3), (20, 8, 17), (5, 4, 9)),         # interior, large z offset
                             ((0, 0, 0), (24, 10, 19), (0, 0, 0))):       # congruent part: no transfers at all
        whole_d = po.Vol(po.synth(Dd, fmt, 3), fmt)
        ref = whole_d.copy()
        o.copy_range(ref, whole_a, first, last, off)
        a = sharded.ShardedVolume(torch, Sd, fmt, rank, world, device=device)
        d = sharded.ShardedVolume(torch, Dd, fmt, rank, world, device=device)
        fill_slab(a, whole_a); fill_slab(d, whole_d)
        compute = None if gpu else (lambda dd, sd, f, l, of: o.copy_range(as_vol(dd, fmt), as_vol(sd, fmt), f, l, of))
        total_halo += sharded.copy_range_sharded(d, a, first, last, off, dist, compute=compute)
        if gpu:
            torch.cuda.synchronize()
        check_slab(d, ref, f"CopyRange {first} {last} {off}")
    for op, first, last, off in ((po.SAFESUM, (1, 2, 3), (23, 9, 18), (4, 3, 7)),
                                 (po.DIFF, (0, 0, 0), (24, 10, 19), (2, 1, 0)),        # no z offset: no transfers
                                 (po.SAFEPROD, (0, 0, 5), (24, 10, 6), (0, 0, 20))):   # a single plane far away
        whole_d = po.Vol(po.synth(Dd, fmt, 3), fmt)
        ref = whole_d.copy()
        o.arith_range(op, ref, whole_a, whole_b, first, last, off)
        a = sharded.ShardedVolume(torch, Sd, fmt, rank, world, device=device)
        b = sharded.ShardedVolume(torch, Sd, fmt, rank, world, device=device)
        d = sharded.ShardedVolume(torch, Dd, fmt, rank, world, device=device)
        fill_slab(a, whole_a); fill_slab(b, whole_b); fill_slab(d, whole_d)
        compute = None if gpu else (lambda op_, dd, ad, bd, f, l, of:
                                    o.arith_range(op_, as_vol(dd, fmt), as_vol(ad, fmt), as_vol(bd, fmt), f, l, of))
        total_halo += sharded.arith_range_sharded(op, d, a, b, first, last, off, dist, compute=compute)
        if gpu:
            torch.cuda.synchronize()
        check_slab(d, ref, f"arith {op} {first} {last} {off}")

    # Histogram: per-rank bins + one SUM all-reduce (the only collective on the path)
    for hfmt, lo, hi in ((po.UINT16, 0.0, 1.0), (po.UINT8, -1.0, 2.0)):
        whole = po.Vol(po.synth((33, 12, 21), hfmt, 4), hfmt, lo, hi)
        hv = sharded.ShardedVolume(torch, (33, 12, 21), hfmt, rank, world, lo, hi, device=device)
        fill_slab(hv, whole)
        for first, last, nb in (((0, 0, 0), (33, 12, 21), 256), ((2, 1, 5), (30, 11, 17), 37)):
            ref_bins, dropped = o.histogram_range(whole, first, last, nb)
            assert dropped == 0

            def cpu_hist(desc, f, l, bins_ptr, n):
                b, _ = o.histogram_range(as_vol(desc, hfmt, (lo, hi)), f, l, n)
                ctypes.memmove(bins_ptr, b.ctypes.data, n * 8)

            got = sharded.histogram_sharded(hv, first, last, nb, dist, compute=None if gpu else cpu_hist)
            if gpu:
                torch.cuda.synchronize()
            if got.cpu().numpy().astype(np.uint64).tobytes() != ref_bins.tobytes():
                print(f"rank {rank}: MISMATCH in histogram {hfmt} {first} {last} {nb}", flush=True)
                sys.exit(1)

    # Aggregates: one partial per rank, all-gathered and folded in z order
    for afmt, lo, hi in ((po.UINT16, 0.0, 1.0), (po.FLOAT32, -1.0, 2.0)):
        whole = po.Vol(po.synth((33, 12, 21), afmt, 6), afmt, lo, hi)
        if afmt == po.UINT16:
            whole.a[:] = whole.a % 7                 # many ties: the first occurrence must win across slabs
        av = sharded.ShardedVolume(torch, (33, 12, 21), afmt, rank, world, lo, hi, device=device)
        fill_slab(av, whole)
        for first, last in (((0, 0, 0), (33, 12, 21)), ((2, 1, 5), (30, 11, 17))):
            ref_agg = o.aggregates_range(whole, first, last)

            def cpu_partial(desc, f, l, part):
                # test double of vktb_aggregates_partial: float64 moments of the slab's planes
                sub = as_vol(desc, afmt, (lo, hi))
                vals = cases.unmapped(o, sub)[f[2]:l[2], f[1]:l[1], f[0]:l[0]].astype(np.float64)
                flat = vals.ravel()
                part.shift = float(np.float32(flat[0]))
                dd = flat - part.shift
                part.sum_d, part.sum_d2, part.count = float(dd.sum()), float((dd * dd).sum()), flat.size
                m, e = np.frexp(np.prod(flat[:8])) if flat.size <= 8 else (0.0, 0)
                part.prod_mant, part.prod_exp = float(m), int(e)
                nx, ny = l[0] - f[0], l[1] - f[1]
                for name, pos in (("min", int(np.argmin(flat))), ("max", int(np.argmax(flat)))):
                    setattr(part, name, float(flat[pos]))
                    setattr(part, "has_" + name, 1)
                    getattr(part, "arg" + name)[:] = [f[0] + pos % nx, f[1] + (pos // nx) % ny, f[2] + pos // (nx * ny)]

            got = sharded.aggregates_sharded(av, first, last, dist, compute=None if gpu else cpu_partial)
            got = po.Aggregates.from_buffer_copy(bytes(got))
            try:
                if gpu:
                    cases.check_aggregates(o, got, ref_agg, whole, first, last, (afmt, first))
                else:
                    assert cases.agg_exact_words(got).tobytes() == cases.agg_exact_words(ref_agg).tobytes()
                    assert abs(got.sum - ref_agg.sum) <= 1e-4 * abs(ref_agg.sum) and abs(got.var - ref_agg.var) <= 1e-4 * ref_agg.var
            except AssertionError as e:
                print(f"rank {rank}: MISMATCH in aggregates {afmt} {first} {last}: {e}", flush=True)
                sys.exit(1)

    # elementwise ops need no exchange: slabs tile the volume exactly
    Z = 23
    covered = sorted(z for r in range(world) for z in range(*sharded.slab_range(Z, world, r)))
    assert covered == list(range(Z))

    t = torch.tensor([total_halo], dtype=torch.int64, device=device)
    dist.all_reduce(t)
    if rank == 0:
        print(f"ok world={world} backend={backend} halo_bytes_total={int(t.item())}", flush=True)
        assert int(t.item()) > 0 or world == 1
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
