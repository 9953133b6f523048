"""z-slab sharding of StructuredVolumes across the GPUs of one box: one process per GPU (torch.distributed).

The reference has no multi-device code; its only decomposition primitive, BrickDecompose
(reference src/vkt/Decompose_serial.hpp:15-48: bricks with negative/positive halos filled by clamped CopyRange), is the
model: rank g owns planes [g*Z/n, (g+1)*Z/n) of a volume (contiguous bytes, z is the slowest axis) plus, for
Resample only, the few halo planes its dst slab taps outside that range.

* Fill / Transform / Arithmetic (dstOffset 0) / Copy of congruent volumes: every rank runs the ordinary C-ABI call on
  its slab -- no communication.
* Resample: dst is sharded by dst planes; the src planes a dst slab reads come from the planner
  vktb_resample_window_src_range (the reference's own float expression, monotone in z); planes a rank does not own
  are fetched from their owners with point-to-point sends over NCCL (NVLink) before the kernel runs on the window.

All functions are pure host logic over torch tensors, so the same code runs on CPU tensors with the gloo backend
(tests) and on CUDA tensors with NCCL.  The voxel work itself is always the C-ABI (`compute=` is only overridden by
the CPU tests, which plug the oracle in to check the sharding logic without a GPU).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

from . import capi


class bound_stream:
    """Orders the C-ABI kernels with the torch work around them (NCCL receives, copy_, zeros, all_reduce).

    The backend launches on its own non-blocking stream unless told otherwise, and `req.wait()` / `copy_` only order
    torch's CURRENT stream, so a kernel could start before its halo planes arrived or race the collective that
    follows.  Every sharded op therefore runs its C-ABI calls inside `with bound_stream(torch, tensor)`: the backend's
    compute stream of this thread is torch's current stream for the duration (handle 0, the legacy default stream,
    is passed as cudaStreamLegacy = 1, because 0 means "backend-owned" to vktb_set_stream).  CPU tensors (the gloo
    tests, which plug the oracle in) need nothing."""

    def __init__(self, torch, tensor):
        self.active = bool(getattr(tensor, "is_cuda", False))
        self.torch = torch

    def __enter__(self):
        if self.active:
            lib = capi.load()
            self.prev = lib.vktb_get_stream()
            h = self.torch.cuda.current_stream().cuda_stream
            capi.check(lib.vktb_set_stream(h if h else 1), "vktb_set_stream")
        return self

    def __exit__(self, *exc):
        if self.active:
            capi.load().vktb_set_stream(self.prev)
        return False


def slab_range(dim_z: int, world: int, rank: int) -> tuple[int, int]:
    """Planes [z0, z1) owned by `rank`: floor(g*Z/n) .. floor((g+1)*Z/n)."""
    return rank * dim_z // world, (rank + 1) * dim_z // world


def _src_coord(z: int, D: int, S: int) -> int:
    """(int)(z / float(D) * S) in binary32, reference src/vkt/Resample_serial.hpp:60-62 (host twin of the planner)."""
    import numpy as np
    q = np.float32(z) / np.float32(D)
    s = np.float32(q * np.float32(S))
    return int(s)


def resample_src_range(filt: int, dz0: int, dz1: int, D, S, lib=None) -> tuple[int, int]:
    """src planes [s0, s1) the dst slab [dz0, dz1) reads (D, S = full dims); the C-ABI planner when the library is
    given, else its pure-Python twin."""
    if dz1 <= dz0:
        return 0, 0
    if lib is not None:
        a, b = ctypes.c_int32(), ctypes.c_int32()
        capi.check(lib.vktb_resample_window_src_range3(filt, dz0, dz1, capi.I3(*D), capi.I3(*S), ctypes.byref(a),
                                                       ctypes.byref(b)), "vktb_resample_window_src_range3")
        return a.value, b.value
    lo = min(max(_src_coord(dz0, D[2], S[2]), 0), S[2] - 1)
    hi = min(max(_src_coord(dz1 - 1, D[2], S[2]), 0), S[2] - 1)
    end = hi + 1 if filt == capi.NEAREST else hi + 2
    # reference-faithful Linear: the unclamped +1 x tap leaves its plane only from the last column of the last row
    if filt == capi.LINEAR and _src_coord(D[0] - 1, D[0], S[0]) >= S[0] - 1 \
            and _src_coord(D[1] - 1, D[1], S[1]) + 1 >= S[1] - 1:
        end = hi + 3
    return lo, min(end, S[2])


@dataclass
class Transfer:
    src_rank: int
    dst_rank: int
    z0: int      # global plane range [z0, z1) of the src volume
    z1: int


def halo_plan(filt: int, D: tuple, S: tuple, world: int, lib=None) -> tuple[list, list]:
    """For every rank the src window it needs and the list of plane transfers that fill the parts it does not own.
    Deterministic: every rank computes the same plan.  Equal dims need no halo (conversion only)."""
    windows, transfers = [], []
    same = tuple(D) == tuple(S)
    for r in range(world):
        dz0, dz1 = slab_range(D[2], world, r)
        if same:
            need = (dz0, dz1)
        else:
            need = resample_src_range(filt, dz0, dz1, D, S, lib)
        windows.append(need)
        for o in range(world):
            if o == r:
                continue
            oz0, oz1 = slab_range(S[2], world, o)
            a, b = max(need[0], oz0), min(need[1], oz1)
            if a < b:
                transfers.append(Transfer(o, r, a, b))
    return windows, transfers


class ShardedVolume:
    """This rank's slab of a (X, Y, Z) volume as one flat torch.uint8 tensor (CPU or CUDA).  `halo = (below, above)`
    reserves room for that many foreign planes directly before / after the slab, so that a Resample window is
    assembled in place (halo_for() sizes it from a plan)."""

    def __init__(self, torch, dims, fmt, rank, world, lo=0.0, hi=1.0, device="cuda", halo=(0, 0)):
        self.torch, self.dims, self.fmt, self.rank, self.world = torch, tuple(dims), fmt, rank, world
        self.lo, self.hi = float(lo), float(hi)
        self.bpv = capi.BYTES_PER_VOXEL[fmt]
        self.plane_bytes = self.dims[0] * self.dims[1] * self.bpv
        self.z0, self.z1 = slab_range(self.dims[2], world, rank)
        self.halo = (max(0, min(halo[0], self.z0)), max(0, min(halo[1], self.dims[2] - self.z1)))
        n = (self.z1 - self.z0 + self.halo[0] + self.halo[1]) * self.plane_bytes
        self.storage = torch.empty(max(n, 1), dtype=torch.uint8, device=device)[:n]
        self.data = self.planes(self.z0, self.z1)   # the owned slab

    def planes(self, z0, z1):
        """view of the global planes [z0, z1) (owned or halo)"""
        base = self.z0 - self.halo[0]
        assert base <= z0 <= z1 <= self.z1 + self.halo[1], (z0, z1, self.z0, self.z1, self.halo)
        return self.storage[(z0 - base) * self.plane_bytes:(z1 - base) * self.plane_bytes]

    def holds(self, z0, z1):
        return self.z0 - self.halo[0] <= z0 and z1 <= self.z1 + self.halo[1]

    def desc(self, z0=None, z1=None):
        """vktb_vol of the owned slab, or of the global planes [z0, z1)"""
        z0 = self.z0 if z0 is None else z0
        z1 = self.z1 if z1 is None else z1
        return capi.vol(self.planes(z0, z1).data_ptr(), (self.dims[0], self.dims[1], z1 - z0), self.fmt, self.lo, self.hi)


def halo_for(plan, src_dims, world, rank) -> tuple[int, int]:
    """(planes below, planes above) the src slab of `rank` must reserve to assemble its window in place."""
    w0, w1 = plan[0][rank]
    z0, z1 = slab_range(src_dims[2], world, rank)
    if w1 <= w0:
        return 0, 0
    return max(0, z0 - w0), max(0, w1 - z1)


def start_gather(src: ShardedVolume, window, transfers, dist):
    """Posts the plane transfers that complete the src planes `window` of this rank and returns (buf, wait): `buf`
    holds the window -- in place when the slab was allocated with enough halo room (foreign planes are received
    straight into it), else a fresh buffer (owned planes copied) -- and `wait()` blocks the CURRENT STREAM (not the
    host, on CUDA) until the foreign planes have arrived.  Work launched between the two calls that only reads owned
    planes overlaps the exchange.

    NCCL moves CUDA tensors peer to peer over NVLink.  With the gloo backend CUDA planes are staged through host
    memory (the two-ranks-on-one-GPU test set-up of tests/test_sharded_gpu.py; NCCL refuses two ranks per device).
    VKTB_TEST_BREAK_HALO=1 (tests only) receives the first incoming transfer into a scratch buffer, i.e. loses it."""
    import os
    torch = src.torch
    w0, w1 = window
    mine = [t for t in transfers if t.dst_rank == src.rank]
    outgoing = [t for t in transfers if t.src_rank == src.rank]
    if w1 <= w0:
        buf, base = src.storage[:0], w0
    elif src.holds(w0, w1):
        buf, base = src.planes(w0, w1), w0
    else:
        buf = torch.empty((w1 - w0) * src.plane_bytes, dtype=torch.uint8, device=src.storage.device)
        base = w0
        a, b = max(w0, src.z0), min(w1, src.z1)
        if a < b:
            buf[(a - w0) * src.plane_bytes:(b - w0) * src.plane_bytes].copy_(src.planes(a, b))
    staged = src.storage.is_cuda and (outgoing or mine) and dist.get_backend() == "gloo"
    broken = os.environ.get("VKTB_TEST_BREAK_HALO") == "1"
    ops, landed = [], []
    for t in outgoing:
        view = src.planes(t.z0, t.z1)
        ops.append(dist.P2POp(dist.isend, view.cpu() if staged else view, t.dst_rank))
    for i, t in enumerate(mine):
        view = buf[(t.z0 - base) * src.plane_bytes:(t.z1 - base) * src.plane_bytes]
        if broken and i == 0:
            view = torch.empty_like(view)
        if staged:
            tmp = torch.empty(view.numel(), dtype=torch.uint8)
            landed.append((view, tmp))
            view = tmp
        ops.append(dist.P2POp(dist.irecv, view, t.src_rank))
    reqs = dist.batch_isend_irecv(ops) if ops else []

    def wait():
        for req in reqs:
            req.wait()
        for view, tmp in landed:
            view.copy_(tmp)
    return buf, wait


def gather_window(src: ShardedVolume, window, transfers, dist):
    """start_gather + wait: the tensor holding the complete window."""
    buf, wait = start_gather(src, window, transfers, dist)
    wait()
    return buf


def _cabi_resample_window(dst_desc, src_desc, filt, dz0, D, sz0, S):
    lib = capi.load()
    capi.check(lib.vktb_resample_window(ctypes.byref(dst_desc), ctypes.byref(src_desc), filt, dz0, capi.I3(*D), sz0,
                                        capi.I3(*S)), "vktb_resample_window")


_interior_cache: dict = {}


def interior_planes(filt, dz0, dz1, D, S, own0, own1, lib=None):
    """The dst planes [i0, i1) of [dz0, dz1) whose taps all lie in the src planes [own0, own1) this rank owns.  The
    coordinate map is monotone in z, so they are contiguous: the ends are walked inwards plane by plane (a halo is a
    few planes thick)."""
    key = (filt, dz0, dz1, tuple(D), tuple(S), own0, own1)
    hit = _interior_cache.get(key)
    if hit is not None:
        return hit
    same = tuple(D) == tuple(S)

    def owned(z):
        a, b = (z, z + 1) if same else resample_src_range(filt, z, z + 1, D, S, lib)
        return own0 <= a and b <= own1
    i0 = dz0
    while i0 < dz1 and not owned(i0):
        i0 += 1
    i1 = dz1
    while i1 > i0 and not owned(i1 - 1):
        i1 -= 1
    _interior_cache[key] = (i0, i1)
    return i0, i1


def resample_sharded(dst: ShardedVolume, src: ShardedVolume, filt: int, dist, compute=None, planner_lib=None,
                     plan=None, overlap=True):
    """dst slab <- Resample(src) with halo exchange.  The dst planes whose taps this rank owns (all but a few at the
    slab ends) are launched first and run WHILE the halo planes travel; the boundary planes follow once they have
    arrived.  Returns the number of halo bytes this rank received."""
    D, S = dst.dims, src.dims
    windows, transfers = plan if plan is not None else halo_plan(filt, D, S, dst.world, planner_lib)
    window = windows[dst.rank]
    run = compute or _cabi_resample_window
    incoming = [t for t in transfers if t.dst_rank == dst.rank]
    received = sum((t.z1 - t.z0) * src.plane_bytes for t in incoming)
    with bound_stream(dst.torch, dst.storage):
        buf, wait = start_gather(src, window, transfers, dist)
        pieces = [(dst.z0, dst.z1)]
        if incoming and overlap and dst.z1 > dst.z0:
            i0, i1 = interior_planes(filt, dst.z0, dst.z1, D, S, src.z0, src.z1, planner_lib)
            if i1 > i0:
                same = tuple(D) == tuple(S)
                need = (i0, i1) if same else resample_src_range(filt, i0, i1, D, S, planner_lib)
                run(dst.desc(i0, i1), src.desc(need[0], need[1]), filt, i0, D, need[0], S)
                pieces = [(dst.z0, i0), (i1, dst.z1)]
        wait()
        sdesc = capi.vol(buf.data_ptr(), (S[0], S[1], window[1] - window[0]), src.fmt, src.lo, src.hi)
        for a, b in pieces:
            if b > a:
                run(dst.desc(a, b), sdesc, filt, a, D, window[0], S)
    return received


# ---- *Range calls whose dst planes come from other src planes (z offset, clamped Copy) ------------------------------
#
# Reference semantics: arithmetic writes dst[p + off] = f(s1[p], s2[p]) for absolute p in
# [first, last) (Arithmetic_serial.hpp:15-45); Copy writes dst[p - first + off] = src[clamp(p, 0, dims-1)]
# (Copy_serial.hpp:13-82).  Sharded by DST slab: a rank runs the sub-range of p.z whose dst planes it owns on a window of
# the src planes that sub-range reads; planes it does not own arrive by the same point-to-point transfers Resample uses.

def range_plan(kind: str, first_z: int, last_z: int, off_z: int, dst_dimz: int, src_dimz: int, world: int):
    """(windows, subranges, transfers): per rank the src planes [w0, w1) it needs, the p.z sub-range [pz0, pz1) it
    computes, and the plane transfers.  kind = "arith" (dst plane = p.z + off_z, no clamp) or "copy"
    (dst plane = p.z - first_z + off_z, src plane clamped to [0, src_dimz - 1])."""
    assert kind in ("arith", "copy")
    shift = off_z if kind == "arith" else off_z - first_z
    windows, subranges, transfers = [], [], []
    for r in range(world):
        dz0, dz1 = slab_range(dst_dimz, world, r)
        pz0, pz1 = max(first_z, dz0 - shift), min(last_z, dz1 - shift)
        if pz1 <= pz0:
            windows.append((0, 0))
            subranges.append((0, 0))
            continue
        if kind == "arith":
            assert 0 <= pz0 and pz1 <= src_dimz, "arithmetic range outside the source volumes"
            need = (pz0, pz1)
        else:
            need = (min(max(pz0, 0), src_dimz - 1), min(max(pz1 - 1, 0), src_dimz - 1) + 1)
        windows.append(need)
        subranges.append((pz0, pz1))
        for o in range(world):
            if o == r:
                continue
            oz0, oz1 = slab_range(src_dimz, world, o)
            a, b = max(need[0], oz0), min(need[1], oz1)
            if a < b:
                transfers.append(Transfer(o, r, a, b))
    return windows, subranges, transfers


def _window_desc(vol: ShardedVolume, buf, window):
    return capi.vol(buf.data_ptr(), (vol.dims[0], vol.dims[1], window[1] - window[0]), vol.fmt, vol.lo, vol.hi)


def _cabi_copy_range(dst_desc, src_desc, first, last, off):
    lib = capi.load()
    capi.check(lib.vktb_copy_range(ctypes.byref(dst_desc), ctypes.byref(src_desc), capi.I3(*first), capi.I3(*last),
                                   capi.I3(*off)), "vktb_copy_range")


def _cabi_arith_range(op, dst_desc, a_desc, b_desc, first, last, off):
    lib = capi.load()
    capi.check(lib.vktb_arith_range(op, ctypes.byref(dst_desc), ctypes.byref(a_desc), ctypes.byref(b_desc),
                                    capi.I3(*first), capi.I3(*last), capi.I3(*off)), "vktb_arith_range")


def _split_owned(pz0, pz1, src_plane, own0, own1):
    """[pz0, pz1) cut into (before, interior, after): interior = the p.z whose src plane src_plane(p.z) (monotone) lies
    in the owned planes [own0, own1)."""
    i0 = pz0
    while i0 < pz1 and not own0 <= src_plane(i0) < own1:
        i0 += 1
    i1 = pz1
    while i1 > i0 and not own0 <= src_plane(i1 - 1) < own1:
        i1 -= 1
    return (pz0, i0), (i0, i1), (i1, pz1)


def copy_range_sharded(dst: ShardedVolume, src: ShardedVolume, first, last, off, dist, compute=None, overlap=True) -> int:
    """dst slab <- CopyRange(src, first, last, off) with the reference's clamped src indexing.  The part of the
    sub-range that reads owned src planes runs while the foreign planes travel.  Returns the halo bytes this rank
    received."""
    assert dst.storage.data_ptr() != src.storage.data_ptr(), "sharded CopyRange needs distinct volumes"
    windows, sub, transfers = range_plan("copy", first[2], last[2], off[2], dst.dims[2], src.dims[2], dst.world)
    window = windows[dst.rank]
    run = compute or _cabi_copy_range
    incoming = [t for t in transfers if t.dst_rank == dst.rank]
    with bound_stream(dst.torch, dst.storage):
        buf, wait = start_gather(src, window, transfers, dist)
        pz0, pz1 = sub[dst.rank]
        w0 = window[0]

        def piece(a, b, sdesc, base):
            # src z relative to the plane `base` the descriptor starts at (may be negative / past the end: clamped
            # to the descriptor's first / last plane, which is then the volume's)
            if b > a:
                run(dst.desc(), sdesc, (first[0], first[1], a - base), (last[0], last[1], b - base),
                    (off[0], off[1], a - first[2] + off[2] - dst.z0))

        clamp = lambda z: min(max(z, 0), src.dims[2] - 1)
        before, inner, after = (pz0, pz1), (pz1, pz1), (pz1, pz1)
        if incoming and overlap and pz1 > pz0:
            before, inner, after = _split_owned(pz0, pz1, clamp, src.z0, src.z1)
            if inner[1] > inner[0]:
                # owned planes [ia, ib) as a stand-alone volume: a p.z outside [0, dims) clamps to the volume's first /
                # last plane, which is then ia / ib - 1, i.e. the descriptor's own first / last plane
                ia, ib = clamp(inner[0]), clamp(inner[1] - 1) + 1
                piece(inner[0], inner[1], src.desc(ia, ib), ia)
        wait()
        for a, b in (before, after):
            piece(a, b, _window_desc(src, buf, window), w0)
    return sum((t.z1 - t.z0) * src.plane_bytes for t in incoming)


def arith_range_sharded(op: int, dst: ShardedVolume, s1: ShardedVolume, s2: ShardedVolume, first, last, off, dist,
                        compute=None, overlap=True) -> int:
    """dst slab <- op(s1, s2) over [first, last) written at + off.  With off[2] == 0 and congruent volumes no plane
    moves; otherwise the p.z whose source planes this rank owns run while the others' planes travel.  Returns the
    halo bytes this rank received."""
    assert s1.dims[2] == s2.dims[2], "sources of a sharded arithmetic call must have the same number of planes"
    assert dst.storage.data_ptr() not in (s1.storage.data_ptr(), s2.storage.data_ptr()) or off[2] == 0
    windows, sub, transfers = range_plan("arith", first[2], last[2], off[2], dst.dims[2], s1.dims[2], dst.world)
    window = windows[dst.rank]
    run = compute or _cabi_arith_range
    incoming = [t for t in transfers if t.dst_rank == dst.rank]
    with bound_stream(dst.torch, dst.storage):
        buf1, wait1 = start_gather(s1, window, transfers, dist)
        buf2, wait2 = (buf1, lambda: None) if s2 is s1 else start_gather(s2, window, transfers, dist)
        pz0, pz1 = sub[dst.rank]

        def piece(a, b, d1, d2, base):
            if b > a:
                run(op, dst.desc(), d1, d2, (first[0], first[1], a - base), (last[0], last[1], b - base),
                    (off[0], off[1], base + off[2] - dst.z0))

        before, inner, after = (pz0, pz1), (pz1, pz1), (pz1, pz1)
        if incoming and overlap and pz1 > pz0:
            before, inner, after = _split_owned(pz0, pz1, lambda z: z, s1.z0, s1.z1)
            piece(inner[0], inner[1], s1.desc(inner[0], inner[1]) if inner[1] > inner[0] else None,
                  s2.desc(inner[0], inner[1]) if inner[1] > inner[0] else None, inner[0])
        wait1()
        wait2()
        for a, b in (before, after):
            piece(a, b, _window_desc(s1, buf1, window), _window_desc(s2, buf2, window), window[0])
    recv = sum((t.z1 - t.z0) for t in incoming)
    return recv * (s1.plane_bytes + (0 if s2 is s1 else s2.plane_bytes))


# ---- Transform of a sharded volume: the op must see ABSOLUTE coordinates ----------------------------------------------

def _cabi_transform_range1(vol_desc, first, last, op):
    lib = capi.load()
    capi.check(lib.vktb_transform_range1(ctypes.byref(vol_desc), capi.I3(*first), capi.I3(*last), op),
               "vktb_transform_range1")


def transform_sharded(vol: ShardedVolume, first, last, op: int, compute=None):
    """TransformRange (reference src/vkt/Transform_serial.hpp:15-50) on this rank's planes of [first, last).  The op
    receives the voxel's coordinates in the WHOLE volume, so the slab is described with the full dims and a data
    pointer displaced by z0 planes (only the owned planes are ever dereferenced); no communication."""
    z0, z1 = max(first[2], vol.z0), min(last[2], vol.z1)
    if z1 <= z0 or last[0] <= first[0] or last[1] <= first[1]:
        return
    base = vol.planes(vol.z0, vol.z1).data_ptr() - vol.z0 * vol.plane_bytes
    desc = capi.vol(base, vol.dims, vol.fmt, vol.lo, vol.hi)
    with bound_stream(vol.torch, vol.storage):
        (compute or _cabi_transform_range1)(desc, (first[0], first[1], z0), (last[0], last[1], z1), op)


# ---- Histogram of a sharded volume: the one real exchange step of the path -- a SUM all-reduce of the bins -------------

def _cabi_histogram_range(vol_desc, first, last, bins_ptr, num_bins):
    lib = capi.load()
    capi.check(lib.vktb_histogram_range(ctypes.byref(vol_desc), capi.I3(*first), capi.I3(*last), bins_ptr, num_bins, None),
               "vktb_histogram_range")


def histogram_sharded(vol: ShardedVolume, first, last, num_bins: int, dist, compute=None):
    """ComputeHistogramRange (reference src/vkt/Histogram_serial.hpp:18-51) of a z-slab sharded volume: every rank
    counts the planes of [first, last) it owns into its own 64-bit bins (int64 tensor on the volume's device), then
    one all-reduce (NCCL on GPUs, SUM) gives every rank the global histogram.  Returns the bins tensor."""
    torch = vol.torch
    bins = torch.zeros(num_bins, dtype=torch.int64, device=vol.storage.device)
    z0, z1 = max(first[2], vol.z0), min(last[2], vol.z1)
    if z1 > z0 and last[0] > first[0] and last[1] > first[1]:
        f = (first[0], first[1], z0 - vol.z0)
        l = (last[0], last[1], z1 - vol.z0)
        with bound_stream(torch, vol.storage):    # after the zero fill, before the all-reduce, on torch's stream
            (compute or _cabi_histogram_range)(vol.desc(), f, l, bins.data_ptr(), num_bins)
    if dist is not None and vol.world > 1:
        dist.all_reduce(bins, op=dist.ReduceOp.SUM)
    return bins


# ---- Aggregates of a sharded volume: one 88-byte partial per rank, all-gathered, folded in rank (= z) order ------------

def _cabi_aggregates_partial(vol_desc, first, last, part):
    lib = capi.load()
    capi.check(lib.vktb_aggregates_partial(ctypes.byref(vol_desc), capi.I3(*first), capi.I3(*last), ctypes.byref(part)),
               "vktb_aggregates_partial")


def aggregates_sharded(vol: ShardedVolume, first, last, dist, compute=None) -> "capi.VktbAggregates":
    """ComputeAggregatesRange (reference src/vkt/Aggregates_serial.hpp:20-83) of a z-slab sharded volume: every rank
    reduces the planes of [first, last) it owns to one vktb_agg_partial (moments, product, extrema with positions),
    the partials travel through ONE all-gather (88 bytes per rank) and every rank folds them in rank order with
    vktb_aggregates_finish (host arithmetic), so ties between slabs go to the lower z like the serial walk."""
    torch = vol.torch
    part = capi.VktbAggPartial()
    part.prod_mant, part.min, part.max = 1.0, 3.4028234663852886e38, -3.4028234663852886e38
    z0, z1 = max(first[2], vol.z0), min(last[2], vol.z1)
    if z1 > z0 and last[0] > first[0] and last[1] > first[1]:
        f = (first[0], first[1], z0 - vol.z0)
        l = (last[0], last[1], z1 - vol.z0)
        with bound_stream(torch, vol.storage):
            (compute or _cabi_aggregates_partial)(vol.desc(), f, l, part)
        part.argmin[2] += vol.z0 if part.has_min else 0          # slab -> whole-volume coordinates
        part.argmax[2] += vol.z0 if part.has_max else 0
    world = vol.world if dist is not None else 1
    raw = bytes(part)
    if world > 1:
        mine = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(vol.storage.device)
        everyone = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(everyone, mine)
        raw = b"".join(bytes(t.cpu().numpy().tobytes()) for t in everyone)
    parts = (capi.VktbAggPartial * world).from_buffer_copy(raw)
    out = capi.VktbAggregates()
    num_elems = vol.dims[0] * vol.dims[1] * vol.dims[2]
    capi.check(capi.load().vktb_aggregates_finish(parts, world, num_elems, ctypes.byref(out)), "vktb_aggregates_finish")
    return out
