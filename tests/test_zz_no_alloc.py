"""vktb_copy_range_batch_noalloc (no counterpart in the reference): the
batched CopyRange with its box table in a file-scope device array instead of a per-call allocation.

Collected LAST on purpose.  Checks, through the C-ABI, that two calls in a row --
which share the static scratch array -- give byte for byte the bricks of the existing vktb_copy_range_batch, that the
device-memory pool does not grow across the calls, and that a batch beyond the stated maximum is refused.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bricks(dims, brick, halo):
    """BrickDecompose layout: (firsts, lasts, brick dims) with `halo` voxels on every side (clamped CopyRange)."""
    firsts, lasts, bdims = [], [], []
    nb = [(d + brick - 1) // brick for d in dims]
    for z in range(nb[2]):
        for y in range(nb[1]):
            for x in range(nb[0]):
                f = [i * brick - halo for i in (x, y, z)]
                l = [min((i + 1) * brick, d) + halo for i, d in zip((x, y, z), dims)]
                firsts.append(f); lasts.append(l); bdims.append([b - a for a, b in zip(f, l)])
    return firsts, lasts, bdims


def _run(entry, lib, capi, src, fmt, firsts, lasts, bdims, prefill):
    """one batched call into freshly pre-filled bricks; returns the bricks' bytes"""
    from volkit_b200.device import DeviceVolume
    n = len(firsts)
    bricks = [DeviceVolume.from_numpy(np.full((d[2], d[1], d[0]), prefill, src.dtype), fmt) for d in bdims]
    s = DeviceVolume.from_numpy(src, fmt)
    descs = (capi.VktbVol * n)(*[b.desc() for b in bricks])
    flat = lambda t: (ctypes.c_int32 * (3 * n))(*[c for v in t for c in v])  # noqa: E731
    capi.check(entry(n, descs, ctypes.byref(s.desc()), flat(firsts), flat(lasts), flat([[0, 0, 0]] * n)), "batch")
    out = [b.download().tobytes() for b in bricks]
    for b in bricks:
        b.free()
    s.free()
    return out


def test_copy_range_batch_noalloc_matches_the_allocating_entry_point():
    from oracle import pyoracle as po
    from volkit_b200 import capi
    lib = capi.load()
    capi.check(lib.vktb_set_device(0), "vktb_set_device")
    assert lib.vktb_copy_range_batch_noalloc_max() == 4096

    # two calls in a row on different volumes, formats and brick counts: the second call finds the first one's table
    # in the shared scratch array (more records than it uses itself) and must not be affected by it
    cases = [((70, 45, 38), capi.UINT16, po.UINT16, 16, 1),     # 45 bricks, halos clamp at the borders, odd row lengths
             ((64, 32, 32), capi.UINT8, po.UINT8, 32, 0)]       # 2 bricks, no halo
    for round_no in range(2):
        for dims, fmt, pfmt, brick, halo in cases:
            src = po.synth(dims, pfmt, 11 + round_no)
            firsts, lasts, bdims = _bricks(dims, brick, halo)
            want = _run(lib.vktb_copy_range_batch, lib, capi, src, fmt, firsts, lasts, bdims, 7)
            got = _run(lib.vktb_copy_range_batch_noalloc, lib, capi, src, fmt, firsts, lasts, bdims, 7)
            assert got == want, f"noalloc batch differs from vktb_copy_range_batch ({dims}, round {round_no})"
            # and both equal the oracle's BrickDecompose (bit-exact: raw voxel moves)
            ref = po.oracle().brick_decompose(po.Vol(src, pfmt), (brick,) * 3, (halo,) * 3, (halo,) * 3)
            assert got == [r.tobytes() for r in ref]


def test_copy_range_batch_noalloc_does_not_allocate_and_rejects_oversize():
    from oracle import pyoracle as po
    from volkit_b200 import capi
    from volkit_b200.device import DeviceVolume
    lib = capi.load()
    capi.check(lib.vktb_set_device(0), "vktb_set_device")
    dims = (48, 48, 48)
    src = DeviceVolume.from_numpy(po.synth(dims, po.UINT16, 3), capi.UINT16)
    firsts, lasts, bdims = _bricks(dims, 16, 1)
    n = len(firsts)
    bricks = [DeviceVolume(d, capi.UINT16) for d in bdims]
    descs = (capi.VktbVol * n)(*[b.desc() for b in bricks])
    flat = lambda t: (ctypes.c_int32 * (3 * n))(*[c for v in t for c in v])  # noqa: E731
    f, l, o = flat(firsts), flat(lasts), flat([[0, 0, 0]] * n)
    capi.check(lib.vktb_sync())

    def stats():
        a, b = ctypes.c_size_t(), ctypes.c_size_t()
        capi.check(lib.vktb_mem_stats(ctypes.byref(a), ctypes.byref(b), None, None))
        return a.value, b.value

    before = stats()
    for _ in range(3):
        capi.check(lib.vktb_copy_range_batch_noalloc(n, descs, ctypes.byref(src.desc()), f, l, o), "noalloc batch")
    capi.check(lib.vktb_sync())
    assert stats() == before, "device memory in use / reserved changed across no-alloc calls"

    # beyond the static table: refused with a message, nothing launched
    big = 4097
    many = (capi.VktbVol * big)(*([bricks[0].desc()] * big))
    zeros = (ctypes.c_int32 * (3 * big))()
    launches = lib.vktb_launch_count()
    rc = lib.vktb_copy_range_batch_noalloc(big, many, ctypes.byref(src.desc()), zeros, zeros, zeros)
    assert rc != 0 and b"4096" in lib.vktb_last_error()
    assert lib.vktb_launch_count() == launches
    # a converting box is refused too (it would need vktb_copy_range_batch)
    other = DeviceVolume(bdims[0], capi.UINT8)
    one = (capi.VktbVol * 1)(other.desc())
    rc = lib.vktb_copy_range_batch_noalloc(1, one, ctypes.byref(src.desc()), (ctypes.c_int32 * 3)(*firsts[0]),
                                           (ctypes.c_int32 * 3)(*lasts[0]), (ctypes.c_int32 * 3)(0, 0, 0))
    assert rc != 0 and b"format" in lib.vktb_last_error()
    for b in bricks + [other]:
        b.free()
    src.free()
