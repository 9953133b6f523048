"""Several GPUs behind the drop-in C-ABI in ONE process: vktb_set_devices + striped (VMM) volumes + fan-out.

On a one-GPU box the group names the same device several times ({0, 0, 0}): a volume is then striped in three parts
on one GPU, each part with its own stream -- the whole placement / fork-join / per-stripe-copy logic runs, only the
peer loads stay local.  With two or more GPUs visible the same tests also run on real devices (gpurun --gpus 2).
Every result is compared with the oracle; volume shapes are chosen so that plane boundaries do NOT coincide with the
2 MiB stripe granularity (planes straddle stripes)."""
import ctypes
import os

import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


def groups():
    import torch
    n = torch.cuda.device_count()
    g = [[0, 0, 0]]
    if n >= 2:
        g.append(list(range(min(n, 4))))
    return g


@pytest.fixture()
def striped_env():
    from volkit_b200 import capi
    old = os.environ.get("VKTB_SHARD_MIN_MB")
    os.environ["VKTB_SHARD_MIN_MB"] = "1"          # volumes of 1 MiB and more are striped (read per allocation)
    lib = capi.load()
    yield lib, capi
    one = (ctypes.c_int * 1)(0)
    capi.check(lib.vktb_set_devices(one, 1))
    if old is None:
        os.environ.pop("VKTB_SHARD_MIN_MB", None)
    else:
        os.environ["VKTB_SHARD_MIN_MB"] = old


def set_group(lib, capi, ids):
    arr = (ctypes.c_int * len(ids))(*ids)
    capi.check(lib.vktb_set_devices(arr, len(ids)), "vktb_set_devices")
    n = ctypes.c_int()
    got = (ctypes.c_int * 8)()
    capi.check(lib.vktb_get_devices(got, 8, ctypes.byref(n)))
    assert list(got[:n.value]) == ids


def up(capi, a, fmt, lo=0.0, hi=1.0):
    from volkit_b200.device import DeviceVolume
    return DeviceVolume.from_numpy(a, fmt, lo, hi)


@pytest.mark.parametrize("gi", [0, 1])
def test_striped_volumes_fan_out_and_match_the_oracle(striped_env, gi):
    lib, capi = striped_env
    gs = groups()
    if gi >= len(gs):
        pytest.skip("needs two GPUs")
    ids = gs[gi]
    set_group(lib, capi, ids)
    o = po.oracle()
    I3 = capi.I3
    # 300 x 220 x 97 UInt16 = 12.8 MB: six 2 MiB granules, planes of 132000 bytes straddle every stripe boundary
    dims = (300, 220, 97)
    fmt, cf = po.UINT16, capi.UINT16
    a, b = po.synth(dims, fmt, 1), po.synth(dims, fmt, 2)
    da, db = up(capi, a, cf), up(capi, b, cf)
    dd = up(capi, po.synth(dims, fmt, 3), cf)
    st = ctypes.c_int()
    assert lib.vktb_is_striped(da.ptr, ctypes.byref(st)) == 1 and st.value == len(ids)
    assert da.download().tobytes() == a.tobytes()                       # striped H2D / D2H round trip
    launches = lib.vktb_launch_count()

    # arithmetic, whole volume
    want = po.empty(dims, fmt)
    o.arith_range(po.SAFESUM, want, po.Vol(a, fmt), po.Vol(b, fmt), (0, 0, 0), dims)
    capi.check(lib.vktb_arith_range(capi.SAFESUM, ctypes.byref(dd.desc()), ctypes.byref(da.desc()), ctypes.byref(db.desc()),
                                    I3(0, 0, 0), I3(*dims), I3(0, 0, 0)))
    assert lib.vktb_launch_count() - launches >= len(ids)              # one launch per stripe
    assert dd.download().tobytes() == want.a.tobytes()
    # *Range with a z offset: dst planes read source planes of the neighbouring stripe
    pre = po.synth(dims, fmt, 3)
    want = po.Vol(pre.copy(), fmt)
    first, last, off = (3, 5, 7), (290, 210, 80), (8, 1, 13)
    o.arith_range(po.DIFF, want, po.Vol(a, fmt), po.Vol(b, fmt), first, last, off)
    dd.upload(pre)
    capi.check(lib.vktb_arith_range(capi.DIFF, ctypes.byref(dd.desc()), ctypes.byref(da.desc()), ctypes.byref(db.desc()),
                                    I3(*first), I3(*last), I3(*off)))
    assert dd.download().tobytes() == want.a.tobytes()
    # in place (dst is a source, no offset)
    want = po.Vol(a.copy(), fmt)
    o.arith_range(po.SAFEPROD, want, want, po.Vol(b, fmt), (0, 0, 0), dims)
    tmp = up(capi, a, cf)
    capi.check(lib.vktb_arith_range(capi.SAFEPROD, ctypes.byref(tmp.desc()), ctypes.byref(tmp.desc()), ctypes.byref(db.desc()),
                                    I3(0, 0, 0), I3(*dims), I3(0, 0, 0)))
    assert tmp.download().tobytes() == want.a.tobytes()
    tmp.free()
    # CopyRange with clamped borders and a z shift
    want = po.Vol(pre.copy(), fmt)
    first, last, off = (-4, -3, -5), (250, 200, 70), (20, 10, 9)
    o.copy_range(want, po.Vol(a, fmt), first, last, off)
    dd.upload(pre)
    capi.check(lib.vktb_copy_range(ctypes.byref(dd.desc()), ctypes.byref(da.desc()), I3(*first), I3(*last), I3(*off)))
    assert dd.download().tobytes() == want.a.tobytes()
    # Fill range, Transform (absolute coordinates), binary Transform
    want = po.Vol(pre.copy(), fmt)
    o.fill_range(want, (1, 2, 3), (299, 219, 95), 0.3)
    o.transform_range1(want, (0, 0, 2), (300, 220, 96), 0)
    dd.upload(pre)
    capi.check(lib.vktb_fill_range(ctypes.byref(dd.desc()), I3(1, 2, 3), I3(299, 219, 95), 0.3))
    capi.check(lib.vktb_transform_range1(ctypes.byref(dd.desc()), I3(0, 0, 2), I3(300, 220, 96), 0))
    assert dd.download().tobytes() == want.a.tobytes()
    wa, wb = po.Vol(a.copy(), fmt), po.Vol(b.copy(), fmt)
    o.transform_range2(wa, wb, (2, 3, 4), (298, 217, 90), 0)
    ta, tb = up(capi, a, cf), up(capi, b, cf)
    capi.check(lib.vktb_transform_range2(ctypes.byref(ta.desc()), ctypes.byref(tb.desc()), I3(2, 3, 4), I3(298, 217, 90), 0))
    assert ta.download().tobytes() == wa.a.tobytes() and tb.download().tobytes() == wb.a.tobytes()
    ta.free(); tb.free()
    # Resample: up (taps cross stripe boundaries: the fused halo), down, true trilinear, Float32 8-tap
    for D, filt in (((450, 330, 145), capi.LINEAR), ((150, 110, 48), capi.NEAREST), ((400, 300, 130), capi.TRILINEAR)):
        want = po.empty(D, fmt)
        o.resample(want, po.Vol(a, fmt), filt)
        from volkit_b200.device import DeviceVolume
        dr = DeviceVolume(D, cf)
        capi.check(lib.vktb_resample(ctypes.byref(dr.desc()), ctypes.byref(da.desc()), filt))
        assert dr.download().tobytes() == want.a.tobytes(), (D, filt)
        dr.free()
    f = po.synth((160, 128, 60), po.FLOAT32, 5)
    df = up(capi, f, capi.FLOAT32)
    want = po.empty((240, 192, 90), po.FLOAT32)
    o.resample(want, po.Vol(f, po.FLOAT32), po.LINEAR)
    from volkit_b200.device import DeviceVolume
    dr = DeviceVolume((240, 192, 90), capi.FLOAT32)
    assert lib.vktb_is_striped(dr.ptr, None) == 1
    capi.check(lib.vktb_resample(ctypes.byref(dr.desc()), ctypes.byref(df.desc()), capi.LINEAR))
    assert dr.download().tobytes() == want.a.tobytes()
    # an entry point without a fan-out still works on striped memory (runs on the group's first device)
    bins = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(bins), 256 * 8))
    capi.check(lib.vktb_histogram_range(ctypes.byref(da.desc()), I3(0, 0, 0), I3(*dims), bins, 256, None))
    got = np.empty(256, np.uint64)
    capi.check(lib.vktb_memcpy(got.ctypes.data, bins, 2048, capi.D2H))
    ref_bins, _ = o.histogram_range(po.Vol(a, fmt), (0, 0, 0), dims, 256)
    assert got.tobytes() == ref_bins.tobytes()
    capi.check(lib.vktb_free(bins))
    # small volumes stay in the pool; mixed operands run on one device and are still right
    small = up(capi, po.synth((16, 16, 16), fmt, 1), cf)
    assert lib.vktb_is_striped(small.ptr, None) == 0
    small.free()
    for v in (da, db, dd, df, dr):
        v.free()
    # the group cannot change while striped volumes live
    keep = up(capi, a, cf)
    one = (ctypes.c_int * 1)(0)
    assert lib.vktb_set_devices(one, 1) != 0 and b"striped volumes are alive" in lib.vktb_last_error()
    keep.free()


@pytest.mark.parametrize("gi", [0, 1])
def test_histogram_of_a_striped_volume(striped_env, gi):
    """vktb_histogram_range on a striped volume: every device counts the planes that start in its stripe into its own
    bins, the parts are added on the calling device.  Whole volume and a sub-range whose first / last plane fall inside
    stripes, identity and non-identity mapping (voxels outside the bins are dropped and counted), against the oracle."""
    lib, capi = striped_env
    gs = groups()
    if gi >= len(gs):
        pytest.skip("needs two GPUs")
    set_group(lib, capi, gs[gi])
    o = po.oracle()
    I3 = capi.I3
    for fmt, cf, dims in ((po.UINT16, capi.UINT16, (300, 220, 97)), (po.UINT8, capi.UINT8, (512, 300, 61))):
        a = po.synth(dims, fmt, 4)
        for lo, hi, nb in ((0.0, 1.0, 256), (0.25, 0.75, 100), (0.0, 1.0, 5000)):
            dv = up(capi, a, cf, lo, hi)
            st = ctypes.c_int()
            assert lib.vktb_is_striped(dv.ptr, ctypes.byref(st)) == 1 and st.value == len(gs[gi])
            for first, last in (((0, 0, 0), dims), ((3, 2, 5), (dims[0] - 1, dims[1] - 4, dims[2] - 7))):
                want, wdrop = o.histogram_range(po.Vol(a, fmt, lo, hi), first, last, nb)
                bins = ctypes.c_void_p()
                capi.check(lib.vktb_alloc(ctypes.byref(bins), nb * 8))
                dropped = ctypes.c_uint64(123)
                before = lib.vktb_launch_count()
                capi.check(lib.vktb_histogram_range(ctypes.byref(dv.desc()), I3(*first), I3(*last), bins, nb, ctypes.byref(dropped)),
                           "vktb_histogram_range (striped)")
                assert lib.vktb_launch_count() - before >= 3      # at least two stripes counted, and the combine
                got = np.zeros(nb, np.uint64)
                capi.check(lib.vktb_memcpy(got.ctypes.data, bins, nb * 8, capi.D2H))
                capi.check(lib.vktb_free(bins))
                assert got.tobytes() == want.tobytes() and int(dropped.value) == wdrop, (fmt, lo, hi, nb, first)
            dv.free()
