"""The z-window forms of the C restatement (used by bench.py to check full-size volumes plane by plane) against their
whole-volume twins: a window of planes must reproduce exactly the planes of the whole-volume result."""
import numpy as np
import pytest

from oracle import pyoracle as po

FORMATS = [po.UINT8, po.UINT16, po.FLOAT32]


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("op", [0, 1, 2, 3])
def test_transform_window(fmt, op):
    o = po.oracle()
    dims = (19, 11, 23)
    whole = po.Vol(po.synth(dims, fmt, 5), fmt, -1.0, 2.0)
    ref = whole.copy()
    first, last = (2, 1, 0), (17, 10, 23)
    o.transform_range1(ref, first, last, op)
    for z0, z1 in ((0, 1), (7, 10), (22, 23)):
        win = po.Vol(whole.a[z0:z1].copy(), fmt, -1.0, 2.0)
        o.transform_range1_window(win, (first[0], first[1], 0), (last[0], last[1], z1 - z0), op, z0)
        assert win.a.tobytes() == ref.a[z0:z1].tobytes()


@pytest.mark.parametrize("fmt", FORMATS)
def test_affine_window(fmt):
    o = po.oracle()
    dims = (20, 14, 18)
    src = po.Vol(po.synth(dims, fmt, 1), fmt)
    for m in (o.scale_matrix((1.5, 1.5, 1.5)), o.rotate_matrix((0.0, 0.0, 1.0), 0.5), o.rotate_matrix((1.0, 2.0, 0.5), 0.8)):
        c = (dims[0] / 2, dims[1] / 2, dims[2] / 2)
        pre = po.synth(dims, fmt, 3)
        ref = po.Vol(pre.copy(), fmt)
        o.affine_range(ref, src, (0, 0, 0), dims, m, c)
        for z0, z1 in ((0, 2), (8, 9), (17, 18)):
            win = po.Vol(pre[z0:z1].copy(), fmt)
            o.affine_window(win, z0, src, 0, dims, (0, 0, 0), dims, m, c)
            assert win.a.tobytes() == ref.a[z0:z1].tobytes()
        # a genuinely partial src window for the z-preserving matrices (src plane = dst plane, +1 tap)
        if m[2] == 0.0 and m[5] == 0.0 and m[8] == 1.0:
            z = 9
            win = po.Vol(pre[z:z + 1].copy(), fmt)
            sub = po.Vol(src.a[z:z + 2].copy(), fmt)
            o.affine_window(win, z, sub, z, dims, (0, 0, 0), dims, m, c)
            assert win.a.tobytes() == ref.a[z:z + 1].tobytes()


@pytest.mark.parametrize("fmt", FORMATS)
def test_filter_window(fmt):
    o = po.oracle()
    dims = (17, 13, 12)
    src = po.Vol(po.synth(dims, fmt, 1), fmt)
    w = (np.arange(27, dtype=np.float32).reshape(3, 3, 3) - 9.0) / np.float32(27.0)
    ref = po.empty(dims, fmt)
    o.apply_filter(ref, src, w)
    for z0, z1 in ((0, 1), (1, 2), (5, 8), (10, 11), (11, 12)):
        s0, s1 = max(z0 - 1, 0), min(z1 + 1, dims[2])
        win = po.empty((dims[0], dims[1], z1 - z0), fmt)
        sub = po.Vol(src.a[s0:s1].copy(), fmt)
        o.apply_filter_window(win, z0, sub, s0, dims, w)
        assert win.a.tobytes() == ref.a[z0:z1].tobytes()


def test_synth_index_form():
    idx = np.array([[0, 5, 17], [1 << 33, 123456789, 7]], dtype=np.uint64)
    for fmt in (po.UINT8, po.UINT16, po.UINT32, po.FLOAT32):
        got = po.synth_numpy(0, fmt, 4, index=idx)
        for (i, j), v in np.ndenumerate(idx):
            assert got[i, j] == po.synth_numpy(1, fmt, 4, int(v))[0]
