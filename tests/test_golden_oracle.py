"""CPU suite: the C restatement (oracle/vkt_oracle.c) must reproduce every frozen output of the
reference build (tests/golden/golden.json).  Runs anywhere (no reference sources, no GPU)."""
import hashlib
import json
import os

import numpy as np
import pytest

import cases
from oracle import pyoracle as po

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
ALL = cases.cases()


def test_catalogue_matches_golden_file():
    assert sorted(c[0] for c in ALL) == sorted(GOLD["sha256"].keys())


@pytest.mark.parametrize("kind", sorted({c[1] for c in ALL}))
def test_oracle_reproduces_golden(kind):
    o = po.oracle()
    for case in ALL:
        if case[1] != kind:
            continue
        got = [hashlib.sha256(a.tobytes()).hexdigest() for a in cases.run_case(o, case)]
        assert got == GOLD["sha256"][case[0]], case[0]


def test_known_answers():
    o = po.oracle()
    kat = GOLD["kat"]
    names = {"UInt8": po.UINT8, "UInt16": po.UINT16, "UInt32": po.UINT32}
    for fname, table in kat["fill"].items():
        for v, raw in table.items():
            assert int.from_bytes(o.map_voxel(float(v), names[fname], 0.0, 1.0), "little") == raw
    # the mapping WRAPS, it does not saturate
    assert kat["fill"]["UInt16"]["1.0"] == 0 and kat["fill"]["UInt16"]["1.5"] == 32768
    assert kat["fill"]["UInt8"]["1.5"] == 127 and kat["fill"]["UInt32"]["1.0"] == 0
    a = np.array(kat["arith_u16"]["a"], np.uint16).reshape(1, 1, 4)
    b = np.array(kat["arith_u16"]["b"], np.uint16).reshape(1, 1, 4)
    for op, name in enumerate(po.ARITH_NAMES):
        d = po.empty((4, 1, 1), po.UINT16)
        o.arith_range(op, d, po.Vol(a, po.UINT16), po.Vol(b, po.UINT16), (0, 0, 0), (4, 1, 1))
        assert [int(x) for x in d.a.ravel()] == kat["arith_u16"][name], name
    assert kat["arith_u16"]["SafeSum"] == [300, 0, 0, 65535]
    for fname, table in kat["unmap"].items():
        n = np.dtype(po.DTYPES[names[fname]]).itemsize
        for r, hexbits in table.items():
            v = o.unmap_voxel(int(r).to_bytes(n, "little"), names[fname], -1000.0, 3000.0)
            assert np.float32(v).tobytes().hex() == hexbits


def test_synth_numpy_twin():
    for fmt in cases.FMT4:
        a = po.synth(1000, fmt, 7, first_voxel=123)
        b = po.synth_numpy(1000, fmt, 7, first_voxel=123)
        assert a.tobytes() == b.tobytes()


def test_u8_identity_round_trip():
    """map(unmap(v)) == v for every UInt8 and UInt16 value under the default mapping: the one-tap Resample kernels move
    raw voxels for these formats instead of running them through the codec."""
    libs = [po.oracle()] + ([po.reference()] if po.have_reference() else [])
    for lib in libs:
        for v in range(256):
            raw = bytes([v])
            assert lib.map_voxel(lib.unmap_voxel(raw, po.UINT8), po.UINT8) == raw, v
        for v in list(range(0, 65536, 257)) + [1, 2, 65534, 65535, 32767, 32768]:
            raw = int(v).to_bytes(2, "little")
            assert lib.map_voxel(lib.unmap_voxel(raw, po.UINT16), po.UINT16) == raw, v
