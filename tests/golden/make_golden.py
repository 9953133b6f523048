"""Generates tests/golden/golden.json from the UNMODIFIED reference serial CPU path
(oracle/_ref/libvkt_ref.so, built by `make -C oracle ref` from /root/reference).

Run in the build container (the reference sources do not exist on the GPU box):
    python tests/golden/make_golden.py

The file holds, per case of tests/cases.py, the SHA-256 of every output buffer the reference
produced, plus a small table of literal known answers for readability.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import pyoracle as po  # noqa: E402
import cases  # noqa: E402


def main():
    ref = po.reference()
    digests = {}
    for case in cases.cases():
        outs = cases.run_case(ref, case)
        digests[case[0]] = [hashlib.sha256(o.tobytes()).hexdigest() for o in outs]

    # literal known answers
    kat = {"fill": {}, "arith_u16": {}, "unmap": {}}
    for fmt in (po.UINT8, po.UINT16, po.UINT32):
        kat["fill"][po.FORMAT_NAMES[fmt]] = {
            repr(v): int.from_bytes(ref.map_voxel(v, fmt, 0.0, 1.0), "little")
            for v in (0.0, 0.1, 0.5, 0.999, 1.0, 1.5, -0.25)}
    a = np.array([100, 40000, 65535, 30000], np.uint16).reshape(1, 1, 4)
    b = np.array([200, 40000, 1, 35535], np.uint16).reshape(1, 1, 4)
    for op in range(10):
        d = po.empty((4, 1, 1), po.UINT16)
        ref.arith_range(op, d, po.Vol(a, po.UINT16), po.Vol(b, po.UINT16), (0, 0, 0), (4, 1, 1))
        kat["arith_u16"][po.ARITH_NAMES[op]] = [int(x) for x in d.a.ravel()]
    kat["arith_u16"]["a"] = [int(x) for x in a.ravel()]
    kat["arith_u16"]["b"] = [int(x) for x in b.ravel()]
    for fmt, raws in ((po.UINT8, (0, 1, 127, 255)), (po.UINT16, (0, 1, 32768, 65535)), (po.UINT32, (0, 1, 2 ** 31, 2 ** 32 - 1))):
        n = np.dtype(po.DTYPES[fmt]).itemsize
        kat["unmap"][po.FORMAT_NAMES[fmt]] = {
            str(r): np.float32(ref.unmap_voxel(int(r).to_bytes(n, "little"), fmt, -1000.0, 3000.0)).tobytes().hex()
            for r in raws}

    out = {"generator": "tests/golden/make_golden.py", "source": "oracle/_ref/libvkt_ref.so (reference @ ca9f6ec, "
           "g++ -std=c++14 -O2 -ffp-contract=off, x86-64)", "kat": kat, "sha256": digests}
    path = os.path.join(ROOT, "tests", "golden", "golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(f"wrote {path}: {len(digests)} cases")


if __name__ == "__main__":
    main()
