#!/usr/bin/env python
"""Debug helper: runs vktb_resample on one shape and compares with the oracle.
usage: python tools/rs_check.py "Sx,Sy,Sz>Dx,Dy,Dz:fmt:filter" """
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np  # noqa: E402

import cases  # noqa: E402
from gpu_adapter import GpuLib  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

FMT = {"u8": po.UINT8, "u16": po.UINT16, "u32": po.UINT32, "f32": po.FLOAT32}
g = GpuLib()
o = po.oracle()
for spec in sys.argv[1:]:
    shape, fmt, filt = spec.split(":")
    s, d = shape.split(">")
    S, D = tuple(int(v) for v in s.split(",")), tuple(int(v) for v in d.split(","))
    src = cases.mk(S, FMT[fmt], 1)
    a, b = po.empty(D, FMT[fmt]), po.empty(D, FMT[fmt])
    o.resample(a, src, int(filt))
    g.resample(b, src, int(filt))
    bad = np.argwhere(a.a != b.a)
    print(spec, "OK" if len(bad) == 0 else f"MISMATCH {len(bad)} first {bad[:3].tolist()}", flush=True)
