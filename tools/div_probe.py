#!/usr/bin/env python
"""Which mapping ranges pass the exhaustive proof of the reciprocal division (vktb_debug_div_proven), and how long a
proof takes.  usage: div_probe.py"""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from volkit_b200 import capi  # noqa: E402

lib = capi.load()
capi.check(lib.vktb_set_device(0))
for lo, hi in ((-1000.0, 3000.0), (0.0, 3.0), (0.1, 0.8), (1e-3, 1e3), (2.0, -5.0), (0.0, 255.0), (0.0, 7.0), (-1.0, 2.0),
               (0.0, 1.0000001), (0.0, 1e30), (0.0, 1e-30)):
    p = ctypes.c_int(-1)
    t = time.time()
    capi.check(lib.vktb_debug_div_proven(lo, hi, ctypes.byref(p)))
    print(f"range {hi - lo!r:>24}: proven={p.value}  ({(time.time() - t) * 1e3:.1f} ms)", flush=True)
