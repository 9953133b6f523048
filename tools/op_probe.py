#!/usr/bin/env python
"""Times one C-ABI op on device-resident volumes.  usage: op_probe.py OP X Y Z FMT   (OP: filter | scale | rotate | agg | scan | sumrange | sumrange8 | copyrange)"""
import ctypes
import math
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

from volkit_b200 import capi  # noqa: E402
from volkit_b200.device import DeviceVolume  # noqa: E402

FMT = {"u8": capi.UINT8, "u16": capi.UINT16, "u32": capi.UINT32, "f32": capi.FLOAT32}


def main():
    op, dims, fmt = sys.argv[1], tuple(int(a) for a in sys.argv[2:5]), FMT[sys.argv[5]]
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    capi.check(lib.vktb_set_stream(stream.cuda_stream))
    v, d = DeviceVolume(dims, fmt), DeviceVolume(dims, fmt)
    t = torch.randint(0, 255, (min(v.nbytes, 1 << 28),), dtype=torch.uint8, device="cuda")
    if fmt == capi.FLOAT32:
        t = (torch.rand(t.numel() // 4, device="cuda")).view(torch.uint8)
    for off in range(0, v.nbytes, t.numel()):
        n = min(t.numel(), v.nbytes - off)
        capi.check(lib.vktb_memcpy(v.ptr + off, t.data_ptr(), n, capi.D2D))
    sd, dd = v.desc(), d.desc()
    s2 = sd
    if op.startswith("sumrange"):
        v2 = DeviceVolume(dims, fmt)
        capi.check(lib.vktb_memcpy(v2.ptr, v.ptr, v.nbytes, capi.D2D))
        s2 = v2.desc()
    z, l = capi.I3(0, 0, 0), capi.I3(*dims)
    F9, F3 = ctypes.c_float * 9, ctypes.c_float * 3
    half = F3(dims[0] / 2, dims[1] / 2, dims[2] / 2)
    box = (ctypes.c_float * 27)(*([1.0 / 27.0] * 27))
    cs, sn = math.cos(0.5), math.sin(0.5)
    agg = capi.VktbAggregates()
    runs = {
        "filter": lambda: capi.check(lib.vktb_apply_filter(ctypes.byref(dd), ctypes.byref(sd), box, capi.I3(3, 3, 3), 2)),
        "scale": lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(sd), z, l,
                                                          F9(1 / 1.5, 0, 0, 0, 1 / 1.5, 0, 0, 0, 1 / 1.5), half)),
        "rotate": lambda: capi.check(lib.vktb_affine_range(ctypes.byref(dd), ctypes.byref(sd), z, l,
                                                           F9(cs, -sn, 0, sn, cs, 0, 0, 0, 1), half)),
        # sources at another 16-byte phase than the destination rows
        "sumrange": lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(sd), ctypes.byref(s2), z,
                                                            capi.I3(dims[0] - 8, dims[1] - 8, dims[2] - 8), capi.I3(1, 2, 3))),
        "sumrange8": lambda: capi.check(lib.vktb_arith_range(capi.SUM, ctypes.byref(dd), ctypes.byref(sd), ctypes.byref(s2), z,
                                                             capi.I3(dims[0] - 8, dims[1] - 8, dims[2] - 8), capi.I3(8, 8, 8))),
        "copyrange": lambda: capi.check(lib.vktb_copy_range(ctypes.byref(dd), ctypes.byref(sd), capi.I3(1, 1, 1),
                                                            capi.I3(dims[0] - 7, dims[1] - 7, dims[2] - 7), z)),
        "scan": lambda: capi.check(lib.vktb_scan_range(ctypes.byref(sd), ctypes.byref(sd), z, l, z)),
        "agg": lambda: capi.check(lib.vktb_aggregates_range(ctypes.byref(sd), z, l, ctypes.byref(agg))),
    }
    run = runs[op]
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{op} {dims} {sys.argv[5]}: {ms:.3f} ms  {v.nbytes / ms / 1e6:.1f} GB/s per volume pass", flush=True)


if __name__ == "__main__":
    main()
