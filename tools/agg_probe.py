#!/usr/bin/env python
"""Times vktb_aggregates_range on a device-resident volume.  usage: agg_probe.py X Y Z FMT [lo hi]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

from volkit_b200 import capi  # noqa: E402
from volkit_b200.device import DeviceVolume  # noqa: E402

FMT = {"u8": capi.UINT8, "u16": capi.UINT16, "u32": capi.UINT32, "f32": capi.FLOAT32}


def main():
    dims, fmt = tuple(int(a) for a in sys.argv[1:4]), FMT[sys.argv[4]]
    lo, hi = (float(sys.argv[5]), float(sys.argv[6])) if len(sys.argv) > 6 else (0.0, 1.0)
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    capi.check(lib.vktb_set_stream(stream.cuda_stream))
    v = DeviceVolume(dims, fmt, lo, hi)
    t = torch.randint(0, 255, (min(v.nbytes, 1 << 28),), dtype=torch.uint8, device="cuda")
    if fmt == capi.FLOAT32:
        t = (torch.rand(t.numel() // 4, device="cuda")).view(torch.uint8)
    for off in range(0, v.nbytes, t.numel()):
        n = min(t.numel(), v.nbytes - off)
        capi.check(lib.vktb_memcpy(v.ptr + off, t.data_ptr(), n, capi.D2D))
    d = v.desc()
    z, l = capi.I3(0, 0, 0), capi.I3(*dims)
    agg = capi.VktbAggregates()

    def run():
        capi.check(lib.vktb_aggregates_range(ctypes.byref(d), z, l, ctypes.byref(agg)))
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
    print(f"aggregates {dims} {sys.argv[4]} ({lo},{hi}): {ms:.3f} ms  {v.nbytes / ms / 1e6:.1f} GB/s  "
          f"min {agg.min} max {agg.max} mean {agg.mean} var {agg.var} sum {agg.sum} prod {agg.prod}", flush=True)


if __name__ == "__main__":
    main()
