#!/usr/bin/env python
"""Times vktb_histogram_range / vktb_flip_range on a device-resident volume.  usage: hist_probe.py EDGE FMT BINS"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

from volkit_b200 import capi  # noqa: E402
from volkit_b200.device import DeviceVolume  # noqa: E402

FMT = {"u8": capi.UINT8, "u16": capi.UINT16, "u32": capi.UINT32, "f32": capi.FLOAT32}


def main():
    edge, fmt, nb = int(sys.argv[1]), FMT[sys.argv[2]], int(sys.argv[3])
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    capi.check(lib.vktb_set_stream(stream.cuda_stream))
    v = DeviceVolume((edge,) * 3, fmt)
    t = torch.randint(0, 255, (min(v.nbytes, 1 << 28),), dtype=torch.uint8, device="cuda")
    if fmt == capi.FLOAT32:
        t = (torch.rand(t.numel() // 4, device="cuda")).view(torch.uint8)
    for off in range(0, v.nbytes, t.numel()):
        n = min(t.numel(), v.nbytes - off)
        capi.check(lib.vktb_memcpy(v.ptr + off, t.data_ptr(), n, capi.D2D))
    bins = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(bins), nb * 8))
    d = v.desc()
    z, l = capi.I3(0, 0, 0), capi.I3(edge, edge, edge)

    def run():
        capi.check(lib.vktb_histogram_range(ctypes.byref(d), z, l, bins, nb, None))
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
    print(f"histogram {edge}^3 {sys.argv[2]} {nb} bins: {ms:.3f} ms  {v.nbytes / ms / 1e6:.1f} GB/s", flush=True)


if __name__ == "__main__":
    main()
