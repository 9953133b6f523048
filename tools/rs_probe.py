#!/usr/bin/env python
"""Times vktb_resample on a list of (src dims, dst dims, format, filter) shapes (device-resident, CUDA events).
usage: python tools/rs_probe.py "2048,2048,2048>1024,1024,1024:u16:0" ...   (filter 0 nearest, 1 linear, 2 trilinear)"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

from volkit_b200 import capi  # noqa: E402
from volkit_b200.device import DeviceVolume  # noqa: E402

FMT = {"u8": capi.UINT8, "u16": capi.UINT16, "u32": capi.UINT32, "f32": capi.FLOAT32}


def main():
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    capi.check(lib.vktb_set_stream(stream.cuda_stream))
    for spec in sys.argv[1:]:
        shape, fmt, filt = spec.split(":")
        s, d = shape.split(">")
        S, D = tuple(int(v) for v in s.split(",")), tuple(int(v) for v in d.split(","))
        f = FMT[fmt]
        bpv = capi.BYTES_PER_VOXEL[f]
        src, dst = DeviceVolume(S, f), DeviceVolume(D, f)
        t = torch.randint(0, 255, (min(src.nbytes, 1 << 28),), dtype=torch.uint8, device="cuda")
        for off in range(0, src.nbytes, t.numel()):
            n = min(t.numel(), src.nbytes - off)
            capi.check(lib.vktb_memcpy(src.ptr + off, t.data_ptr(), n, capi.D2D))
        torch.cuda.synchronize()
        ds, dd = src.desc(), dst.desc()

        def run():
            capi.check(lib.vktb_resample(ctypes.byref(dd), ctypes.byref(ds), int(filt)))
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K = 5
        for _ in range(K):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        nd, ns = D[0] * D[1] * D[2], S[0] * S[1] * S[2]
        # src rows actually hit by a one-tap filter
        hit = S[0] * min(S[1], D[1]) * min(S[2], D[2])
        print(f"{spec:60s} {ms:9.3f} ms  dst {nd / ms / 1e6:8.1f} GVox/s  whole-src+dst {(ns + nd) * bpv / ms / 1e6:8.1f} GB/s  "
              f"hit-rows+dst {(hit + nd) * bpv / ms / 1e6:8.1f} GB/s", flush=True)
        src.free(); dst.free()


if __name__ == "__main__":
    main()
