#!/usr/bin/env python
"""File <-> device throughput of vktb_stream_in / vktb_stream_out against the read-everything-then-migrate sequence the
reference's InputStream::read + ManagedBuffer::migrate() amounts to.  usage: io_probe.py GIB [DIR]   (DIR: /dev/shm)"""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402

from volkit_b200 import capi  # noqa: E402


def main():
    gib = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
    d = sys.argv[2] if len(sys.argv) > 2 else "/dev/shm"
    n = int(gib * (1 << 30))
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    path, path2 = os.path.join(d, "vktb_io_probe.raw"), os.path.join(d, "vktb_io_probe_out.raw")
    block = np.random.default_rng(0).integers(0, 256, 1 << 26, dtype=np.uint8)
    with open(path, "wb") as f:
        for off in range(0, n, block.size):
            f.write(block[:min(block.size, n - off)].tobytes())
    dev = ctypes.c_void_p()
    capi.check(lib.vktb_alloc(ctypes.byref(dev), n))
    res = {}
    for rep in range(2):                                   # the second round is the warm one
        # (a) streamed: the file is read straight into the pinned staging slots
        f = open(path, "rb", buffering=0)

        def produce(ctx, buf, length):
            return f.readinto((ctypes.c_char * length).from_address(buf))
        t = time.perf_counter()
        total = ctypes.c_size_t(0)
        capi.check(lib.vktb_stream_in(dev, n, capi.READ_FN(produce), None, ctypes.byref(total)))
        capi.check(lib.vktb_sync())
        res["stream_in"] = n / (time.perf_counter() - t) / 1e9
        f.close()
        assert total.value == n
        # (b) the reference's order: read the whole file into (pageable) host memory, then migrate
        t = time.perf_counter()
        host = np.empty(n, np.uint8)
        with open(path, "rb", buffering=0) as g:
            g.readinto(host)
        capi.check(lib.vktb_memcpy(dev, host.ctypes.data, n, capi.H2D))
        capi.check(lib.vktb_sync())
        res["read_then_migrate"] = n / (time.perf_counter() - t) / 1e9
        # (c) streamed out
        o = open(path2, "wb", buffering=0)

        def consume(ctx, buf, length):
            return o.write((ctypes.c_char * length).from_address(buf))
        t = time.perf_counter()
        capi.check(lib.vktb_stream_out(dev, n, capi.WRITE_FN(consume), None, ctypes.byref(total)))
        o.close()
        res["stream_out"] = n / (time.perf_counter() - t) / 1e9
        # (d) migrate back, then write
        t = time.perf_counter()
        capi.check(lib.vktb_memcpy(host.ctypes.data, dev, n, capi.D2H))
        with open(path2, "wb", buffering=0) as g:
            g.write(host)
        res["migrate_then_write"] = n / (time.perf_counter() - t) / 1e9
        del host
    same = open(path, "rb").read(1 << 20) == open(path2, "rb").read(1 << 20)
    os.remove(path); os.remove(path2)
    print(f"io_probe {gib} GiB in {d}: " + "  ".join(f"{k} {v:.2f} GB/s" for k, v in res.items()) + f"  round trip ok: {same}",
          flush=True)


if __name__ == "__main__":
    main()
