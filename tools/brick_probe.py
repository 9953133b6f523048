"""BrickDecompose alone (one vktb_copy_range_batch launch) for profiling: python tools/brick_probe.py [edge] [brick] [halo]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from volkit_b200 import capi  # noqa: E402
from volkit_b200.device import DeviceVolume  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 64
halo = int(sys.argv[3]) if len(sys.argv) > 3 else 1
lib = capi.load()
capi.check(lib.vktb_set_device(0))
src = DeviceVolume((S, S, S), capi.UINT16)
capi.check(lib.vktb_fill_range(ctypes.byref(src.desc()), capi.I3(0, 0, 0), capi.I3(S, S, S), 0.3))
nb = (S + bs - 1) // bs
edge = lambda i: halo + min(bs, S - i * bs) + halo
bricks, firsts, lasts, total = [], [], [], 0
for z in range(nb):
    for y in range(nb):
        for x in range(nb):
            bd = (edge(x), edge(y), edge(z))
            bricks.append((total, bd))
            firsts += [x * bs - halo, y * bs - halo, z * bs - halo]
            lasts += [min((x + 1) * bs, S) + halo, min((y + 1) * bs, S) + halo, min((z + 1) * bs, S) + halo]
            total += bd[0] * bd[1] * bd[2]
pool = ctypes.c_void_p()
capi.check(lib.vktb_alloc(ctypes.byref(pool), total * 2))
descs = (capi.VktbVol * len(bricks))(*[capi.vol(pool.value + 2 * o, bd, capi.UINT16) for o, bd in bricks])
i32 = lambda v: (ctypes.c_int32 * len(v))(*v)
bf, bl, bo = i32(firsts), i32(lasts), i32([0] * len(firsts))
sd = src.desc()
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
capi.check(lib.vktb_set_stream(stream.cuda_stream))
for rep in range(4):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    capi.check(lib.vktb_copy_range_batch(len(bricks), descs, ctypes.byref(sd), bf, bl, bo))
    b.record()
    b.synchronize()
    ms = a.elapsed_time(b)
    print(f"{len(bricks)} bricks of {bs}^3+{halo}: {ms:.3f} ms, {4 * total / ms / 1e6:.1f} GB/s (2 bpv per brick voxel)")
