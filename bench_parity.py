"""Full-size parity for bench.py / bench_ops.py / bench_sharded.py (CHECKER code: besides bench.py's CPU legs the
only benchmark module that touches oracle/).

Every op the benchmark times at a BASELINE size is compared with the oracle at that size on sampled z-planes: the
first and last plane, eight seeded-random ones and every slab-boundary plane +- 1.  The inputs of a plane are
regenerated on the host from the counter-based synthetic stream, the oracle (oracle/vkt_oracle.c)
computes the plane through a *Range call or a z-window entry point, and the bytes are compared with the plane
downloaded from the device.  Oracle work runs on a thread pool (ctypes calls release the GIL).

Where an op's result is a function of a small domain the oracle can enumerate (UInt8 operand pairs, UInt16 values),
the whole device volume can additionally be compared against the oracle's table: torch only indexes.
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from oracle import pyoracle as po

NP = {po.UINT8: np.uint8, po.UINT16: np.uint16, po.UINT32: np.uint32, po.FLOAT32: np.float32}


def sample_planes(nz, boundaries=(), n_random=8, seed=12345, lo=0):
    """first, last, n_random seeded-random planes and the planes either side of every boundary, as sorted global
    plane numbers in [lo, lo + nz)."""
    rng = np.random.RandomState(seed)
    s = {lo, lo + nz - 1}
    if nz > 2:
        s |= {int(lo + v) for v in rng.randint(1, nz - 1, size=n_random)}
    for b in boundaries:
        for z in (b - 2, b - 1, b, b + 1):
            if lo <= z < lo + nz:
                s.add(int(z))
    return sorted(s)


class Parity:
    def __init__(self, lib, capi, threads=None):
        self.lib, self.capi = lib, capi
        self.o = po.oracle()
        self.pool = ThreadPoolExecutor(threads or min(32, os.cpu_count() or 1))
        self.results = {}     # name -> {"ok": bool, "planes": n}

    # ---- plumbing
    def regen(self, seed, fmt, dims, z, nz=1, first_voxel=0):
        """planes [z, z + nz) of the synthetic volume `seed` of x/y extent dims[0] x dims[1] as a (nz, Y, X) array"""
        plane = dims[0] * dims[1]
        return po.synth_numpy(plane * nz, fmt, seed, first_voxel + z * plane).reshape(nz, dims[1], dims[0])

    def fetch(self, ptr, dims, fmt, z, nz=1):
        """planes [z, z + nz) of the device volume at `ptr` (D2H through the C-ABI)"""
        out = np.empty((nz, dims[1], dims[0]), NP[fmt])
        plane_bytes = dims[0] * dims[1] * out.itemsize
        self.capi.check(self.lib.vktb_memcpy(out.ctypes.data, ptr + z * plane_bytes, out.nbytes, self.capi.D2H),
                        "parity D2H")
        return out

    def check(self, name, planes, got, expected):
        """got(z) -> ndarray downloaded on the calling thread; expected(z) -> ndarray computed on the pool.  Equal bytes
        (Float32: equal bytes, or NaN in the same places and equal bytes elsewhere)."""
        futures = [(z, got(z), self.pool.submit(expected, z)) for z in planes]
        ok, bad = True, []
        for z, g, f in futures:
            e = f.result()
            same = g.tobytes() == e.tobytes()
            if not same and g.dtype == np.float32 and g.shape == e.shape:
                gn, en = np.isnan(g), np.isnan(e)
                same = bool((gn == en).all()) and g[~gn].tobytes() == e[~en].tobytes()
            if not same:
                ok = False
                bad.append(z if isinstance(z, (tuple, list)) else int(z))
        self.results[name] = {"ok": ok, "planes": len(planes)}
        if bad:
            self.results[name]["mismatched_planes"] = bad[:8]
        return ok

    def record(self, name, ok, what):
        self.results[name] = {"ok": bool(ok), "check": what}
        return bool(ok)

    def all_ok(self):
        return all(r["ok"] for r in self.results.values())

    def summary(self):
        """compact form kept in the bench line"""
        return {"all": self.all_ok(), "ops": len(self.results),
                "planes": int(sum(r.get("planes", 0) for r in self.results.values())),
                "failed": [k for k, r in self.results.items() if not r["ok"]]}

    # ---- oracle tables over small domains
    def pair_table_u8(self, op, maps=((0.0, 1.0),) * 3):
        """result byte of an arithmetic op for every UInt8 operand pair, from the oracle: table[a * 256 + b]"""
        a = po.Vol(np.repeat(np.arange(256, dtype=np.uint8), 256).reshape(1, 256, 256), po.UINT8, *maps[1])
        b = po.Vol(np.tile(np.arange(256, dtype=np.uint8), 256).reshape(1, 256, 256), po.UINT8, *maps[2])
        d = po.empty((256, 256, 1), po.UINT8, *maps[0])
        self.o.arith_range(op, d, a, b, (0, 0, 0), (256, 256, 1))
        return d.a.reshape(-1).copy()

    def bin_table_u16(self, num_bins, lo=0.0, hi=1.0):
        """histogram bin of every UInt16 voxel value from the oracle's ComputeHistogram (-1 = dropped)"""
        vol = po.Vol(np.arange(65536, dtype=np.uint16).reshape(1, 1, 65536), po.UINT16, lo, hi)
        table = np.full(65536, -1, np.int64)
        for v in range(65536):
            bins, dropped = self.o.histogram_range(vol, (v, 0, 0), (v + 1, 1, 1), num_bins)
            if not dropped:
                table[v] = int(np.argmax(bins))
        return table
