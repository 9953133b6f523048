"""Multi-rank sharding logic on CPU: world_size 2 and 3 over gloo (no GPU needed).  See tests/dist_worker.py."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def spawn(world, backend, extra_env=None):
    port = free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), GLOO_SOCKET_IFNAME="lo", **(extra_env or {}))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), backend], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    return procs


def launch(world, backend):
    procs = spawn(world, backend)
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
    assert any("ok world=%d" % world in o for o in outs), outs


def launch_expect_mismatch(world, backend):
    """A lost halo plane (VKTB_TEST_BREAK_HALO=1: the first incoming transfer lands in a scratch buffer) must turn the
    slab comparison red: some rank reports MISMATCH and exits non-zero.  The others may then wait in a collective for
    the dead rank; they are killed (by their own PIDs) once the verdict is in."""
    import time
    procs = spawn(world, backend, {"VKTB_TEST_BREAK_HALO": "1"})
    deadline = time.time() + 600
    failed = None
    while time.time() < deadline and failed is None:
        for p in procs:
            if p.poll() is not None and p.returncode != 0:
                failed = p
        if all(p.poll() is not None for p in procs):
            break
        time.sleep(0.2)
    for p in procs:
        if p.poll() is None:
            p.kill()
    outs = [p.communicate()[0] for p in procs]
    assert failed is not None, "a lost halo plane went unnoticed:\n" + "\n".join(outs)
    assert any("MISMATCH" in o for o in outs), outs


@pytest.mark.parametrize("world", [2, 3])
def test_resample_sharded_gloo(world):
    launch(world, "gloo")


def test_broken_halo_is_detected_gloo():
    launch_expect_mismatch(2, "gloo")


def test_halo_plan_properties():
    from volkit_b200 import capi, sharded
    for S, D in (((8, 8, 17), (8, 8, 23)), ((4, 4, 4096), (4, 4, 2048)), ((4, 4, 2048), (4, 4, 3072)), ((4, 4, 9), (4, 4, 31))):
        for world in (1, 2, 3, 4, 8):
            for filt in (capi.NEAREST, capi.LINEAR, capi.TRILINEAR):
                windows, transfers = sharded.halo_plan(filt, D, S, world)
                for r, (w0, w1) in enumerate(windows):
                    z0, z1 = sharded.slab_range(S[2], world, r)
                    got = set(range(max(w0, z0), min(w1, z1)))
                    for t in transfers:
                        assert t.src_rank != t.dst_rank
                        o0, o1 = sharded.slab_range(S[2], world, t.src_rank)
                        assert o0 <= t.z0 < t.z1 <= o1          # the sender owns what it sends
                        if t.dst_rank == r:
                            got |= set(range(t.z0, t.z1))
                    assert got == set(range(w0, w1)), (S, D, world, filt, r)   # window fully covered, nothing extra
    # exact 2:1: the 8 taps of dst plane z are src planes 2z and 2z+1 -> never cross a slab boundary
    _, transfers = sharded.halo_plan(capi.TRILINEAR, (4, 4, 2048), (4, 4, 4096), 8)
    assert transfers == []
    # 2:3 up-sampling does cross
    _, transfers = sharded.halo_plan(capi.TRILINEAR, (4, 4, 3072), (4, 4, 2048), 8)
    assert len(transfers) == 7 and all(t.z1 - t.z0 == 1 for t in transfers)


def test_range_plan_properties():
    """Every p.z of the range is computed by exactly one rank, whose window holds the (clamped) src plane it reads;
    transfers only carry planes their sender owns; a congruent call (no z shift) moves nothing."""
    from volkit_b200 import sharded
    for kind in ("arith", "copy"):
        for Sz, Dz, first, last, off in ((19, 26, 3, 17, 7), (19, 26, 0, 19, 0), (64, 64, 0, 64, 0), (9, 40, 2, 8, 31),
                                         (19, 26, -4, 23, 2)):
            if kind == "arith" and (first < 0 or last > Sz):
                continue
            for world in (1, 2, 3, 4, 8):
                windows, sub, transfers = sharded.range_plan(kind, first, last, off, Dz, Sz, world)
                owner = {}
                for r, (pz0, pz1) in enumerate(sub):
                    dz0, dz1 = sharded.slab_range(Dz, world, r)
                    for pz in range(pz0, pz1):
                        assert pz not in owner
                        owner[pz] = r
                        dplane = pz + off if kind == "arith" else pz - first + off
                        assert dz0 <= dplane < dz1
                        splane = pz if kind == "arith" else min(max(pz, 0), Sz - 1)
                        assert windows[r][0] <= splane < windows[r][1]
                want = [pz for pz in range(first, last) if 0 <= (pz + off if kind == "arith" else pz - first + off) < Dz]
                assert sorted(owner) == want
                for t in transfers:
                    o0, o1 = sharded.slab_range(Sz, world, t.src_rank)
                    assert t.src_rank != t.dst_rank and o0 <= t.z0 < t.z1 <= o1
                for r, (w0, w1) in enumerate(windows):
                    z0, z1 = sharded.slab_range(Sz, world, r)
                    got = set(range(max(w0, z0), min(w1, z1)))
                    for t in transfers:
                        if t.dst_rank == r:
                            got |= set(range(t.z0, t.z1))
                    assert got == set(range(w0, w1))
                if Sz == Dz and off == 0 and first == 0:
                    assert transfers == []
