"""The C++ / C99 host mirror of the reference API (include/vkt, libvkt_b200.so) and the `vkt` CLI, compared with the
oracle.  The programs under tests/cpp read like the reference's examples; they are compiled here with g++/gcc and run
under the CPU policy (always) and the GPU policy (-m gpu)."""
import os
import subprocess

import numpy as np
import pytest

import cases
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "volkit_b200", "lib")
INC = os.path.join(ROOT, "include")
U8, U16, F32 = po.UINT16 - 1, po.UINT16, po.FLOAT32


@pytest.fixture(scope="module")
def programs(tmp_path_factory):
    if not os.path.exists(os.path.join(LIB, "libvkt_b200.so")):
        pytest.fail("volkit_b200/lib/libvkt_b200.so is missing: run __graft_entry__.build()")
    out = tmp_path_factory.mktemp("hostbin")
    link = ["-L" + LIB, "-lvkt_b200", "-lvktb200", "-Wl,-rpath," + LIB, "-lpthread"]
    cpp = str(out / "host_api_cpp")
    subprocess.run(["g++", "-std=c++14", "-O1", "-I" + INC, os.path.join(ROOT, "tests", "cpp", "host_api_cpp.cpp"),
                    "-o", cpp] + link, check=True)
    c = str(out / "host_api_c")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-D_POSIX_C_SOURCE=200809L", "-O1", "-I" + INC,
                    os.path.join(ROOT, "tests", "cpp", "host_api_c.c"), "-o", c] + link, check=True)
    return {"cpp": cpp, "c": c}


def raw(path, dims, dtype):
    return np.fromfile(path, dtype).reshape(dims[2], dims[1], dims[0])


def run_cpp(programs, policy, tmp):
    o = po.oracle()
    dims = (48, 40, 36)
    a, b = cases.mk(dims, U16, 1), cases.mk(dims, U16, 2)
    a.a.tofile(tmp / "a.raw")
    b.a.tofile(tmp / "b.raw")
    r = subprocess.run([programs["cpp"], policy, str(tmp)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout, r.stderr)

    full = ((0, 0, 0), dims)
    # SafeSum whole volume, then one voxel set through setValue(1,2,3, .75)
    ref = po.empty(dims, U16)
    o.arith_range(po.SAFESUM, ref, a, b, *full)
    got = raw(tmp / "out_sum.raw", dims, np.uint16)
    ref.a[3, 2, 1] = int.from_bytes(o.map_voxel(0.75, U16, 0.0, 1.0), "little")
    assert got.tobytes() == ref.a.tobytes()
    # Fill(.25) + SafeDiffRange with dstOffset
    ref = po.empty(dims, U16)
    o.fill_range(ref, *full, 0.25)
    o.arith_range(po.SAFEDIFF, ref, a, b, (3, 5, 7), (47, 38, 33), (1, 2, 3))
    assert raw(tmp / "out_diff.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    # Prod into Float32 with mapping (-1000, 3000)
    ref = po.empty(dims, F32, -1000.0, 3000.0)
    o.arith_range(po.PROD, ref, a, b, *full)
    assert raw(tmp / "out_prod.raw", dims, np.float32).tobytes() == ref.a.tobytes()
    # Fill(.5) + CopyRange with clamped src and dstOffset
    cd = (64, 48, 44)
    ref = po.empty(cd, U16)
    o.fill_range(ref, (0, 0, 0), cd, 0.5)
    o.copy_range(ref, a, (-3, -2, -1), (51, 42, 37), (1, 2, 3))
    assert raw(tmp / "out_cpy.raw", cd, np.uint16).tobytes() == ref.a.tobytes()
    # TransformRange unary on a copy of a, then binary with a copy of b
    t1, t2 = a.copy(), b.copy()
    o.transform_range1(t1, (1, 2, 3), (47, 39, 35), 0)
    o.transform_range2(t1, t2, *full, 0)
    assert raw(tmp / "out_t1.raw", dims, np.uint16).tobytes() == t1.a.tobytes()
    assert raw(tmp / "out_t2.raw", dims, np.uint16).tobytes() == t2.a.tobytes()
    # Resample up (Linear) and down (Nearest, UInt8 with mapping (-1, 2))
    ref = po.empty((72, 60, 54), U16)
    o.resample(ref, a, po.LINEAR)
    assert raw(tmp / "out_up.raw", (72, 60, 54), np.uint16).tobytes() == ref.a.tobytes()
    ref = po.empty((24, 20, 18), U8, -1.0, 2.0)
    o.resample(ref, a, po.NEAREST)
    assert raw(tmp / "out_down.raw", (24, 20, 18), np.uint8).tobytes() == ref.a.tobytes()
    # Flip along X (whole), Y (range + offset into a 64x48x44 volume filled with .5), Z in place; histogram of a range
    ref = po.empty(dims, U16)
    o.flip_range(ref, a, (0, 0, 0), dims, (0, 0, 0), 0)
    assert raw(tmp / "out_fx.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    ref = po.empty(cd, U16)
    o.fill_range(ref, (0, 0, 0), cd, 0.5)
    o.flip_range(ref, a, (2, 1, 3), (46, 39, 35), (5, 4, 3), 1)
    assert raw(tmp / "out_fy.raw", cd, np.uint16).tobytes() == ref.a.tobytes()
    ref = a.copy()
    o.flip_range(ref, ref, (0, 0, 0), dims, (0, 0, 0), 2)
    assert raw(tmp / "out_fz.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    bins, dropped = o.histogram_range(a, (1, 2, 3), (47, 39, 35), 300)
    assert dropped == 0 and np.fromfile(tmp / "out_hist.raw", np.uint64).tobytes() == bins.tobytes()
    ref = po.empty(dims, U16)
    o.fill_range(ref, *full, 0.125)
    o.scale(ref, a, (1.5, 0.75, 2.0), (20.0, 18.0, 16.0))
    assert raw(tmp / "out_scl.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    ref = po.empty(cd, U16)
    o.fill_range(ref, (0, 0, 0), cd, 0.25)
    o.rotate(ref, a, (1.0, 2.0, 0.5), 0.7, (24.0, 20.0, 18.0))
    assert raw(tmp / "out_rot.raw", cd, np.uint16).tobytes() == ref.a.tobytes()
    box = np.full((3, 3, 3), np.float32(1.0) / np.float32(27.0), np.float32)
    ref = po.empty(dims, U16)
    o.apply_filter(ref, a, box)
    assert raw(tmp / "out_flt.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    assert raw(tmp / "out_fin.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    # stream I/O: the file holds a's bytes, they come back unchanged; range rows as the reference addresses them
    assert (tmp / "io_48x40x36_uint16.raw").read_bytes() == a.a.tobytes()
    assert raw(tmp / "out_ioback.raw", dims, np.uint16).tobytes() == a.a.tobytes()
    part = o.rows_gather(b, (3, 1, 2), (41, 38, 30))
    assert (tmp / "io_part.raw").read_bytes() == part
    ref = po.empty(dims, U16)
    o.fill_range(ref, *full, 0.75)
    o.rows_scatter(ref, (3, 1, 2), (41, 38, 30), part)
    assert raw(tmp / "out_iorange.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    got_agg = po.Aggregates.from_buffer_copy((tmp / "out_agg.raw").read_bytes())
    ref_agg = o.aggregates_range(a, (1, 2, 3), (47, 39, 35))
    if policy == "cpu":
        assert got_agg.words() == ref_agg.words()            # the reference's sequential accumulation, word for word
    else:
        cases.check_aggregates(o, got_agg, ref_agg, a, (1, 2, 3), (47, 39, 35), policy)
    ref = po.empty(cd, U16)
    o.fill_range(ref, (0, 0, 0), cd, 0.5)
    o.scan_range(ref, a, (2, 1, 3), (46, 39, 35), (5, 4, 3))
    assert raw(tmp / "out_sc.raw", cd, np.uint16).tobytes() == ref.a.tobytes()
    ref = b.copy()
    o.scan_range(ref, ref, (45, 2, 1), (3, 38, 30))
    assert raw(tmp / "out_sci.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
    want = b"".join(x.tobytes() for x in o.brick_decompose(a, (20, 16, 16), (1, 2, 0), (2, 1, 1)))
    assert (tmp / "out_bricks.raw").read_bytes() == want
    if policy.startswith("gpu-multi"):
        # the device-group prelude: SafeSum, SumRange with a z offset and Resample on striped 256 x 200 x 90 volumes
        md = (256, 200, 90)
        i = np.arange(md[0] * md[1] * md[2], dtype=np.uint64)
        ma = po.Vol(((i * np.uint64(2654435761)) >> np.uint64(11)).astype(np.uint16).reshape(md[2], md[1], md[0]), U16)
        mb = po.Vol(((i * np.uint64(40503) + np.uint64(12345)) >> np.uint64(3)).astype(np.uint16).reshape(md[2], md[1], md[0]), U16)
        ref = po.empty(md, U16)
        o.arith_range(po.SAFESUM, ref, ma, mb, (0, 0, 0), md)
        o.arith_range(po.SUM, ref, ma, mb, (2, 3, 4), (250, 190, 80), (1, 2, 9))
        assert raw(tmp / "out_m_sum.raw", md, np.uint16).tobytes() == ref.a.tobytes()
        ref = po.empty((384, 300, 135), U16)
        o.resample(ref, ma, po.LINEAR)
        assert raw(tmp / "out_m_up.raw", (384, 300, 135), np.uint16).tobytes() == ref.a.tobytes()
    if policy == "gpu-stream":
        # the host-streaming prelude: same voxels as the migrating calls
        ref = po.empty(dims, U16)
        o.arith_range(po.SAFESUM, ref, a, b, *full)
        assert raw(tmp / "out_s_sum.raw", dims, np.uint16).tobytes() == ref.a.tobytes()
        ref = po.empty((72, 60, 54), U16)
        o.resample(ref, a, po.LINEAR)
        assert raw(tmp / "out_s_up.raw", (72, 60, 54), np.uint16).tobytes() == ref.a.tobytes()
        t = a.copy()
        o.transform_range1(t, (1, 2, 3), (47, 39, 35), 0)
        assert raw(tmp / "out_s_t.raw", dims, np.uint16).tobytes() == t.a.tobytes()


def run_c(programs, policy, tmp):
    o = po.oracle()
    r = subprocess.run([programs["c"], policy, str(tmp)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout, r.stderr)
    dims = (40, 32, 24)
    full = ((0, 0, 0), dims)
    v, v2 = po.empty(dims, U8), po.empty(dims, U8)
    o.fill_range(v, *full, 0.1)
    o.fill_range(v2, *full, 0.3)
    o.fill_range(v2, (4, 4, 4), (20, 20, 20), 0.9)
    o.copy_range(v, v2, (10, 4, 2), (34, 28, 20), (0, 0, 0))
    o.transform_range1(v, (2, 2, 2), (30, 30, 20), 0)
    v3 = v.copy()
    o.arith_range(po.SAFESUM, v3, v, v2, *full)
    o.arith_range(po.ABSDIFF, v2, v, v3, (1, 1, 1), (39, 31, 23), (0, 0, 0))
    assert raw(tmp / "c_out_volume.raw", dims, np.uint8).tobytes() == v.a.tobytes()
    assert raw(tmp / "c_out_volume2.raw", dims, np.uint8).tobytes() == v2.a.tobytes()
    assert raw(tmp / "c_out_volume3.raw", dims, np.uint8).tobytes() == v3.a.tobytes()
    want = b"".join(x.tobytes() for x in o.brick_decompose(v3, (16, 16, 16), (1, 1, 1), (1, 1, 1)))
    assert (tmp / "c_out_bricks.raw").read_bytes() == want


def cli_pipeline(device):
    """BASELINE config 1: declare-sv 64^3 uint8 | fill 0.1 | fill-range 1 [0,8)^3 | resample --dims 32^3 uint16."""
    vkt = os.path.join(LIB, "vkt")
    env = dict(os.environ, VKT_B200_DEVICE=device)
    cmd = (f"{vkt} declare-sv --dims 64 64 64 --data-format uint8 | {vkt} fill --value 0.1 | "
           f"{vkt} fill-range --value 1 --first 0 0 0 --last 8 8 8 | {vkt} resample --dims 32 32 32 --data-format uint16")
    r = subprocess.run(["bash", "-o", "pipefail", "-c", cmd], capture_output=True, env=env)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    hdr = np.frombuffer(out[:24], np.uint32)
    assert hdr.tolist() == [1, 0, 32, 32, 32, po.UINT16]
    assert np.frombuffer(out[24:44], np.float32).tolist() == [1.0, 1.0, 1.0, 0.0, 1.0]
    case = [c for c in cases.cases() if c[0] == "cli/config1"][0]
    ref = cases.run_case(po.oracle(), case)[1]
    assert out[44:] == ref.tobytes()


REF_CLI = os.path.join(ROOT, "oracle", "_ref", "vkt_ref_cli")
CLI_STAGES = ["declare-sv --dims 64 64 64 --data-format uint8", "fill --value 0.1",
              "fill-range --value 1 --first 0 0 0 --last 8 8 8", "resample --dims 32 32 32 --data-format uint16"]
# sha256 of the 65580 bytes the REFERENCE's own tool (oracle/_ref/vkt_ref_cli = src/cli/main.cpp, unmodified) writes at the
# end of BASELINE config 1, and its 44-byte header; made by tests/golden/make_cli_golden.py where /root/reference exists
CLI_GOLDEN = os.path.join(ROOT, "tests", "golden", "cli_config1.json")


def run_pipe(tools, device="cpu"):
    """stage i runs with tools[i]; returns the bytes after every stage"""
    env = dict(os.environ, VKT_B200_DEVICE=device)
    outs, data = [], b""
    for tool, stage in zip(tools, CLI_STAGES):
        r = subprocess.run([tool] + stage.split(), input=data, capture_output=True, env=env, timeout=300)
        assert r.returncode == 0, (tool, stage, r.stderr)
        data = r.stdout
        outs.append(data)
    return outs


@pytest.mark.needs_ref
def test_cli_pipe_bytes_equal_the_reference_tool():
    """the pipe format, checked against bytes the reference CLI wrote (not against a reading of its source): after every
    command of config 1, and with the two tools mixed in one pipeline in both orders"""
    if not os.path.exists(REF_CLI):
        pytest.skip("oracle/_ref/vkt_ref_cli not built")
    mine = os.path.join(LIB, "vkt")
    ref = run_pipe([REF_CLI] * 4)
    got = run_pipe([mine] * 4)
    for i, (a, b) in enumerate(zip(ref, got)):
        assert a == b, f"stage {i} ({CLI_STAGES[i]}) differs from the reference tool"
    assert run_pipe([REF_CLI, mine, REF_CLI, mine])[-1] == ref[-1]
    assert run_pipe([mine, REF_CLI, mine, REF_CLI])[-1] == ref[-1]
    import hashlib, json
    g = json.load(open(CLI_GOLDEN))
    assert hashlib.sha256(ref[-1]).hexdigest() == g["sha256"] and ref[-1][:44].hex() == g["header_hex"], \
        "tests/golden/cli_config1.json is stale: rerun tests/golden/make_cli_golden.py"


def _pipe(tool, stages, data=b"", device="cpu"):
    env = dict(os.environ, VKT_B200_DEVICE=device)
    for st in stages:
        r = subprocess.run([tool] + st.split(), input=data, capture_output=True, env=env, timeout=300)
        assert r.returncode == 0, (tool, st, r.stderr)
        data = r.stdout
    return data


# a longer pipeline over every command both tools have, on a volume with dims.x == dims.z (see the test below for why)
CLI_LONG = ["declare-sv --dims 20 12 20 --data-format uint16", "fill-range -val 0.7 --first 2 1 3 --last 17 9 8",
            "fill-range -val 0.2 --first 0 0 0 --last 5 5 5", "flip -ax x", "flip-range -ax z --first 1 1 1 --last 19 11 19",
            "resample --dims 30 7 30 -df uint8", "flip --axis Y", "flip -ax z --first 0 0 0 --last 30 7 15"]


@pytest.mark.needs_ref
def test_cli_flip_dump_and_option_spellings_equal_the_reference_tool():
    """flip / flip-range (all axes, with and without a range), dump / dump-range text, the short options (-val, -ax, -df),
    error messages and exit codes: byte for byte what the reference's own tool does"""
    if not os.path.exists(REF_CLI):
        pytest.skip("oracle/_ref/vkt_ref_cli not built")
    mine = os.path.join(LIB, "vkt")
    for k in range(1, len(CLI_LONG) + 1):
        assert _pipe(REF_CLI, CLI_LONG[:k]) == _pipe(mine, CLI_LONG[:k]), CLI_LONG[k - 1]
    vol = _pipe(REF_CLI, CLI_LONG)
    for st in ("dump", "dump-range --first 1 1 1 --last 4 3 3", "set-header"):
        assert _pipe(REF_CLI, [st], vol) == _pipe(mine, [st], vol), st
    env = dict(os.environ, VKT_B200_DEVICE="cpu")
    for bad in ("flip-range -ax y", "flip -ax q", "flip-range -ax x --first -1 0 0 --last 3 3 3",
                "flip-range --first 0 0 0 --last 31 7 30", "dump-range", "frobnicate"):
        a = subprocess.run([REF_CLI] + bad.split(), input=vol, capture_output=True, env=env, timeout=60)
        b = subprocess.run([mine] + bad.split(), input=vol, capture_output=True, env=env, timeout=60)
        assert a.returncode == b.returncode != 0 and a.stdout == b.stdout == b"", bad
        assert a.stderr.split(b"\n")[0].split(b":")[0] == b.stderr.split(b"\n")[0].split(b":")[0], (bad, a.stderr, b.stderr)


@pytest.mark.needs_ref
def test_cli_raw_files_equal_the_reference_tool(tmp_path):
    """`read -i` and `dump -i` of raw files (dims and format parsed from the name) byte for byte as the reference's own
    tool -- built with the reference's VolumeFile.cpp and RawFile.cpp; `write -o` round trip; the two cases where the
    reference tool crashes (a missing raw file, and every `write`: its VolumeFile opens raw files read-only) are errors /
    work here"""
    if not os.path.exists(REF_CLI):
        pytest.skip("oracle/_ref/vkt_ref_cli not built")
    mine = os.path.join(LIB, "vkt")
    env = dict(os.environ, VKT_B200_DEVICE="cpu")
    rng = np.random.RandomState(3)
    files = {"vol_20x12x20_uint16.raw": rng.randint(0, 65536, 20 * 12 * 20).astype(np.uint16),
             "small_9x7x5_uint8.raw": rng.randint(0, 256, 9 * 7 * 5).astype(np.uint8),
             "f_6x5x4_float32.raw": rng.standard_normal(6 * 5 * 4).astype(np.float32)}
    for name, data in files.items():
        data.tofile(tmp_path / name)
        for cmd in ("read", "dump", "dump-range --first 1 1 1 --last 4 3 3"):
            a = subprocess.run([REF_CLI] + cmd.split() + ["-i", name], capture_output=True, env=env, cwd=tmp_path, timeout=60)
            b = subprocess.run([mine] + cmd.split() + ["-i", name], capture_output=True, env=env, cwd=tmp_path, timeout=60)
            assert a.returncode == b.returncode == 0 and a.stdout == b.stdout and len(a.stdout) > 44, (name, cmd, b.stderr)
    # write: the volume comes back from the file it was written to
    vol = subprocess.run([mine, "read", "-i", "vol_20x12x20_uint16.raw"], capture_output=True, env=env, cwd=tmp_path).stdout
    w = subprocess.run([mine, "write", "-o", "out_20x12x20_uint16.raw"], input=vol, capture_output=True, env=env, cwd=tmp_path)
    assert w.returncode == 0 and (tmp_path / "out_20x12x20_uint16.raw").read_bytes() == files["vol_20x12x20_uint16.raw"].tobytes()
    # errors: same message and exit code where the reference tool reports one; an error instead of its crash otherwise
    for args in (["read"], ["read", "-i", "foo.txt"], ["write"]):
        a = subprocess.run([REF_CLI] + args, input=vol, capture_output=True, env=env, cwd=tmp_path, timeout=60)
        b = subprocess.run([mine] + args, input=vol, capture_output=True, env=env, cwd=tmp_path, timeout=60)
        assert a.returncode == b.returncode == 1 and a.stderr == b.stderr, (args, a.stderr, b.stderr)
    a = subprocess.run([REF_CLI, "read", "-i", "nothere_4x4x4_uint8.raw"], capture_output=True, env=env, cwd=tmp_path, timeout=60)
    b = subprocess.run([mine, "read", "-i", "nothere_4x4x4_uint8.raw"], capture_output=True, env=env, cwd=tmp_path, timeout=60)
    assert a.returncode < 0 and b.returncode == 1 and b"Cannot open input file" in b.stderr      # the reference: SIGSEGV


@pytest.mark.needs_ref
def test_cli_reads_non_cubic_volumes_where_the_reference_tool_does_not():
    """The reference's pipe reader builds the volume as (dims.x, dims.y, dims.X) (src/cli/main.cpp:66): any stage that reads a
    volume with dims.x != dims.z gets the wrong depth (20 x 12 x 9 comes out of `fill` as 20 x 12 x 20 with a tail of
    uninitialised voxels).  This tool reads the header as written; the defect is not copied."""
    if not os.path.exists(REF_CLI):
        pytest.skip("oracle/_ref/vkt_ref_cli not built")
    stages = ["declare-sv --dims 20 12 9 --data-format uint16", "fill --value 0.25", "flip -ax z --first 0 0 0 --last 20 12 5"]
    ref = _pipe(REF_CLI, stages)
    assert np.frombuffer(ref[8:20], np.int32).tolist() == [20, 12, 20]          # the reference's defect, as observed
    out = _pipe(os.path.join(LIB, "vkt"), stages)
    assert np.frombuffer(out[8:20], np.int32).tolist() == [20, 12, 9] and len(out) == 44 + 20 * 12 * 9 * 2
    o = po.oracle()
    v = po.empty((20, 12, 9), U16)
    o.fill_range(v, (0, 0, 0), (20, 12, 9), 0.25)
    w = v.copy()
    o.flip_range(w, v, (0, 0, 0), (20, 12, 5), (0, 0, 0), 2)
    assert out[44:] == w.a.tobytes()


def test_cli_config1_cpu_matches_the_reference_tools_bytes():
    import hashlib, json
    g = json.load(open(CLI_GOLDEN))
    out = run_pipe([os.path.join(LIB, "vkt")] * 4)[-1]
    assert len(out) == g["bytes"] and out[:44].hex() == g["header_hex"] and hashlib.sha256(out).hexdigest() == g["sha256"]


SWEEPS = [("sweep_fill_transform", 400), ("sweep_copy_range", 400), ("sweep_arithmetic", 500), ("sweep_resample", 400),
          ("sweep_resample_aligned", 20), ("sweep_flip_scan_histogram_bricks", 500), ("sweep_scale_rotate_filter", 400)]


@pytest.mark.parametrize("sweep,n", SWEEPS)
def test_random_sweep_through_the_c_api_cpu_policy(sweep, n):
    """tests/fuzz_sweeps.py through the host mirror's C API (vkt*SV calls on vktStructuredVolume handles) under the CPU
    policy: the serial bodies of volkit_b200/csrc/host/serial.cpp against the restatement, random cases"""
    import fuzz_sweeps as fz
    from host_adapter import HostMirror
    getattr(fz, sweep)(po.oracle(), HostMirror("cpu"), n, False)


def test_random_aggregates_through_the_c_api_cpu_policy():
    """ComputeAggregatesRange of the host mirror's CPU policy: the reference's sequential binary32 accumulation, word for word"""
    import fuzz_sweeps as fz
    from host_adapter import HostMirror
    fz.sweep_aggregates(po.oracle(), HostMirror("cpu"), 400, False)


@pytest.mark.gpu
@pytest.mark.parametrize("sweep,n", SWEEPS)
def test_random_sweep_through_the_c_api_gpu_policy(sweep, n):
    """the same cases under Device::GPU: create on the host, switch the policy, vkt*SV (migrate -> C-ABI -> kernels), switch
    back, read -- the whole drop-in stack as a reference user drives it, against the restatement"""
    import fuzz_sweeps as fz
    from host_adapter import HostMirror
    from volkit_b200 import capi
    lib = capi.load()
    before = lib.vktb_launch_count()
    getattr(fz, sweep)(po.oracle(), HostMirror("gpu"), n, False)
    assert lib.vktb_launch_count() - before >= n // 2, "the GPU policy did not launch kernels"


def test_cpp_api_cpu_policy(programs, tmp_path):
    run_cpp(programs, "cpu", tmp_path)


def test_c99_api_cpu_policy(programs, tmp_path):
    run_c(programs, "cpu", tmp_path)


def test_cli_config1_cpu():
    cli_pipeline("cpu")


def test_gpu_policy_fails_loudly_without_gpu(programs, tmp_path):
    """No CPU fallback: on a machine without a usable B200 the GPU policy must fail, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cases.mk((48, 40, 36), U16, 1).a.tofile(tmp_path / "a.raw")
    cases.mk((48, 40, 36), U16, 2).a.tofile(tmp_path / "b.raw")
    r = subprocess.run([programs["cpp"], "gpu", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no usable CUDA device" in r.stderr or "FAILED" in r.stderr


@pytest.mark.gpu
def test_cpp_api_gpu_policy(programs, tmp_path):
    run_cpp(programs, "gpu", tmp_path)


@pytest.mark.gpu
def test_cpp_api_gpu_policy_host_streaming(programs, tmp_path, monkeypatch):
    monkeypatch.setenv("VKTB_HOSTED_SLAB_KB", "16")      # several slabs even on the 48x40x36 test volumes
    run_cpp(programs, "gpu-stream", tmp_path)


@pytest.mark.gpu
def test_cpp_api_gpu_policy_device_group(programs, tmp_path, monkeypatch):
    """vkt:: calls of ONE process on a device group: real GPUs when two are visible, else the same GPU named twice
    (striped volumes, fan-out and per-stripe copies all run; only the peer loads stay local)"""
    import torch
    monkeypatch.setenv("VKTB_SHARD_MIN_MB", "1")
    n = torch.cuda.device_count()
    ids = ",".join(str(i) for i in range(min(n, 4))) if n >= 2 else "0,0,0"
    run_cpp(programs, "gpu-multi:" + ids, tmp_path)


@pytest.mark.gpu
def test_cpp_api_gpu_policy_allocation_failure(programs, tmp_path):
    """fault injection (VKTB_FAIL_ALLOC_AT): a migrate() that cannot allocate leaves the volume on the host, the vkt::
    call returns InvalidValue, and the same call succeeds once memory is available again"""
    r = subprocess.run([programs["cpp"], "gpu-oom", str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "oom ok" in r.stdout, (r.returncode, r.stdout, r.stderr)
    assert "injected by VKTB_FAIL_ALLOC_AT" in r.stderr      # the backend's message reaches the caller's stderr


@pytest.mark.gpu
def test_c_abi_allocation_failure_is_reported(monkeypatch):
    import ctypes
    from volkit_b200 import capi
    lib = capi.load()
    capi.check(lib.vktb_set_device(0))
    monkeypatch.setenv("VKTB_FAIL_ALLOC_AT", "2")
    p1, p2, p3 = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
    assert lib.vktb_alloc(ctypes.byref(p1), 1 << 20) == 0
    assert lib.vktb_alloc(ctypes.byref(p2), 1 << 20) != 0 and not p2.value
    assert b"out of memory (injected" in lib.vktb_last_error()
    assert lib.vktb_alloc(ctypes.byref(p3), 1 << 20) == 0      # only the n-th call fails
    monkeypatch.delenv("VKTB_FAIL_ALLOC_AT")
    capi.check(lib.vktb_free(p1)); capi.check(lib.vktb_free(p3))


@pytest.mark.gpu
def test_c99_api_gpu_policy(programs, tmp_path):
    run_c(programs, "gpu", tmp_path)


@pytest.mark.gpu
def test_cli_config1_gpu():
    cli_pipeline("gpu")
    import hashlib, json
    g = json.load(open(CLI_GOLDEN))
    out = run_pipe([os.path.join(LIB, "vkt")] * 4, "gpu")[-1]
    assert out[:44].hex() == g["header_hex"] and hashlib.sha256(out).hexdigest() == g["sha256"], \
        "the GPU policy's pipe differs from the bytes the reference tool writes"
    if os.path.exists(REF_CLI):      # travels with the snapshot: the reference tool itself, run on this box
        assert run_pipe([REF_CLI] * 4)[-1] == out
