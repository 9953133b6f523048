"""Sanitizer runs of the HOST side of the boundary (`-fsanitize=address,undefined` on the host
mirror, a thread sanitizer run for the policy map and the resource registry that the reference mutates without locks).
The host mirror's sources are compiled again with the sanitizer (the shipped libvkt_b200.so is not instrumented) and
linked with the test programs; everything runs under the CPU policy, so no GPU is needed.  compute-sanitizer for the
device side is closed on this pool."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "volkit_b200", "lib")
INC = os.path.join(ROOT, "include")
HOST = os.path.join(ROOT, "volkit_b200", "csrc", "host")
HOST_SRCS = ["policy.cpp", "memory.cpp", "volume.cpp", "serial.cpp", "algorithms.cpp", "io.cpp", "c_api.cpp"]


def build(out, program, flags):
    exe = str(out / (os.path.splitext(program)[0] + "_" + flags[0].split("=")[1].split(",")[0]))
    cmd = ["g++", "-std=c++14", "-O1", "-g", "-ffp-contract=off", "-fno-omit-frame-pointer", "-I" + INC] + flags + \
          [os.path.join(HOST, s) for s in HOST_SRCS] + [os.path.join(ROOT, "tests", "cpp", program), "-o", exe,
           "-L" + LIB, "-lvktb200", "-Wl,-rpath," + LIB, "-lpthread"]
    subprocess.run(cmd, check=True, timeout=600)
    return exe


@pytest.fixture(scope="module")
def need_backend():
    if not os.path.exists(os.path.join(LIB, "libvktb200.so")):
        pytest.fail("volkit_b200/lib/libvktb200.so is missing: run __graft_entry__.build()")


def test_thread_sanitizer_policy_map_and_registry(need_backend, tmp_path):
    exe = build(tmp_path, "host_threads.cpp", ["-fsanitize=thread"])
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1 exitcode=66", VKT_B200_PINNED="0")
    r = subprocess.run([exe, "8", "40"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "threads ok" in r.stdout, (r.returncode, r.stdout[-2000:], r.stderr[-4000:])
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-4000:]


def test_address_and_undefined_sanitizer_cpu_policy(need_backend, tmp_path):
    """the whole C++ API walk of tests/cpp/host_api_cpp.cpp under the CPU policy (serial bodies, streams, bricks, scan ...)"""
    import numpy as np
    import cases
    from oracle import pyoracle as po
    exe = build(tmp_path, "host_api_cpp.cpp", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"])
    a, b = cases.mk((48, 40, 36), po.UINT16, 1), cases.mk((48, 40, 36), po.UINT16, 2)
    a.a.tofile(tmp_path / "a.raw")
    b.a.tofile(tmp_path / "b.raw")
    # detect_leaks=0: the backend library keeps per-process singletons alive on purpose (memory.cu, CopyPool)
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:halt_on_error=1:exitcode=67", UBSAN_OPTIONS="print_stacktrace=1",
               VKT_B200_PINNED="0")
    r = subprocess.run([exe, "cpu", str(tmp_path)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.returncode, r.stdout[-2000:], r.stderr[-4000:])
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-4000:]
