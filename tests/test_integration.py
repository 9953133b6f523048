"""The drop-in boundary, compiled (integration/Makefile): the reference's own L3 front files built with
VKT_HAVE_CUDA=1, linked with integration/B200Backend.cpp against libvktb200.so, and a user program written against
the REFERENCE's public headers that runs every hot-path algorithm under Device::CPU and Device::GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "oracle", "_ref_gpu")
PROG = os.path.join(OUT, "both_policies")

needs_build = pytest.mark.skipif(not os.path.exists(PROG), reason="oracle/_ref_gpu not built (reference sources absent)")


@needs_build
def test_shim_defines_every_cuda_backend_symbol_the_fronts_reference():
    """libvkt_ref_gpu.so links: no `*_cuda` symbol and no CUDA runtime call is left undefined; what it imports from
    the backend is exactly C-ABI entry points declared in include/vktb/backend.h"""
    nm = subprocess.run(["nm", "-D", "-C", os.path.join(OUT, "libvkt_ref_gpu.so")], capture_output=True, text=True, check=True).stdout
    undefined = [l.split(" U ", 1)[1].strip() for l in nm.splitlines() if " U " in l]
    assert not [u for u in undefined if "_cuda" in u or u.startswith("cuda")], undefined
    imported = sorted(u for u in undefined if u.startswith("vktb_"))
    header = open(os.path.join(ROOT, "include", "vktb", "backend.h")).read()
    assert imported and all(f"{name}(" in header for name in imported), imported
    for need in ("vktb_alloc", "vktb_free", "vktb_memcpy", "vktb_arith_range", "vktb_copy_range", "vktb_fill_range",
                 "vktb_resample", "vktb_transform_range1"):
        assert need in imported
    defined = [l for l in nm.splitlines() if " T " in l]
    for sym in ("vkt::SafeSumRange_cuda", "vkt::CopyRange_cuda", "vkt::FillRange_cuda", "vkt::Resample_cuda",
                "vkt::MemsetRange_cuda", "vkt::TransformRange_cuda", "vkt::CudaTimer::CudaTimer"):
        assert any(sym in l for l in defined), sym


@needs_build
def test_program_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    r = subprocess.run([PROG], capture_output=True, text=True)
    assert r.returncode == 99 and "no usable GPU" in r.stderr


@pytest.mark.gpu
def test_reference_fronts_agree_under_both_policies():
    """the reference's header-inline ManagedBuffer<T>::migrate() drives vktb_alloc / vktb_memcpy / vktb_free on the
    B200, its fronts dispatch into the shim, and Device::CPU (reference serial code) == Device::GPU byte for byte"""
    assert os.path.exists(PROG), "oracle/_ref_gpu/both_policies missing: run __graft_entry__.build() where /root/reference exists"
    r = subprocess.run([PROG], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ALL IDENTICAL" in r.stdout and "MISMATCH" not in r.stdout, r.stdout
    assert r.stdout.count("identical") >= 14
