"""The sharded paths on real GPUs, every rank checking its slab (boundary planes included) against the whole-volume
oracle result (tests/dist_worker.py).

* NCCL: one rank per visible GPU, capped at 4 -- needs at least two GPUs (`gpurun --gpus 2`); NCCL refuses two ranks
  on one device.
* gloo-cuda: TWO ranks on GPU 0, the real C-ABI kernels on CUDA slabs, halo planes carried by gloo through host memory:
  the slab / halo / window logic and the kernels' window entry points are exercised on the single-GPU boxes the driver
  tests on.
"""
import pytest

from test_sharded_gloo import launch, launch_expect_mismatch

pytestmark = pytest.mark.gpu


def test_resample_sharded_nccl():
    import torch
    n = torch.cuda.device_count()
    assert n >= 1, "no GPU visible"
    if n < 2:
        pytest.skip("one GPU visible: the NCCL exchange needs two (covered by the gloo-cuda test on this box)")
    launch(min(n, 4), "nccl")


def test_sharded_two_ranks_one_gpu():
    launch(2, "gloo-cuda")


def test_sharded_three_ranks_one_gpu():
    launch(3, "gloo-cuda")


def test_broken_halo_is_detected_on_gpu():
    launch_expect_mismatch(2, "gloo-cuda")
