"""The sharded Resample path on real GPUs: NCCL point-to-point halo exchange + vktb_resample_window per rank.
world = number of visible GPUs capped at 4 (1 on the single-GPU boxes the driver uses; `gpurun --gpus 2` for more)."""
import pytest

from test_sharded_gloo import launch

pytestmark = pytest.mark.gpu


def test_resample_sharded_nccl():
    import torch
    n = torch.cuda.device_count()
    assert n >= 1, "no GPU visible"
    launch(min(n, 4), "nccl")
