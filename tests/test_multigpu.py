"""Multi-GPU runs (one rank per GPU, torchrun) of tests/_dist_worker.py: the
same world-size-invariance assertions as the CPU runs, now through the fused
kernels with the in-kernel NVLink exchange, plus the collective kernels."""
import functools
from pathlib import Path

import pytest
import torch

from .conftest import run_distributed
from .test_distributed_cpu import CASES as CPU_CASES

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

WORKER = Path(__file__).with_name("_dist_worker.py")
CASES = CPU_CASES + ["gpu_allreduce_kernels", "gpu_fused_step_uses_one_launch",
                     "gpu_linear_link_world_invariance",
                     "gpu_scatter_nd_over_nvlink", "gpu_group_device_steps",
                     "gpu_bfgs_rank_skew_stress", "gpu_collective_soak",
                     "gpu_linked_model_world_invariance",
                     "gpu_linear_link_bfgs_and_group"]


def _worlds():
    import os
    if os.environ.get("MGB_TEST_WORLDS"):
        return [int(w) for w in os.environ["MGB_TEST_WORLDS"].split(",")]
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n] or [2]


@functools.lru_cache(maxsize=None)
def _results(world):
    res, out = run_distributed(
        WORKER, world, timeout=900,
        env={"MULTIGRAD_B200_SPIN_TIMEOUT_MS": "10000"})
    return res


@pytest.mark.parametrize("world", _worlds())
@pytest.mark.parametrize("case", CASES)
def test_case(world, case):
    res = _results(world)
    assert res.get(case) == "ok", res.get(case)
