"""Multi-process CPU runs (gloo) of tests/_dist_worker.py at several world
sizes -- the analogue of running the reference suite under
``mpiexec -n {1,2,...}`` (reference ``tests/test_mpi.py:1-7``)."""
import functools
from pathlib import Path

import pytest

from .conftest import run_distributed

WORKER = Path(__file__).with_name("_dist_worker.py")
CASES = ["reduce_sum_scalar_and_array", "comm_object_collectives",
         "scatter_nd_chunks", "split_subcomms_mapping", "pipeline_generic",
         "pipeline_native", "world_size_invariance",
         "bfgs_all_ranks_same_result", "group_of_models",
         "experimental_scan_gd"]


@functools.lru_cache(maxsize=None)
def _results(world):
    res, out = run_distributed(
        WORKER, world, env={"CUDA_VISIBLE_DEVICES": ""}, timeout=900)
    return res


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", CASES)
def test_case(world, case):
    res = _results(world)
    assert res.get(case) == "ok", res.get(case)
