"""Multi-GPU tests (need >= 2 B200s: `gpurun --gpus 2`): rows sharded over the devices of one
process with an all-reduce of the result block, and shard scopes landing on distinct devices.
Each case runs tests/multi_gpu_worker.py in a fresh process (the device set is bound once per
process)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _run(*args):
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "multi_gpu_worker.py"), *args],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("family", ["poisson_log", "normal_id"])
def test_in_process_row_sharding(family):
    _run("sharding", family)


def test_widened_families_row_sharded():
    _run("widened")


def test_shard_scopes_use_distinct_devices():
    _run("scopes")


def test_torchrun_two_ranks_bench_line():
    """bench.py under torchrun with 2 ranks: one JSON line, device and host-level paths agree."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    import json
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "bench.py"),
                        "--gpus", "2", "--steps", "5", "--warmup", "3", "--rows", "400000", "--cols", "64",
                        "--no-cpu-baseline"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["consistent_device_vs_host_api"] and line["value"] > 0
