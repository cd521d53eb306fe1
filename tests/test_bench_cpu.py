"""CPU tests of bench.py's host side: the reference arm (the reference's CPU path timed by `--impl reference`)
must print one JSON line and EXIT 0 -- round 1's arm printed its line and then died in interpreter teardown
(SIGSEGV), which voided the driver's ratio -- and both arms must describe the workload with the same `config`."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], stdout=subprocess.PIPE,
                          stderr=subprocess.PIPE, text=True, timeout=600, env=e)


def test_reference_arm_exits_zero_with_one_line():
    r = _bench("--impl", "reference", "--rows", "64000", "--cols", "16", "--steps", "3", "--warmup", "3")
    assert r.returncode == 0, (r.returncode, r.stderr[-2000:])
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["value"] > 0 and line["steps_run"] == 3
    assert line["cpu_baseline"]["sample"].startswith("each step = all 64000 rows")
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["gpu_launches"] == 0
    assert line["config"]["n"] == 64000 and line["config"]["k"] == 16


def test_reference_arm_other_ranks_exit_zero_without_work():
    r = _bench("--impl", "reference", "--gpus", "2", "--rows", "64000", "--cols", "16", env={"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_both_arms_print_the_same_config():
    sys.path.insert(0, ROOT)
    import bench
    fam, n, k, desc = bench.WORKLOADS["c3"]
    for world in (1, 2, 8):
        a = bench.workload_config(desc, fam, n, k, world)
        assert a == bench.workload_config(desc, fam, n, k, world)
        assert a["rows_per_gpu"] * world >= n and a["workload"] == desc


@pytest.mark.parametrize("n,cores", [(10_000_000, 32), (10_000_000, 48), (10_000_000, 224), (20_000_000, 64), (64000, 8)])
def test_shard_count_covers_all_rows(n, cores):
    sys.path.insert(0, ROOT)
    import bench
    d = bench.shard_count(n, cores)
    assert cores <= d <= 4 * cores and n % d == 0


def test_summarize_percentiles():
    sys.path.insert(0, ROOT)
    import bench
    s = bench.summarize([5.0, 1.0, 2.0, 3.0, 4.0])
    assert s["min"] == 1.0 and s["max"] == 5.0 and s["median"] == 3.0 and s["p10"] == 1.0 and s["p90"] == 5.0
