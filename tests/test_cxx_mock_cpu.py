"""CPU tests of the C++ HOST LAYER against a test double of the device library.

The product library has no CPU path, so without a GPU nothing behind the C ABI can run.  The host layer in front
of it can: the include-path overlay and its routing, argument marshalling, the mapping of ABI return codes onto
the reference's exceptions, the tape node built from the returned partials, map_rect's shard scopes.  These tests
run the SAME gtest binaries as tests/test_cxx_gpu.py (the reference's own unit tests compiled against the overlay,
plus test/unit/math/cuda) with tests/mock_device/mock_stan_cuda.c in front of them on LD_LIBRARY_PATH: that
library forwards the C ABI to the plain-C oracle.  It is test infrastructure (see its header); the GPU tests run
the binaries against the real library, found through their RUNPATH."""
import os
import subprocess

import pytest

from tests.test_cxx_gpu import BINARIES, CXX_DIR, ROOT

MOCK_DIR = os.path.join(ROOT, "build", "mock")


@pytest.fixture(scope="module")
def mock_env():
    os.makedirs(MOCK_DIR, exist_ok=True)
    lib = os.path.join(MOCK_DIR, "libstanmath_cuda.so")
    subprocess.run(["gcc", "-O2", "-std=gnu11", "-fPIC", "-shared", "-Wall", "-Wl,-soname,libstanmath_cuda.so", "-o", lib,
                    os.path.join(ROOT, "tests", "mock_device", "mock_stan_cuda.c"),
                    os.path.join(ROOT, "oracle", "glm_oracle.c"), "-lm"], check=True)
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = MOCK_DIR + os.pathsep + os.path.join(ROOT, "oracle", "_ref", "tbb")
    return env


def test_mock_exports_what_the_binaries_import(mock_env):
    """every stan_cuda_* symbol a C++ test binary imports is one the ABI header declares and the mock defines"""
    built = [os.path.join(CXX_DIR, b) for b in BINARIES if os.path.exists(os.path.join(CXX_DIR, b))]
    if not built:
        pytest.skip("C++ test binaries not built (test/Makefile needs /root/reference)")
    wanted = set()
    for exe in built:
        out = subprocess.run(["nm", "-D", "--undefined-only", exe], stdout=subprocess.PIPE, text=True).stdout
        wanted |= {line.split()[-1] for line in out.splitlines() if "stan_cuda_" in line}
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(MOCK_DIR, "libstanmath_cuda.so")],
                         stdout=subprocess.PIPE, text=True).stdout
    have = {line.split()[-1] for line in out.splitlines() if "stan_cuda_" in line}
    header = open(os.path.join(ROOT, "include", "stan_cuda.h")).read()
    assert wanted and wanted <= have, wanted - have
    assert all(name + "(" in header for name in wanted)


@pytest.mark.parametrize("name", BINARIES)
def test_cxx_binary_against_mock_device(mock_env, name):
    exe = os.path.join(CXX_DIR, name)
    if not os.path.exists(exe):
        pytest.skip("%s not built (test/Makefile needs /root/reference)" % name)
    ldd = subprocess.run(["ldd", exe], stdout=subprocess.PIPE, text=True, env=mock_env).stdout
    if "libstanmath_cuda.so" in ldd:   # (tests whose argument kinds never route to the device do not link it)
        assert os.path.join("build", "mock", "libstanmath_cuda.so") in ldd, ldd
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, env=mock_env)
    tail = "\n".join(r.stdout.splitlines()[-40:])
    assert r.returncode == 0, tail
    assert "PASSED" in r.stdout and "FAILED" not in r.stdout, tail


def test_arena_node_equals_operands_and_partials(mock_env):
    """bench/host_phases.cpp evaluates normal_lpdf with a var vector both ways -- the reference's operands_and_partials
    edges + build(), and the in-place arena node of stan/math/cuda/arena_partials.hpp -- and prints the largest
    difference over the value and all N + 1 adjoints: the two tape nodes must be identical (N above and below the
    threshold of the parallel gather)."""
    import json
    exe = os.path.join(CXX_DIR, "host_phases")
    if not os.path.exists(exe):
        pytest.skip("host_phases not built (test/Makefile needs /root/reference)")
    for n in ("1000", "70001", "300001"):   # the last one is above the threshold of the parallel chain()
        r = subprocess.run([exe, n, "2"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, env=mock_env)
        assert r.returncode == 0, r.stdout
        out = json.loads(r.stdout)
        assert out["max_abs_difference_of_results"] == 0.0, r.stdout
        assert "mock device" in out["device"]


def test_map_rect_dispatch_against_mock_device(mock_env):
    """test/unit/math/cuda/map_rect_dispatch_mock_test.cpp: job i of a map_rect call evaluates inside the shard scope
    (call_id, i, i mod G); the scope is closed when a job throws; evaluations outside map_rect carry no scope.  The
    mock reports 3 devices and records the scope of every evaluation."""
    exe = os.path.join(CXX_DIR, "cuda_map_rect_dispatch_mock_test")
    if not os.path.exists(exe):
        pytest.skip("not built (test/Makefile needs /root/reference)")
    env = dict(mock_env, STAN_CUDA_MOCK_DEVICES="3", STAN_NUM_THREADS="4")
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "PASSED  ] 4 tests" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("config", ["c1", "c2", "c3", "c4", "c5", "d1", "d2", "v2"])
def test_stan_bench_harness_against_mock_device(mock_env, config):
    """bench/stan_bench.cpp, every config at a small size: the harness itself (problem generation, the map_rect
    functors, the reading of adjoints) agrees with the reference bodies it times beside the device path."""
    import json
    exe = os.path.join(CXX_DIR, "stan_bench")
    if not os.path.exists(exe):
        pytest.skip("stan_bench not built (test/Makefile needs /root/reference)")
    env = dict(mock_env, STAN_NUM_THREADS="4")
    n = "2000" if config == "c5" else "20000"
    r = subprocess.run([exe, "--config", config, "--n", n, "--reps", "2", "--cpu-reps", "1"], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    errs = [v for k, v in out.items() if k.startswith("parity_rel_err")]
    assert errs and max(errs) <= 1e-10, out
