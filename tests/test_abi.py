"""CPU tests (no GPU): the C-ABI library builds, loads and exports exactly what include/stan_cuda.h
declares; argument validation that needs no device; loud failure without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from math_b200 import _abi
from math_b200.build import LIB, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "stan_cuda.h")).read()
    return sorted(set(re.findall(r"STAN_CUDA_API\s+[\w\s\*]+?\b(stan_cuda_\w+)\s*\(", text)))


@pytest.fixture(scope="module")
def cdll():
    build()
    return ctypes.CDLL(LIB)


def test_header_and_ctypes_mirror_agree():
    assert declared_functions() == sorted(_abi.PROTOTYPES)


def test_library_exports_every_declared_symbol(cdll):
    for name in declared_functions():
        assert hasattr(cdll, name), name
    _abi.bind(cdll)


def test_struct_layout_matches_header(cdll):
    # sizes implied by the header's field order on LP64
    assert ctypes.sizeof(_abi.GlmArgs) == 4 + 4 + 8 + 8 + 8 + 8 + 4 + 4 + 8 * 16 + 4 + 4 + 8 + 8
    assert ctypes.sizeof(_abi.NormalArgs) == 4 + 4 + 8 * 11
    assert ctypes.sizeof(_abi.StudentTArgs) == 4 + 4 + 8 * 14
    assert ctypes.sizeof(_abi.Stats) == 8 * 9


def test_argument_validation_needs_no_device(cdll):
    lib = _abi.bind(cdll)
    assert lib.stan_cuda_glm_eval(None) == _abi.ERR_ARG
    a = _abi.GlmArgs()
    a.family, a.n, a.k = 7, 3, 2
    assert lib.stan_cuda_glm_eval(ctypes.byref(a)) == _abi.ERR_ARG
    assert b"family" in lib.stan_cuda_last_error()
    # n == 0 returns 0 like size_zero(y) in the reference, before any device work
    lp = np.full(1, 5.0)
    al = np.zeros(1)
    a = _abi.GlmArgs()
    a.family, a.n, a.k, a.alpha, a.alpha_len, a.lp = 1, 0, 0, al.ctypes.data, 1, lp.ctypes.data
    assert lib.stan_cuda_glm_eval(ctypes.byref(a)) == 0 and lp[0] == 0.0


def test_fails_loudly_without_gpu(cdll):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _abi.bind(cdll)
    assert lib.stan_cuda_init(0) == _abi.ERR_CUDA
    assert b"no CPU fallback" in lib.stan_cuda_last_error()
    from math_b200 import glm
    with pytest.raises(glm.StanCudaError):
        glm.poisson_log_glm_lpmf([1, 2], np.ones((2, 2)), 0.1, [0.1, 0.2])


def test_host_wrapper_error_contract_without_device():
    """Eager parameter and size checks run before the ABI call, exactly where the reference has
    them (the outcome range check is answered by the device: tests/test_glm_gpu.py)."""
    from math_b200 import glm
    x = np.ones((3, 2))
    with pytest.raises(glm.DomainError):
        glm.normal_id_glm_lpdf([1.0, 0.0, 2.0], x, 0.3, [0.1, 0.2], [0.0])
    with pytest.raises(ValueError):
        glm.poisson_log_glm_lpmf([1, 0], x, 0.3, [0.1, 0.2])
    with pytest.raises(ValueError):
        glm.poisson_log_glm_lpmf([1, 0, 2], x, 0.3, [0.1, 0.2, 0.3])
    with pytest.raises(ValueError):
        glm.poisson_log_glm_lpmf([1, 0, 2], x, [0.3, 0.1], [0.1, 0.2])
    assert glm.poisson_log_glm_lpmf(np.zeros(0, int), np.zeros((0, 2)), 0.3, [0.1, 0.2])[0] == 0.0
    with pytest.raises(glm.DomainError):
        glm.normal_lpdf([np.nan], [0.0], [1.0])
    with pytest.raises(glm.DomainError):
        glm.normal_lpdf([0.0], [np.inf], [1.0])
    with pytest.raises(glm.DomainError):
        glm.normal_lpdf([0.0], [0.0], [0.0])
    assert glm.normal_lpdf([], [0.0], [1.0])[0] == 0.0


def test_header_is_plain_c():
    """The drop-in boundary is a C ABI: include/stan_cuda.h must compile as C99 without any C++ or CUDA header."""
    import shutil
    import subprocess
    import tempfile
    cc = shutil.which("gcc") or shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        with open(src, "w") as f:
            f.write('#include "stan_cuda.h"\nint main(void) { stan_cuda_glm_args a; stan_cuda_density_args b; '
                    '(void)a; (void)b; return sizeof(stan_cuda_stats_t) == 72 ? 0 : 1; }\n')
        r = subprocess.run([cc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", src,
                            "-o", os.path.join(d, "t.o")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, r.stdout


def test_cxx_routing_table_compiles_and_runs_without_a_gpu():
    """test/unit/math/cuda/routing_static_test.cpp: static_asserts over the eligibility traits (which instantiations
    of the reference's templates go to the device) plus two calls that must stay on the reference loop."""
    import subprocess
    exe = os.path.join(ROOT, "build", "cxx", "cuda_routing_static_test")
    if not os.path.exists(exe):
        pytest.skip("not built (test/Makefile needs /root/reference)")
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert r.returncode == 0 and "PASSED" in r.stdout, r.stdout


def test_product_never_references_the_oracle_or_the_test_double():
    """oracle/ and tests/mock_device/ are test infrastructure: nothing the product is made of includes, links, loads or
    names a path into them (comments that cite a test file by name are fine), and the built library neither needs nor
    defines an oracle symbol."""
    import re
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pat = re.compile(r"(#\s*include[^\n]*(oracle|mock_device))|(oracle/|liboracle|libstanref|mock_device|build/mock)")
    offenders = []
    for top in ("math_b200", "stan", "overlay", "include"):
        for d, _, files in os.walk(os.path.join(root, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                    text = open(os.path.join(d, f), errors="replace").read()
                    if pat.search(text):
                        offenders.append(os.path.relpath(os.path.join(d, f), root))
    assert not offenders, offenders
    lib = os.path.join(root, "math_b200", "lib", "libstanmath_cuda.so")
    if os.path.exists(lib):
        out = subprocess.run(["nm", "-D", lib], stdout=subprocess.PIPE, text=True).stdout
        assert "oracle_" not in out and "stanref_" not in out and "stan_cuda_mock" not in out
        needed = subprocess.run(["readelf", "-d", lib], stdout=subprocess.PIPE, text=True).stdout
        assert "liboracle" not in needed and "libstanref" not in needed
