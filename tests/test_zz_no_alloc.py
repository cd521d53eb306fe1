"""GPU test of stan_cuda_glm_launch_noalloc (include/stan_cuda.h): the fused GLM evaluation with its reduction
scratch in a file-scope __device__ array of the library instead of a cudaMalloc'ed workspace.  Called through the
C ABI, twice in a row per family, and compared with the existing entry point stan_cuda_glm_launch on the same device
buffers (which must agree bit for bit: same kernel, same grid, same summation order) and with the oracle at the
suite's tolerance of 1e-10.  The file sorts last on purpose: a failure here
cannot hide another test.  It runs alone in a fresh process as well as behind the other tests (it only needs the
library's one-time device binding, which glm.init(1) repeats harmlessly)."""
import ctypes as C

import numpy as np
import pytest

from tests.helpers import assert_close, synth_glm

pytestmark = pytest.mark.gpu


def _launch_both(glm, lib, check, torch, family, n, k, seed):
    from math_b200 import _abi
    s = synth_glm(family, n, k, seed=seed)
    fam = s["family"]
    data = glm.GlmData(fam, s["x"], s["y"])
    flags = glm.glm_flags(fam, False)
    stream = torch.cuda.Stream()
    out = []
    try:
        for call in range(2):                        # two calls in a row: the shared scratch keeps the first call's contents
            beta = s["beta"] * (1.0 + 0.05 * call)
            params = np.zeros(data.params_len)
            params[:k], params[data.kpad] = beta, s["alpha"][0]
            params[data.kpad + 1] = s["sigma"][0] if s["sigma"] is not None else 1.0
            d_params = torch.from_numpy(params).cuda()
            d_old = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
            d_new = torch.full((data.result_len,), float("nan"), dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            data.launch(flags, d_params.data_ptr(), d_old.data_ptr(), stream.cuda_stream)
            check(lib.stan_cuda_glm_launch_noalloc(data.h, flags, d_params.data_ptr(), d_new.data_ptr(), stream.cuda_stream))
            stream.synchronize()
            out.append((d_old.cpu().numpy(), d_new.cpu().numpy(), beta))
    finally:
        data.close()
    return s, out


@pytest.mark.parametrize("family,n,k", [("poisson_log", 30_011, 128), ("bernoulli_logit", 150_001, 32),
                                        ("normal_id", 20_003, 200), ("poisson_log", 257, 5)])
def test_noalloc_launch_matches_the_existing_entry_point(family, n, k):
    import torch
    from math_b200 import glm
    from math_b200._lib import check, lib
    from oracle import loader as L
    glm.init(1)
    lib().stan_cuda_set_kernel(1)   # the default (warp) tile layout, whatever an earlier test module selected
    s, calls = _launch_both(glm, lib(), check, torch, family, n, k, seed=41)
    for i, (old, new, beta) in enumerate(calls):
        what = "%s %dx%d call %d" % (family, n, k, i)
        assert_close(new[0], old[0], what + " lp vs stan_cuda_glm_launch")
        assert_close(new[4:4 + k], old[4:4 + k], what + " dbeta vs stan_cuda_glm_launch")
        assert_close(new[1], old[1], what + " dalpha vs stan_cuda_glm_launch")
        assert np.array_equal(new[:4 + k], old[:4 + k]), what + ": same kernel and grid, must agree bit for bit"
        ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], beta, s["sigma"])
        assert_close(new[0], ref["lp"], what + " lp vs oracle")
        assert_close(new[4:4 + k], ref["dbeta"], what + " dbeta vs oracle")
        assert_close(new[1], ref["dalpha"][0], what + " dalpha vs oracle")
        if family == "normal_id":
            assert_close(new[2], ref["dsigma"][0], what + " dsigma vs oracle")


def test_noalloc_launch_rejects_what_it_cannot_serve():
    import torch  # noqa: F401
    from math_b200 import _abi, glm
    from math_b200._lib import lib
    glm.init(1)
    lib().stan_cuda_set_kernel(1)
    s = synth_glm("neg_binomial_2_log", 5000, 16, seed=3)
    data = glm.GlmData(s["family"], s["x"], s["y"])
    try:
        rc = lib().stan_cuda_glm_launch_noalloc(data.h, glm.glm_flags(s["family"], False), 8, 8, None)
        assert rc == _abi.ERR_ARG and b"only" in lib().stan_cuda_last_error()
    finally:
        data.close()
    assert lib().stan_cuda_glm_launch_noalloc(None, 0, None, None, None) == _abi.ERR_ARG
    s = synth_glm("poisson_log", 300, 600, seed=3)      # wider than one warp tile: several launches, not served here
    data = glm.GlmData(s["family"], s["x"], s["y"])
    try:
        rc = lib().stan_cuda_glm_launch_noalloc(data.h, glm.glm_flags(s["family"], False), 8, 8, None)
        assert rc == _abi.ERR_ARG and b"512" in lib().stan_cuda_last_error()
    finally:
        data.close()
