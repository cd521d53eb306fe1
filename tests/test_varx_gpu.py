"""GPU parity tests of x as a parameter: d/dx = d * beta^T
(stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:125-128, poisson_log_glm_lpmf.hpp:119-122,
normal_id_glm_lpdf.hpp:121-124, neg_binomial_2_log_glm_lpmf.hpp:163-168), with x uploaded on every call."""
import numpy as np
import pytest

from tests.helpers import assert_close, check_glm_result, synth_glm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    from math_b200._lib import lib
    g.init(1)
    lib().stan_cuda_set_kernel(1)
    yield g
    g.cache_clear()


def _dx_ref(L, family, s, alpha=None):
    """d per row from the oracle (its per-row d/dalpha with a vector intercept), times beta."""
    n = s["x"].shape[0]
    al = np.full(n, s["alpha"][0]) if alpha is None else alpha
    o = L.oracle_glm(family, s["x"], s["y"], al, s["beta"], s["sigma"])
    return np.outer(o["dalpha"], s["beta"])


@pytest.mark.parametrize("family", ["bernoulli_logit", "poisson_log", "normal_id", "neg_binomial_2_log"])
@pytest.mark.parametrize("n,k", [(3, 2), (153, 71), (2000, 128), (1201, 600)])
def test_var_x_partials(glm, family, n, k):
    from oracle import loader as L
    s = synth_glm(family, n, k, seed=n + k)
    if k > 128:
        s["beta"] = s["beta"] * 0.3
    ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
    want = _dx_ref(L, family, s)
    uploads = []
    for call in range(2):
        st0 = glm.stats()
        got = glm.glm_eval(family, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"], need_dx=True)
        uploads.append(glm.stats()["bytes_uploaded"] - st0["bytes_uploaded"])
        check_glm_result(got, ref, "%s %dx%d var x" % (family, n, k))
        assert got["dx"].shape == (n, k)
        assert_close(got["dx"].ravel(), want.ravel(), "%s %dx%d dx" % (family, n, k))
    assert uploads[0] >= s["x"].nbytes and uploads[1] >= s["x"].nbytes   # x travels on every call
    # new values at the same address and shape: nothing stale
    s["x"][:] = 0.5 * s["x"]
    ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
    got = glm.glm_eval(family, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"], need_dx=True)
    check_glm_result(got, ref, "%s %dx%d var x, new values" % (family, n, k))
    assert_close(got["dx"].ravel(), _dx_ref(L, family, s).ravel(), "dx, new values")
    glm.cache_clear()


def test_var_x_with_vector_intercept_and_row_major(glm):
    from oracle import loader as L
    n, k = 5001, 40
    s = synth_glm("poisson_log", n, k, seed=77)
    alpha = 0.3 + 0.05 * np.random.default_rng(2).standard_normal(n)
    ref = L.oracle_glm("poisson_log", s["x"], s["y"], alpha, s["beta"])
    got = glm.glm_eval("poisson_log", s["y"], np.ascontiguousarray(s["x"]), alpha, s["beta"], need_dx=True)
    check_glm_result(got, ref, "var x + vector alpha, row-major x")
    assert_close(got["dx"].ravel(), _dx_ref(L, "poisson_log", s, alpha).ravel(), "dx with vector alpha")
    glm.cache_clear()
