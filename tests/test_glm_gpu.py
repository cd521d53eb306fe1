"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against
the golden vectors produced by the reference and against the plain-C oracle on seeded inputs.
Bar: 1e-10 relative on the log-density and on every gradient component (tests/helpers.rel_err)."""
import numpy as np
import pytest

from math_b200 import _abi
from tests.helpers import (RTOL, assert_close, cat_case, check_cat_result, check_glm_result, check_ord_result, glm_case,
                           normal_case, ord_case, synth_categorical, synth_glm, synth_ordered)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["warp", "cta"])
def glm(request):
    """The whole module runs twice: the categorical_logit cases run on the DMMA kernel in the first pass (the default)
    and on the vector kernel in the second; the second pass also selects the superseded CTA-synchronous GLM kernel
    (csrc/glm_fused.cu) where the library was built with it (STAN_CUDA_WITH_CTA_KERNEL=1; ignored by the product build)."""
    from math_b200 import glm as g
    from math_b200._lib import lib
    g.init(1)
    g.cache_clear()
    lib().stan_cuda_set_kernel(1 if request.param == "warp" else 0)
    lib().stan_cuda_set_categorical_kernel(1 if request.param == "warp" else 0)
    yield g
    g.cache_clear()
    lib().stan_cuda_set_kernel(1)
    lib().stan_cuda_set_categorical_kernel(1)


@pytest.fixture(scope="module")
def L():
    from oracle import loader
    return loader


def run_case(glm, c, **kw):
    return glm.glm_eval(c["family"], c["y"], c["x"], c["alpha"], c["beta"], c["sigma"],
                        propto=c.get("propto", False), need_dy=c.get("need_dy", False), **kw)


def test_library_reports_b200(glm):
    d = glm.describe()
    assert "sm_100" in d or "sm_103" in d, d


def test_golden_glm_cases(glm, golden):
    for name in golden["glm_cases"]:
        c = glm_case(golden, str(name))
        r = run_case(glm, c)
        assert r["rc"] == 0, name
        check_glm_result(r, c, "cuda vs golden " + str(name))


def test_golden_ordered_logistic_cases(glm, golden):
    for name in golden["ord_cases"]:
        c = ord_case(golden, str(name))
        flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
        r = glm.ordered_logistic_glm_eval(c["y"], c["x"], c["beta"], c["cuts"], flags)
        assert r["rc"] == 0, name
        check_ord_result(r, c, "cuda vs golden ord " + str(name))
        lp, g = glm.ordered_logistic_glm_lpmf(c["y"], c["x"], c["beta"], c["cuts"], propto=c["propto"])
        assert lp == r["lp"] and np.array_equal(g["cuts"], r["dcuts"]) and np.array_equal(g["beta"], r["dbeta"])


ORD_SHAPES = [(1, 1, 2), (5, 2, 3), (33, 7, 1), (1000, 16, 4), (4097, 33, 7), (100000, 32, 5), (30011, 128, 10),
              (9000, 256, 3), (2100, 500, 6), (50000, 8, 200), (3000, 40, 257)]


@pytest.mark.parametrize("n,k,c", ORD_SHAPES)
def test_ordered_logistic_against_oracle(glm, L, n, k, c):
    s = synth_ordered(n, k, c, seed=11 * n + k + c)
    flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
    ref = L.oracle_ordered_logistic_glm(s["x"], s["y"], s["beta"], s["cuts"])
    got = glm.ordered_logistic_glm_eval(s["y"], s["x"], s["beta"], s["cuts"], flags)
    assert got["rc"] == ref["rc"] == 0
    check_ord_result(got, ref, "ordered %dx%d C=%d" % (n, k, c))
    again = glm.ordered_logistic_glm_eval(s["y"], s["x"], s["beta"], s["cuts"], flags)
    assert again["lp"] == got["lp"] and np.array_equal(again["dcuts"], got["dcuts"])   # fixed summation order
    # value only / one edge constant: the other partials are left alone
    only = glm.ordered_logistic_glm_eval(s["y"], s["x"], s["beta"], s["cuts"], _abi.LP_MAIN | _abi.NEED_DBETA)
    assert only["lp"] == got["lp"] and np.array_equal(only["dbeta"], got["dbeta"])


def test_ordered_logistic_error_contract(glm):
    """ordered_logistic_glm_lpmf_test.cpp:410-470 (glm_errors)."""
    x = np.array([[1, 2], [3, 4], [5, 6], [7, 8], [9, 0]], float)
    y, beta, cuts = [1, 1, 2, 4, 4], [1.1, 0.4], [0.9, 1.1, 7]
    f = glm.ordered_logistic_glm_lpmf
    f(y, x, beta, cuts)
    for bad in ([1, 1, 2, 4, 4, 2],):
        with pytest.raises(ValueError) as e:
            f(bad, x, beta, cuts)
        assert not isinstance(e.value, glm.DomainError)
    with pytest.raises(ValueError) as e:
        f(y, x, [1.1, 0.4, 0], cuts)
    assert not isinstance(e.value, glm.DomainError)
    for bad_y in ([1, 1, 2, 4, 5], [1, 1, 2, 4, 0]):
        with pytest.raises(glm.DomainError):
            f(bad_y, x, beta, cuts)
    xv = x.copy()
    xv[3, 1] = np.inf
    with pytest.raises(glm.DomainError):
        f(y, xv, beta, cuts)
    with pytest.raises(glm.DomainError):
        f(y, x, [1.1, np.inf], cuts)
    for bad_cuts in ([0.9, 1.1, np.inf], [-np.inf, 1.1, 4], [0, 1.1, 0.5], [0, np.nan, 0.5]):
        with pytest.raises(glm.DomainError):
            f(y, x, beta, bad_cuts)
    # single class / no instances: 0 after the checks (:88-89)
    assert f([1, 1, 1, 1, 1], x, beta, [])[0] == 0.0
    assert f([], np.zeros((0, 2)), beta, cuts)[0] == 0.0
    # 300 cut-points (round 1 stopped at 256): parity with the oracle
    from oracle import loader as L
    so = synth_ordered(4001, 24, 301, seed=21)
    lp, g = f(so["y"], so["x"], so["beta"], so["cuts"])
    o = L.oracle_ordered_logistic_glm(so["x"], so["y"], so["beta"], so["cuts"])
    assert_close(lp, o["lp"], "ordered 301 classes lp")
    assert_close(g["beta"], o["dbeta"], "ordered 301 classes dbeta")
    assert_close(g["cuts"], o["dcuts"], "ordered 301 classes dcuts")
    # more cut-points than the fused kernel's class table holds: the general path, same answer as the reference's loop
    lp, g = f([1, 1100, 2, 1101, 570], x, beta, np.arange(1100.0) * 0.01)
    o = L.oracle_ordered_logistic_glm(x, [1, 1100, 2, 1101, 570], beta, np.arange(1100.0) * 0.01)
    assert_close(lp, o["lp"], "ordered 1101 classes lp")
    assert_close(g["cuts"], o["dcuts"], "ordered 1101 classes dcuts")


# beyond 1024 cut-points the evaluation runs as location / rows / gradient passes with the d/dcuts bins in global memory
# (glm_ordered_wide.cu); k > 512 adds the column blocks.  The reference has no limit (ordered_logistic_glm_lpmf.hpp:69-102).
@pytest.mark.parametrize("n,k,ncuts", [(20000, 24, 1024), (20000, 24, 1025), (100003, 16, 2000), (6000, 600, 1500),
                                        (300, 3, 5000), (50, 1100, 1100), (1, 1, 1025)])
def test_ordered_logistic_many_cut_points(glm, L, n, k, ncuts):
    rng = np.random.default_rng(n + k + ncuts)
    x = np.asfortranarray(rng.standard_normal((n, k)))
    beta = rng.standard_normal(k) / np.sqrt(k)
    gap = 8.0 / ncuts
    cuts = np.linspace(-4.0, 4.0, ncuts) + rng.uniform(0.0, 0.3 * gap, ncuts)   # strictly ascending, gaps >= 0.7 * gap
    y = (1 + np.searchsorted(cuts, x @ beta + rng.logistic(size=n))).astype(np.int32)
    y[0] = ncuts + 1 if n > 1 else y[0]   # the open-ended last class
    if n > 2:
        y[1] = 1
    beta = beta + 0.05 * rng.standard_normal(k)
    flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
    ref = L.oracle_ordered_logistic_glm(x, y, beta, cuts)
    got = glm.ordered_logistic_glm_eval(y, x, beta, cuts, flags)
    assert got["rc"] == ref["rc"] == 0
    check_ord_result(got, ref, "ordered %dx%d, %d cut-points" % (n, k, ncuts))
    again = glm.ordered_logistic_glm_eval(y, x, beta, cuts, flags)
    assert again["lp"] == got["lp"] and np.array_equal(again["dcuts"], got["dcuts"]) and np.array_equal(again["dbeta"], got["dbeta"])
    only = glm.ordered_logistic_glm_eval(y, x, beta, cuts, _abi.LP_MAIN | _abi.NEED_DBETA)
    assert only["lp"] == got["lp"] and np.array_equal(only["dbeta"], got["dbeta"])
    lp_only = glm.ordered_logistic_glm_eval(y, x, beta, cuts, _abi.LP_MAIN)
    assert lp_only["lp"] == got["lp"]


CAT_FLAGS = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA


def test_golden_categorical_logit_cases(glm, golden):
    for name in golden["cat_cases"]:
        c = cat_case(golden, str(name))
        r = glm.categorical_logit_glm_eval(c["y"], c["x"], c["alpha"], c["beta"], CAT_FLAGS)
        assert r["rc"] == 0, name
        check_cat_result(r, c, "cuda vs golden cat " + str(name))
        lp, g = glm.categorical_logit_glm_lpmf(c["y"], c["x"], c["alpha"], c["beta"], propto=c["propto"])
        assert lp == r["lp"] and np.array_equal(g["beta"], r["dbeta"]) and np.array_equal(g["alpha"], r["dalpha"])


# (2000, 500, 6) rather than (2100, 500, 6): at that seed one of the 3000 gradient components cancels to 3e-8 of the
# largest, and two CPU evaluations (plain-C oracle, compiled reference) already differ by 2e-10 on it
CAT_SHAPES = [(1, 1, 2), (5, 2, 3), (33, 7, 1), (1000, 16, 4), (4097, 33, 7), (100000, 32, 5), (30011, 128, 10),
              (9000, 256, 16), (2000, 500, 6), (50000, 8, 16), (3000, 40, 2), (70000, 32, 3)]


@pytest.mark.parametrize("n,k,c", CAT_SHAPES)
def test_categorical_logit_against_oracle(glm, L, n, k, c):
    s = synth_categorical(n, k, c, seed=13 * n + k + c)
    ref = L.oracle_categorical_logit_glm(s["x"], s["y"], s["alpha"], s["beta"])
    got = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    assert got["rc"] == ref["rc"] == 0
    check_cat_result(got, ref, "categorical %dx%d C=%d" % (n, k, c))
    again = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    assert again["lp"] == got["lp"] and np.array_equal(again["dbeta"], got["dbeta"])   # fixed summation order
    only = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], _abi.LP_MAIN)
    assert only["lp"] == got["lp"]


def test_categorical_logit_error_contract(glm):
    """categorical_logit_glm_lpmf_test.cpp:551-620 (glm_errors)."""
    x = np.array([[-12, 46], [-42, 24], [25, 27], [-14, -11], [5, 18]], float)
    beta = np.array([[0.3, 2, 0.4], [-0.1, -1.3, 1]], float)
    y, alpha = [1, 3, 1, 2, 2], [0.5, -2, 4]
    f = glm.categorical_logit_glm_lpmf
    f(y, x, alpha, beta)

    def size_error(*args):
        with pytest.raises(ValueError) as e:
            f(*args)
        assert not isinstance(e.value, glm.DomainError)

    size_error([1, 3, 1, 2, 2, 1], x, alpha, beta)
    size_error(y, np.vstack([x, [[1, 2]]]), alpha, beta)
    size_error(y, np.hstack([x, x[:, :1]]), alpha, beta)
    size_error(y, x, [0.5, -2, 4, 3], beta)
    size_error(y, x, alpha, np.vstack([beta, beta[:1]]))
    size_error(y, x, alpha, np.hstack([beta, beta[:, :1]]))
    for bad_y in ([1, 3, 1, 2, 4], [1, 3, 1, 2, 0]):
        with pytest.raises(glm.DomainError):
            f(bad_y, x, alpha, beta)
    for bad in (np.nan, np.inf):
        xv, bv, av = x.copy(), beta.copy(), np.array(alpha, float)
        xv[4, 1], bv[1, 2], av[2] = bad, bad, bad
        with pytest.raises(glm.DomainError):
            f(y, xv, alpha, beta)
        with pytest.raises(glm.DomainError):
            f(y, x, alpha, bv)
        with pytest.raises(glm.DomainError):
            f(y, x, av, beta)
    assert f([1] * 5, x, [0.3], beta[:, :1])[0] == 0.0      # single class (:66-68)
    # shapes beyond the fused kernels (round 1 raised here): the general three-pass path, parity with the oracle
    from oracle import loader as L
    for n, k, ncls in ((5, 2, 17), (4001, 40, 17), (3001, 24, 64), (2001, 600, 5), (1501, 300, 16)):
        sc = synth_categorical(n, k, ncls, seed=ncls + k)
        lp, g = f(sc["y"], sc["x"], sc["alpha"], sc["beta"])
        o = L.oracle_categorical_logit_glm(sc["x"], sc["y"], sc["alpha"], sc["beta"])
        what = "categorical %dx%d, %d classes" % (n, k, ncls)
        assert_close(lp, o["lp"], what + " lp")
        assert_close(g["alpha"], o["dalpha"], what + " dalpha")
        assert_close(np.asarray(g["beta"]).ravel(order="F"), np.asarray(o["dbeta"]).ravel(order="F"), what + " dbeta")
        lp2, g2 = f(sc["y"], sc["x"], sc["alpha"], sc["beta"])
        assert lp2 == lp and np.array_equal(g2["beta"], g["beta"])   # bit-reproducible


def test_golden_normal_cases(glm, golden):
    for name in golden["normal_cases"]:
        c = normal_case(golden, str(name))
        lp, g = glm.normal_lpdf(c["y"], c["mu"], c["sigma"], propto=c["propto"], need_dy=c["need_dy"])
        assert_close(lp, c["lp"], str(name) + " lp")
        assert_close(g["mu"], c["dmu"], str(name) + " dmu")
        assert_close(g["sigma"], c["dsigma"], str(name) + " dsigma")
        if c["need_dy"]:
            assert_close(g["y"], c["dy"], str(name) + " dy")
        if "known_answer" in c:
            assert abs(lp - c["known_answer"]) <= 1e-14 * max(1.0, abs(c["known_answer"]))


SHAPES = [(1, 1), (5, 2), (33, 7), (1000, 16), (4097, 33), (100000, 32), (30011, 128), (9000, 256), (2100, 500),
          (777, 100)]


@pytest.mark.parametrize("family", ["bernoulli_logit", "poisson_log", "normal_id", "neg_binomial_2_log"])
@pytest.mark.parametrize("n,k", SHAPES)
def test_against_oracle(glm, L, family, n, k):
    s = synth_glm(family, n, k, seed=7 * n + k)
    for propto in (False, True):
        ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"], propto=propto)
        got = glm.glm_eval(family, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"], propto=propto)
        assert got["rc"] == ref["rc"] == 0
        check_glm_result(got, ref, "%s %dx%d propto=%s" % (family, n, k, propto))


# 150001 rows: every CTA gets several tiles, so the two-tiles-per-iteration loop runs with per-row operands too
@pytest.mark.parametrize("n", [5003, 150001])
@pytest.mark.parametrize("family", ["bernoulli_logit", "poisson_log", "normal_id", "neg_binomial_2_log"])
def test_vector_alpha_sigma_and_dy(glm, L, family, n):
    k = 40
    s = synth_glm(family, n, k, seed=99)
    rng = np.random.default_rng(3)
    alpha = 0.3 + 0.1 * rng.standard_normal(n)
    sigma = 0.5 + rng.random(n) if family in ("normal_id", "neg_binomial_2_log") else None
    need_dy = family == "normal_id"
    ref = L.oracle_glm(family, s["x"], s["y"], alpha, s["beta"], sigma, need_dy=need_dy)
    got = glm.glm_eval(family, s["y"], s["x"], alpha, s["beta"], sigma, need_dy=need_dy)
    check_glm_result(got, ref, family + " vector alpha/sigma")
    if need_dy:
        assert_close(got["dy"], ref["dy"], "dy")


def test_row_major_x_and_leading_dimension(glm, L):
    s = synth_glm("poisson_log", 3001, 19, seed=11)
    ref = L.oracle_glm("poisson_log", s["x"], s["y"], s["alpha"], s["beta"])
    got = glm.glm_eval("poisson_log", s["y"], np.ascontiguousarray(s["x"]), s["alpha"], s["beta"])
    check_glm_result(got, ref, "row-major x")


def test_zero_instances_and_zero_attributes(glm, L):
    # zero instances -> 0 (bernoulli_logit_glm_lpmf_test.cpp:110-145)
    lp, _ = glm.poisson_log_glm_lpmf(np.zeros(0, np.int32), np.zeros((0, 3)), 0.3, [0.1, 0.2, 0.3])
    assert lp == 0.0
    # zero attributes -> eta = alpha (bernoulli_logit_glm_lpmf_test.cpp:147-175)
    for fam, y, sg in (("bernoulli_logit", [1, 0, 1], None), ("poisson_log", [2, 0, 5], None),
                       ("normal_id", [0.5, -1.0, 2.0], [1.7])):
        ref = L.oracle_glm(fam, np.zeros((3, 0)), y, 0.3, np.zeros(0), sg)
        got = glm.glm_eval(fam, y, np.zeros((3, 0)), 0.3, np.zeros(0), sg)
        assert_close(got["lp"], ref["lp"], fam + " K=0 lp")
        assert_close(got["dalpha"], ref["dalpha"], fam + " K=0 dalpha")


def test_cutoff_branches_bit_exact_selection(glm, L):
    """|ytheta| > 20 on both sides and near the cutoffs: same branch as the reference."""
    eta = np.array([-40.0, -20.000001, -20.0, -19.999999, 0.0, 19.999999, 20.0, 20.000001, 40.0, 700.0, -700.0])
    x = eta.reshape(-1, 1)
    for y in ([1] * len(eta), [0] * len(eta)):
        ref = L.oracle_glm("bernoulli_logit", x, y, np.zeros(len(eta)), [1.0])
        got = glm.glm_eval("bernoulli_logit", y, x, np.zeros(len(eta)), [1.0])
        assert_close(got["lp"], ref["lp"], "cutoff lp")
        assert_close(got["dalpha"], ref["dalpha"], "cutoff d per row", 1e-13)


def test_error_contract_lazy_checks(glm):
    x = np.ones((3, 2))
    with pytest.raises(glm.DomainError, match="Weight vector"):
        glm.poisson_log_glm_lpmf([1, 0, 1], x, 0.3, [0.1, np.inf])
    with pytest.raises(glm.DomainError, match="Intercept"):
        glm.bernoulli_logit_glm_lpmf([1, 0, 1], x, np.nan, [0.1, 0.2])
    xi = x.copy()
    xi[0, 1] = np.nan
    with pytest.raises(glm.DomainError, match="Matrix of independent variables"):
        glm.bernoulli_logit_glm_lpmf([1, 0, 1], xi, 0.3, [0.1, 0.2])
    with pytest.raises(glm.DomainError, match="Matrix of independent variables"):
        glm.normal_id_glm_lpdf([1.0, 0.0, 1.0], xi, 0.3, [0.1, 0.2], [1.0])
    # the reference's one-element check: NaN in a later row returns NaN without throwing
    xi = x.copy()
    xi[2, 0] = np.nan
    lp, _ = glm.bernoulli_logit_glm_lpmf([1, 0, 1], xi, 0.3, [0.1, 0.2])
    assert np.isnan(lp)


def test_neg_binomial_2_log_histogram_and_per_row_paths_agree(glm, L):
    """Scalar phi uses the histogram of y built at fill (sum lgamma / digamma over distinct counts);
    a constant vector phi forces the per-row special functions.  Same numbers, and the wrapper's
    error contract (neg_binomial_2_log_glm_lpmf.hpp:88-92,120)."""
    s = synth_glm("neg_binomial_2_log", 20011, 24, seed=31)
    ref = L.oracle_glm("neg_binomial_2_log", s["x"], s["y"], s["alpha"], s["beta"], [1.7])
    a = glm.glm_eval("neg_binomial_2_log", s["y"], s["x"], s["alpha"], s["beta"], [1.7])
    b = glm.glm_eval("neg_binomial_2_log", s["y"], s["x"], s["alpha"], s["beta"], np.full(20011, 1.7))
    check_glm_result(a, ref, "nb scalar phi")
    assert_close(b["lp"], ref["lp"], "nb vector phi lp")
    assert_close(b["dbeta"], ref["dbeta"], "nb vector phi dbeta")
    assert_close(np.sum(b["dsigma"]), ref["dsigma"][0], "nb sum of per-row dphi", 1e-9)
    lp, g = glm.neg_binomial_2_log_glm_lpmf(s["y"], s["x"], 0.3, s["beta"], 1.7)
    assert_close(lp, ref["lp"], "nb wrapper lp")
    assert_close(g["phi"], ref["dsigma"], "nb wrapper dphi")
    lp_p, _ = glm.neg_binomial_2_log_glm_lpmf(s["y"], s["x"], 0.3, s["beta"], 1.7, propto=True)
    ref_p = L.oracle_glm("neg_binomial_2_log", s["x"], s["y"], s["alpha"], s["beta"], [1.7], propto=True)
    assert_close(lp_p, ref_p["lp"], "nb propto lp")
    with pytest.raises(glm.DomainError, match="Precision"):
        glm.neg_binomial_2_log_glm_lpmf(s["y"], s["x"], 0.3, s["beta"], 0.0)
    with pytest.raises(glm.DomainError, match="Weight"):
        glm.neg_binomial_2_log_glm_lpmf(s["y"], s["x"], 0.3, s["beta"] * np.inf, 1.0)
    ybad = s["y"].copy()
    ybad[5] = -1
    with pytest.raises(glm.DomainError, match="Failures"):
        glm.neg_binomial_2_log_glm_lpmf(ybad, s["x"], 0.3, s["beta"], 1.0)
    xbad = s["x"].copy()
    xbad[100, 3] = np.inf
    with pytest.raises(glm.DomainError, match="independent"):
        glm.neg_binomial_2_log_glm_lpmf(s["y"], xbad, 0.3, s["beta"], 1.0)
    glm.cache_clear()


def test_outcome_domain_is_checked_on_the_device_at_fill(glm):
    """check_bounded(y, 0, 1) / check_nonnegative(y) (bernoulli/poisson :70): the range of y is
    computed once when y is cached; the call reports BAD_OUTCOME and the wrapper raises the
    reference's domain error -- no O(N) host loop per evaluation."""
    x = np.ones((3, 2))
    with pytest.raises(glm.DomainError, match="interval"):
        glm.bernoulli_logit_glm_lpmf([1, 0, 23], x, 0.3, [0.1, 0.2])
    with pytest.raises(glm.DomainError, match=">= 0"):
        glm.poisson_log_glm_lpmf([1, 0, -1], x, 0.3, [0.1, 0.2])
    # outcome errors win over parameter errors, as in the reference (eager before lazy)
    with pytest.raises(glm.DomainError, match=">= 0"):
        glm.poisson_log_glm_lpmf([1, 0, -1], x, 0.3, [0.1, np.inf])
    s = synth_glm("poisson_log", 50000, 8, seed=5)
    y = s["y"].copy()
    y[31337] = -4
    r = glm.glm_eval("poisson_log", y, s["x"], s["alpha"], s["beta"])
    assert r["rc"] == _abi.BAD_OUTCOME
    d = glm.GlmData("bernoulli_logit", s["x"], np.minimum(s["y"], 1))
    import ctypes
    lo, hi = ctypes.c_int64(), ctypes.c_int64()
    from math_b200._lib import lib
    assert lib().stan_cuda_data_y_range(d.h, ctypes.byref(lo), ctypes.byref(hi)) == 0
    assert (lo.value, hi.value) == (0, 1)
    d.close()
    glm.cache_clear()


def test_device_cache_hit_refill_and_invalidate(glm):
    s = synth_glm("poisson_log", 20000, 24, seed=1)
    glm.cache_clear()
    st0 = glm.stats()
    a = glm.glm_eval("poisson_log", s["y"], s["x"], s["alpha"], s["beta"])
    b = glm.glm_eval("poisson_log", a["_keep"][1], a["_keep"][0], s["alpha"], s["beta"])
    st1 = glm.stats()
    assert st1["cache_misses"] - st0["cache_misses"] == 1 and st1["cache_hits"] - st0["cache_hits"] == 1
    assert a["lp"] == b["lp"] and np.array_equal(a["dbeta"], b["dbeta"])  # deterministic, bit for bit
    # same address, new content -> fingerprint mismatch -> refill, new value
    x, y = a["_keep"]
    x[:] = x * 0.5
    c = glm.glm_eval("poisson_log", y, x, s["alpha"], s["beta"])
    st2 = glm.stats()
    assert st2["cache_misses"] - st1["cache_misses"] == 1 and c["lp"] != a["lp"]
    assert glm.cache_invalidate(x) == 1
    glm.cache_clear()


def test_repeat_evaluations_are_bit_identical(glm):
    s = synth_glm("bernoulli_logit", 250000, 32, seed=21)
    r = [glm.glm_eval("bernoulli_logit", s["y"], s["x"], s["alpha"], s["beta"]) for _ in range(3)]
    assert r[0]["lp"] == r[1]["lp"] == r[2]["lp"]
    assert np.array_equal(r[0]["dbeta"], r[1]["dbeta"]) and np.array_equal(r[0]["dbeta"], r[2]["dbeta"])


def test_large_shape_row_additivity(glm, L):
    """Size-independent property at a large shape: value and gradient are sums over rows, so the
    full evaluation equals the sum of the evaluations of two row blocks; a 4096-row sample is
    also checked against the oracle."""
    n, k = 1_000_000, 128
    s = synth_glm("poisson_log", n, k, seed=1003)
    full = glm.glm_eval("poisson_log", s["y"], s["x"], s["alpha"], s["beta"])
    h = n // 2 + 12345
    xa, xb = np.asfortranarray(s["x"][:h]), np.asfortranarray(s["x"][h:])
    a = glm.glm_eval("poisson_log", s["y"][:h], xa, s["alpha"], s["beta"])
    b = glm.glm_eval("poisson_log", s["y"][h:], xb, s["alpha"], s["beta"])
    assert_close(a["lp"] + b["lp"], full["lp"], "additivity lp", 1e-12)
    assert_close(a["dbeta"] + b["dbeta"], full["dbeta"], "additivity dbeta", 1e-11)
    assert_close(a["dalpha"] + b["dalpha"], full["dalpha"], "additivity dalpha", 1e-11)
    xs = np.asfortranarray(s["x"][:4096])
    ref = L.oracle_glm("poisson_log", xs, s["y"][:4096], s["alpha"], s["beta"])
    got = glm.glm_eval("poisson_log", s["y"][:4096], xs, s["alpha"], s["beta"])
    check_glm_result(got, ref, "sample of the large shape")
    glm.cache_clear()


def _raw_eval(glm, fam, x, ldx, n, k, y, alpha, beta):
    """stan_cuda_glm_eval on a row block of a column-major matrix, zero-copy (x points at the block's first
    row, ldx is the full matrix's row count)."""
    import ctypes as C
    from math_b200._lib import check, lib
    a = _abi.GlmArgs()
    lp, dbeta, dalpha = np.zeros(1), np.zeros(k), np.zeros(1)
    al, be = np.array([alpha], float), np.ascontiguousarray(beta, dtype=np.float64)
    a.family, a.n, a.k = fam, n, k
    a.flags = glm.glm_flags(fam, False)
    a.x, a.ldx, a.x_row_major = x, ldx, 0
    a.y, a.y_is_double = y, 0
    a.alpha, a.alpha_len, a.beta = al.ctypes.data, 1, be.ctypes.data
    a.lp, a.dbeta, a.dalpha = lp.ctypes.data, dbeta.ctypes.data, dalpha.ctypes.data
    assert check(lib().stan_cuda_glm_eval(C.byref(a))) == 0
    return float(lp[0]), dbeta, float(dalpha[0])


def test_full_size_c3_row_additivity_and_sample_parity(glm, L, request):
    """BASELINE configs[2] at its full size (N = 1e7, K = 128, 10.28 GB): value and gradient are sums over rows,
    so the evaluation of the whole matrix must equal the sum over three uneven row blocks (passed zero-copy
    through ldx); rows sampled across the whole range are checked against the oracle."""
    import torch
    from bench import synth_shard
    if request.node.callspec.params.get("glm") != "warp":
        pytest.skip("full size once, on the default kernel")
    if torch.cuda.mem_get_info()[0] < 40e9:
        pytest.skip("needs 40 GB of free device memory")
    n, k = 10_000_000, 128
    x, y, alpha, beta, _ = synth_shard("poisson_log", n, k, seed=1003)
    full = _raw_eval(glm, _abi.POISSON_LOG, x.ctypes.data, n, n, k, y.ctypes.data, alpha[0], beta)
    from tests import config_cases as cc
    golden, man = cc.load()                      # outputs of the compiled reference on these same inputs
    reproduced = cc.inputs_reproduce(man, "c3", x, y)
    if reproduced:
        assert_close(full[0], golden["c3/lp"][0], "C3 lp vs compiled reference")
        assert_close(full[1], golden["c3/dbeta"], "C3 dbeta vs compiled reference")
        assert_close(full[2], golden["c3/dalpha"][0], "C3 dalpha vs compiled reference")
    cuts = [0, 3_333_337, 3_333_337 + 4_096_001, n]
    parts = [_raw_eval(glm, _abi.POISSON_LOG, x.ctypes.data + 8 * r0, n, r1 - r0, k, y.ctypes.data + 4 * r0, alpha[0], beta)
             for r0, r1 in zip(cuts[:-1], cuts[1:])]
    assert_close(sum(p[0] for p in parts), full[0], "C3 additivity lp", 1e-12)
    assert_close(sum(p[1] for p in parts), full[1], "C3 additivity dbeta", 1e-11)
    assert_close(sum(p[2] for p in parts), full[2], "C3 additivity dalpha", 1e-11)
    idx = np.random.default_rng(5).choice(n, 20000, replace=False)
    xs, ys = np.asfortranarray(x[idx]), np.ascontiguousarray(y[idx])
    ref = L.oracle_glm("poisson_log", xs, ys, alpha, beta)
    got = glm.glm_eval("poisson_log", ys, xs, alpha, beta)
    check_glm_result(got, ref, "rows sampled from C3")
    glm.cache_clear()
    if not reproduced:   # the additivity and sampled-row checks above ran; say loudly that the 1e-10 check did not
        pytest.skip("this host does not regenerate the stored C3 inputs bit for bit: comparison with the stored "
                    "outputs of the compiled reference NOT run (additivity and sampled-row parity passed)")


def test_device_level_launch_matches_host_level(glm):
    import torch
    s = synth_glm("normal_id", 60000, 64, seed=8)
    host = glm.glm_eval("normal_id", s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
    d = glm.GlmData("normal_id", s["x"], s["y"])
    params = np.zeros(d.params_len)
    params[:64], params[d.kpad], params[d.kpad + 1] = s["beta"], s["alpha"][0], s["sigma"][0]
    dp = torch.from_numpy(params).cuda()
    dr = torch.zeros(d.result_len, dtype=torch.float64, device="cuda")
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        d.launch(glm.glm_flags(_abi.NORMAL_ID, False), dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    res = dr.cpu().numpy()
    assert res[0] == host["lp"] and np.array_equal(res[4:4 + 64], host["dbeta"])
    assert res[1] == host["dalpha"][0] and res[2] == host["dsigma"][0]
    d.close()


def test_normal_lpdf_against_oracle_large(glm, L):
    rng = np.random.default_rng(5)
    n = 1_000_000
    y, mu = rng.standard_normal(n), 0.1 * rng.standard_normal(n)
    ref = L.oracle_normal_lpdf(y, mu, [1.3])
    lp, g = glm.normal_lpdf(y, mu, [1.3])
    assert_close(lp, ref["lp"], "normal_lpdf lp")
    assert_close(g["mu"], ref["dmu"], "normal_lpdf dmu")
    assert_close(g["sigma"], ref["dsigma"], "normal_lpdf dsigma")
    glm.cache_clear()


def test_concurrent_host_threads(glm, L):
    """TBB-style concurrency: several host threads evaluate different GLMs at once."""
    import threading
    cases = [synth_glm("poisson_log", 20000 + 1000 * i, 16 + i, seed=50 + i) for i in range(6)]
    refs = [L.oracle_glm("poisson_log", c["x"], c["y"], c["alpha"], c["beta"]) for c in cases]
    out, errs = [None] * len(cases), []

    def work(i):
        try:
            for _ in range(5):
                out[i] = glm.glm_eval("poisson_log", cases[i]["y"], cases[i]["x"], cases[i]["alpha"], cases[i]["beta"])
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(cases))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    for got, ref in zip(out, refs):
        check_glm_result(got, ref, "threaded")
    glm.cache_clear()


def test_shard_scope_keys_by_call_id_and_job(glm, L):
    """Inside a map_rect shard scope the data identity is (call_id, job): a functor that rebuilds
    X at a new address on every call still hits the cache."""
    s = synth_glm("bernoulli_logit", 10000, 32, seed=77)
    ref = L.oracle_glm("bernoulli_logit", s["x"], s["y"], s["alpha"], s["beta"])
    st0 = glm.stats()
    for rep in range(3):
        x_tmp = np.array(s["x"], order="F", copy=True)  # fresh temporary, like the functor's Eigen matrix
        with glm.shard_scope(call_id=5, job=2, device=0):
            got = glm.glm_eval("bernoulli_logit", s["y"].copy(), x_tmp, s["alpha"], s["beta"])
        check_glm_result(got, ref, "shard scope rep %d" % rep)
    st1 = glm.stats()
    assert st1["cache_misses"] - st0["cache_misses"] == 1 and st1["cache_hits"] - st0["cache_hits"] == 2
    glm.cache_clear()


def test_prepared_call_matches_glm_eval(glm, L):
    """math_b200.glm.GlmCall (argument block built once; what bench.py's end-to-end loop uses) against glm_eval and the oracle."""
    for family in ("poisson_log", "normal_id"):
        s = synth_glm(family, 30_001, 48, seed=31)
        call = glm.GlmCall(family, s["y"], s["x"])
        sg = None if s["sigma"] is None else float(s["sigma"][0])
        for scale in (1.0, 1.05):
            assert call.eval(s["alpha"][0], s["beta"] * scale, sg) == 0
            ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"] * scale, s["sigma"])
            assert_close(call.lp[0], ref["lp"], family + " prepared call lp")
            assert_close(call.dbeta[:48], ref["dbeta"], family + " prepared call dbeta")
            assert_close(call.dalpha[0], ref["dalpha"][0], family + " prepared call dalpha")
            same = glm.glm_eval(family, call.y, call.x, s["alpha"], s["beta"] * scale, s["sigma"])
            assert same["lp"] == call.lp[0] and np.array_equal(same["dbeta"], call.dbeta[:48])
    glm.cache_clear()
