"""GPU parity tests of the vectorised elementwise densities (stan_cuda_density_eval, csrc/density.cu):
bernoulli_logit_lpmf / poisson_log_lpmf / poisson_lpmf / bernoulli_lpmf against the golden vectors of the compiled reference (with the literal
known answers of the reference's own tests), the plain-C oracle on seeded inputs, and the reference's error
contract.  Bar: 1e-10 relative on the log-density and on every partial (tests/helpers.check_dens_result)."""
import numpy as np
import pytest

from math_b200 import _abi
from tests.helpers import assert_close, check_dens_result, dens_case

pytestmark = pytest.mark.gpu

FLAGS = _abi.LP_CONST | _abi.LP_MAIN | _abi.NEED_DALPHA


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    g.init(1)
    g.cache_clear()
    yield g
    g.cache_clear()


@pytest.fixture(scope="module")
def L():
    from oracle import loader
    return loader


def test_golden_density_cases(glm, L, golden):
    for name in golden["dens_cases"]:
        c = dens_case(golden, str(name))
        f = glm.bernoulli_logit_lpmf if c["density"] == _abi.DENSITY_BERNOULLI_LOGIT else glm.poisson_log_lpmf
        lp, g = f(c["y"], c["theta"], propto=c["propto"])
        check_dens_result({"lp": lp, "dtheta": g["theta"]}, c, "cuda vs golden dens " + str(name))
        if "known_answer" in c:
            assert abs(lp - c["known_answer"]) <= 1e-12 * max(1.0, abs(c["known_answer"])), name
        # same numbers through the raw ABI call
        r = glm.density_eval(c["density"], c["y"], c["theta"], L.density_flags(c["propto"]))
        assert r["rc"] == 0 and r["lp"] == lp and np.array_equal(r["dtheta"], g["theta"])


def test_golden_poisson_and_bernoulli_cases(glm, L, golden_d2):
    by_id = {_abi.DENSITY_POISSON: glm.poisson_lpmf, _abi.DENSITY_BERNOULLI: glm.bernoulli_lpmf}
    for name in golden_d2["dens_cases"]:
        c = dens_case(golden_d2, str(name))
        lp, g = by_id[c["density"]](c["y"], c["theta"], propto=c["propto"])
        check_dens_result({"lp": lp, "dtheta": g["theta"]}, c, "cuda vs golden dens " + str(name))
        if "known_answer" in c:
            assert abs(lp - c["known_answer"]) <= 1e-12 * max(1.0, abs(c["known_answer"])), name
        r = glm.density_eval(c["density"], c["y"], c["theta"], L.density_flags(c["propto"]))
        assert r["rc"] == 0 and r["lp"] == lp and np.array_equal(r["dtheta"], g["theta"])


def synth(density, n, seed):
    rng = np.random.default_rng(seed)
    if density == "bernoulli_logit":
        theta = rng.standard_normal(n) * 4.0
        y = (rng.random(n) < 1.0 / (1.0 + np.exp(-theta))).astype(np.int32)
    elif density == "bernoulli":
        theta = rng.uniform(0.02, 0.98, n)
        y = (rng.random(n) < theta).astype(np.int32)
        return y, np.clip(theta + 0.01 * rng.standard_normal(n), 1e-3, 1 - 1e-3)
    elif density == "poisson":
        theta = rng.gamma(2.0, 2.0, n) + 0.05
        y = rng.poisson(theta).astype(np.int32)
        return y, theta * np.exp(0.1 * rng.standard_normal(n))
    else:
        theta = 0.5 + rng.standard_normal(n) * 1.2
        y = rng.poisson(np.exp(np.minimum(theta, 4.0))).astype(np.int32)
    return y, theta + 0.1 * rng.standard_normal(n)


@pytest.mark.parametrize("density", ["bernoulli_logit", "poisson_log", "poisson", "bernoulli"])
@pytest.mark.parametrize("n", [1, 2, 31, 1024, 1025, 100_003, 1_000_000])
def test_density_against_oracle(glm, L, density, n):
    y, theta = synth(density, n, seed=n + len(density))
    for propto in (False, True):
        ref = L.oracle_density(density, y, theta, propto=propto)
        got = glm.density_eval(density, y, theta, L.density_flags(propto))
        assert got["rc"] == ref["rc"] == 0
        c = {"lp": ref["lp"], "dtheta": ref["dtheta"], "theta": theta, "y": y, "density": _abi.DENSITIES[density]}
        check_dens_result(got, c, "%s n=%d propto=%d" % (density, n, propto))
    again = glm.density_eval(density, y, theta, FLAGS)
    first = glm.density_eval(density, y, theta, FLAGS)
    assert again["lp"] == first["lp"] and np.array_equal(again["dtheta"], first["dtheta"])   # fixed summation order
    # scalar operand: its partial is the sum over the elements (broadcast_array); scalar outcome
    ref = L.oracle_density(density, y, theta[:1])
    got = glm.density_eval(density, y, theta[:1], FLAGS)
    assert_close(got["lp"], ref["lp"], "scalar theta lp")
    assert_close(got["dtheta"], ref["dtheta"], "scalar theta dtheta")
    ref = L.oracle_density(density, y[:1], theta)
    got = glm.density_eval(density, y[:1], theta, FLAGS)
    assert_close(got["lp"], ref["lp"], "scalar n lp")
    assert_close(got["dtheta"], ref["dtheta"], "scalar n dtheta")
    # value only (double operand): no partials requested
    only = glm.density_eval(density, y, theta, _abi.LP_CONST | _abi.LP_MAIN)
    assert only["lp"] == first["lp"]


def test_density_outcome_cache_sees_new_contents(glm, L):
    """The outcomes are cached by host address + fingerprint: same buffer, new contents -> new upload."""
    rng = np.random.default_rng(9)
    theta = rng.standard_normal(5000)
    y = (rng.random(5000) < 0.5).astype(np.int32)
    a = glm.density_eval("bernoulli_logit", y, theta, FLAGS)
    y[:] = 1 - y
    b = glm.density_eval("bernoulli_logit", y, theta, FLAGS)
    ref = L.oracle_density("bernoulli_logit", y, theta)
    assert a["lp"] != b["lp"]
    assert_close(b["lp"], ref["lp"], "lp after the buffer changed")


def test_density_error_contract(glm):
    """bernoulli_logit_lpmf.hpp:36-49, poisson_log_lpmf.hpp:32-62 and the invalid values of the reference's
    distribution tests (test/prob/bernoulli/bernoulli_logit_test.hpp:50-58, test/prob/poisson/poisson_log_test.hpp:33-40)."""
    b, p = glm.bernoulli_logit_lpmf, glm.poisson_log_lpmf
    th = np.linspace(-1, 1, 4000)
    ones = np.ones(4000, dtype=np.int32)
    for bad in (-1, 2):
        y = ones.copy()
        y[1234] = bad
        with pytest.raises(glm.DomainError):
            b(y, th)
        with pytest.raises(glm.DomainError):
            b([bad], th)
    y = ones.copy()
    y[77] = -1
    with pytest.raises(glm.DomainError):
        p(y, th)
    tn = th.copy()
    tn[3999] = np.nan
    for f in (b, p):
        with pytest.raises(glm.DomainError):
            f(ones, tn)
        with pytest.raises(glm.DomainError):
            f(ones, [np.nan])
        with pytest.raises(ValueError) as e:       # size mismatch: std::invalid_argument, after the domain checks
            f(ones[:10], th[:9])
        assert not isinstance(e.value, glm.DomainError)
        assert f([], [])[0] == 0.0                                   # size_zero
        assert f(ones, th, propto=True, **{("theta_var" if f is b else "alpha_var"): False})[0] == 0.0
    # infinite logits are legal for bernoulli_logit (only NaN is checked, :43)
    lp, g = b([1, 0], [np.inf, -np.inf])
    assert lp == 0.0 and np.all(g["theta"] == 0.0)
    # poisson_log: LOG_ZERO without partials (:51-62); -inf next to a zero count contributes nothing (:77)
    lp, g = p([1, 2], [np.inf, 0.1])
    assert lp == -np.inf and np.all(g["theta"] == 0.0)
    lp, g = p([1, 2], [-np.inf, 0.1])
    assert lp == -np.inf
    lp, g = p([0, 2], [-np.inf, 0.1])
    assert np.isfinite(lp) and g["theta"][0] == 0.0
    # raw ABI argument validation
    with pytest.raises(Exception):
        glm.density_eval(7, [1], [0.1], FLAGS)
    with pytest.raises(Exception):
        glm.density_eval(0, [1], [0.1], FLAGS)
    with pytest.raises(Exception):
        glm.density_eval("poisson_log", [1, 2, 3], [0.1, 0.2], FLAGS)


def test_poisson_and_bernoulli_error_contract(glm, L):
    """poisson_lpmf.hpp:27-54, bernoulli_lpmf.hpp:37-50 and the invalid values of the reference's distribution tests
    (test/prob/poisson/poisson_test.hpp:32-43, test/prob/bernoulli/bernoulli_test.hpp:37-51)."""
    p, b = glm.poisson_lpmf, glm.bernoulli_lpmf
    lam = np.linspace(0.5, 3.0, 4000)
    th = np.linspace(0.05, 0.95, 4000)
    ones = np.ones(4000, dtype=np.int32)
    y = ones.copy()
    y[1234] = -1
    for f, op in ((p, lam), (b, th)):
        with pytest.raises(glm.DomainError):
            f(y, op)
        with pytest.raises(glm.DomainError):
            f([-1], op)
        bad = op.copy()
        bad[3999] = np.nan
        with pytest.raises(glm.DomainError):
            f(ones, bad)
        with pytest.raises(ValueError) as e:       # size mismatch: std::invalid_argument, after the domain checks
            f(ones[:10], op[:9])
        assert not isinstance(e.value, glm.DomainError)
        assert f([], [])[0] == 0.0                                   # size_zero
        assert f(ones, op, propto=True, **{("lambda_var" if f is p else "theta_var"): False})[0] == 0.0
    y = ones.copy()
    y[5] = 2
    with pytest.raises(glm.DomainError):
        b(y, th)
    for bad_value in (-1e-5, -1.0):
        bad = lam.copy()
        bad[17] = bad_value
        with pytest.raises(glm.DomainError):
            p(ones, bad)
    for bad_value in (-0.001, 1.001, np.inf):
        bad = th.copy()
        bad[17] = bad_value
        with pytest.raises(glm.DomainError):
            b(ones, bad)
    # poisson: LOG_ZERO without partials (:45-54); a zero rate next to a zero count contributes nothing (:59) and its
    # partial is 0/0 - 1, as in the reference
    lp, g = p([1, 2], [np.inf, 0.1])
    assert lp == -np.inf and np.all(g["theta"] == 0.0)
    assert p([1, 2], [0.0, 0.1])[0] == -np.inf
    lp, g = p([0, 2, 0], [0.0, 0.1, 0.4])
    ref = L.oracle_density("poisson", [0, 2, 0], [0.0, 0.1, 0.4])
    assert_close(lp, ref["lp"], "zero rate, zero count")
    assert np.isnan(g["theta"][0]) and np.allclose(g["theta"][1:], ref["dtheta"][1:], rtol=1e-12)
    # bernoulli: the ends of [0, 1] are legal
    lp, g = b([1, 0], [1.0, 0.0])
    assert lp == 0.0
    assert b([0], [1.0])[0] == -np.inf
    # scalar probability: the closed form from the number of successes, all three branches
    for yy in (np.ones(3000, np.int32), np.zeros(3000, np.int32), (np.arange(3000) % 3 == 0).astype(np.int32)):
        got = glm.density_eval("bernoulli", yy, [0.3], FLAGS)
        ref = L.oracle_density("bernoulli", yy, [0.3])
        assert_close(got["lp"], ref["lp"], "scalar probability lp", 1e-14)
        assert_close(got["dtheta"], ref["dtheta"], "scalar probability dtheta", 1e-14)
