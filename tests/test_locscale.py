"""cauchy_lpdf / lognormal_lpdf / gamma_lpdf / exponential_lpdf (and normal_lpdf through the same entry point) on the device
(stan_cuda_locscale_lpdf_eval, csrc/normal_lpdf.cu): the plain-C oracle against the golden vectors of the compiled
reference (CPU), and the CUDA path against both (GPU).  Bar: 1e-10 relative on the log-density and on every partial."""
import numpy as np
import pytest

from math_b200 import _abi
from oracle import loader as L
from tests.helpers import assert_close

needs_ref = pytest.mark.skipif(not L.have_ref(), reason="oracle/_ref/libstanref.so not built")
KEYS = (("dy", 1), ("dmu", 2), ("dsigma", 4))


@pytest.fixture(scope="module")
def golden_ls():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    return np.load(os.path.join(root, "tests", "golden", "locscale_golden.npz"), allow_pickle=False)


def ls_case(g, name):
    p = "ls/%s/" % name
    c = {k: g[p + k] for k in ("y", "mu", "sigma")}
    c["density"], c["propto"], c["var_mask"] = int(g[p + "density"]), bool(g[p + "propto"]), int(g[p + "var_mask"])
    c["lp"] = float(g[p + "lp"])
    for k, _ in KEYS:
        if p + k in g.files:
            c[k] = g[p + k]
    if p + "known_answer" in g.files:
        c["known_answer"] = float(g[p + "known_answer"])
    return c


def gamma_dalpha_scale(y, alpha, beta, size):
    """d/dalpha of gamma_lpdf is -digamma(alpha) + log(beta) + log(y) (gamma_lpdf.hpp:134-137): three O(1) terms that may
    cancel, so the partial is accurate to a few ulp of their magnitudes, not of itself (the floor student_t's d/dnu has)."""
    from scipy.special import digamma
    y, alpha, beta = (np.asarray(v, float) for v in (y, alpha, beta))
    with np.errstate(divide="ignore"):
        terms = np.abs(digamma(alpha)) + np.abs(np.log(beta)) + np.abs(np.where(y > 0, np.log(np.where(y > 0, y, 1.0)), 0.0))
    terms = np.broadcast_to(terms, (max(y.size, alpha.size, beta.size),))
    return terms if size > 1 else np.array([terms.sum()])


def check_gamma_dalpha(got, ref, c, what, rtol):
    scale = np.maximum(np.abs(ref), gamma_dalpha_scale(c["y"], c["mu"], c["sigma"], ref.size))
    err = np.abs(np.asarray(got, float) - ref) / scale
    assert not (err > rtol).any(), "%s dmu: %d components off, worst %.3e" % (what, int((err > rtol).sum()), err.max())


def check_beta_shape_partial(got, ref, c, key, what, rtol):
    """d/dalpha = log(y) + digamma(alpha + beta) - digamma(alpha), d/dbeta = log1m(y) + ... (beta_lpdf.hpp:170-177)"""
    from scipy.special import digamma
    y, a, b = (np.asarray(c[k], float) for k in ("y", "mu", "sigma"))
    own = a if key == "dmu" else b
    terms = np.abs(np.log(y) if key == "dmu" else np.log1p(-y)) + np.abs(digamma(a + b)) + np.abs(digamma(own))
    terms = np.broadcast_to(terms, (max(y.size, a.size, b.size),))
    scale = np.maximum(np.abs(ref), terms if ref.size > 1 else terms.sum())
    err = np.abs(np.asarray(got, float) - ref) / scale
    assert not (err > rtol).any(), "%s %s: %d components off, worst %.3e" % (what, key, int((err > rtol).sum()), err.max())


def weibull_scale(c, key, size):
    """weibull's partials are products / sums of (y / sigma)^alpha, logs and reciprocals (weibull_lpdf.hpp:125-139): the
    floor is the magnitude of their terms"""
    y, a, s = (np.asarray(c[k], float) for k in ("y", "mu", "sigma"))
    pw = np.power(y / s, a)
    if key == "dy":
        terms = (np.abs(a - 1.0) + a * pw) / y
    elif key == "dmu":
        terms = 1.0 / a + (1.0 + pw) * (np.abs(np.log(y)) + np.abs(np.log(s)))
    else:
        terms = a / s * (pw + 1.0)
    terms = np.broadcast_to(terms, (max(y.size, a.size, s.size),))
    return terms if size > 1 else np.array([terms.sum()])


def check_case(r, c, what, rtol=1e-10):
    assert r["rc"] == 0, what
    assert_close(r["lp"], c["lp"], what + " lp", rtol)
    for k, bit in KEYS:
        if c["var_mask"] & bit and k == "dmu" and c.get("density") in (_abi.LOCSCALE_GAMMA, _abi.LOCSCALE_INV_GAMMA):
            check_gamma_dalpha(r[k], np.asarray(c[k], float), c, what, rtol)
        elif c["var_mask"] & bit and k != "dy" and c.get("density") == _abi.LOCSCALE_BETA:
            check_beta_shape_partial(r[k], np.asarray(c[k], float), c, k, what, rtol)
        elif c["var_mask"] & bit and c.get("density") == _abi.LOCSCALE_WEIBULL:
            ref = np.asarray(c[k], float)
            err = np.abs(np.asarray(r[k], float) - ref) / np.maximum(np.abs(ref), weibull_scale(c, k, ref.size))
            assert not (err > rtol).any(), "%s %s: %d components off, worst %.3e" % (what, k, int((err > rtol).sum()), err.max())
        elif c["var_mask"] & bit:
            assert_close(r[k], c[k], what + " " + k, rtol)
            # element by element: relative to the element, floored at 1e-8 of the largest one
            ref, got = np.asarray(c[k], float), np.asarray(r[k], float)
            err = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-8 * max(np.max(np.abs(ref)), 1e-300))
            assert not (err > rtol).any(), "%s %s: %d components off, worst %.3e" % (what, k, int((err > rtol).sum()), err.max())
        else:
            assert k not in r


def test_oracle_matches_golden_and_known_answers(golden_ls):
    """outputs of the compiled reference and the literal expected values of test/prob/cauchy/cauchy_test.hpp:10-31,
    test/prob/lognormal/lognormal_test.hpp:10-25"""
    assert len(golden_ls["cases"]) >= 144
    known = 0
    for name in golden_ls["cases"]:
        c = ls_case(golden_ls, str(name))
        r = L.oracle_locscale_lpdf(c["density"], c["y"], c["mu"], c["sigma"], propto=c["propto"], var_mask=c["var_mask"])
        check_case(r, c, "oracle vs golden " + str(name), rtol=1e-12)
        if "known_answer" in c:
            known += 1
            assert abs(r["lp"] - c["known_answer"]) <= 1e-13 * max(1.0, abs(c["known_answer"])), name
    assert known == 18   # 3 cauchy, 2 lognormal, 3 gamma, 3 exponential, 2 beta, 3 weibull, 2 inv_gamma


def test_oracle_codes():
    """BAD_OPERAND where the reference throws (cauchy_lpdf.hpp:47-49, lognormal_lpdf.hpp:24-27; the invalid values of the
    reference's fixtures: infinite location, zero / negative / infinite scale), LOG_ZERO for a lognormal outcome of zero."""
    for d in ("cauchy", "lognormal"):
        for mu in (np.inf, -np.inf, np.nan):
            assert L.oracle_locscale_lpdf(d, [1.0], [mu], [1.0])["rc"] == _abi.BAD_OPERAND
        for sg in (0.0, -1.0, np.inf, -np.inf, np.nan):
            assert L.oracle_locscale_lpdf(d, [1.0], [0.0], [sg])["rc"] == _abi.BAD_OPERAND
        assert L.oracle_locscale_lpdf(d, [np.nan], [0.0], [1.0])["rc"] == _abi.BAD_OPERAND
        assert L.oracle_locscale_lpdf(d, [], [], [])["lp"] == 0.0
    for al in (0.0, -1.0, np.inf, -np.inf, np.nan):      # test/prob/gamma/gamma_test.hpp:36-61
        assert L.oracle_locscale_lpdf("gamma", [1.0], [al], [1.0])["rc"] == _abi.BAD_OPERAND
        assert L.oracle_locscale_lpdf("gamma", [1.0], [1.0], [al])["rc"] == _abi.BAD_OPERAND
        assert L.oracle_locscale_lpdf("exponential", [1.0], [0.0], [al])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_locscale_lpdf("gamma", [np.nan], [1.0], [1.0])["rc"] == _abi.BAD_OPERAND
    r = L.oracle_locscale_lpdf("gamma", [1.0, -0.5], [1.0], [1.0])
    assert r["rc"] == _abi.LOG_ZERO and r["lp"] == -np.inf                                  # gamma_lpdf.hpp:69-74
    for bad in (0.0, -1.0, np.inf, -np.inf, np.nan):     # test/prob/beta/beta_test.hpp:36-60, weibull_test.hpp:37-49
        for d in ("beta", "weibull"):
            assert L.oracle_locscale_lpdf(d, [0.5], [bad], [1.0])["rc"] == _abi.BAD_OPERAND
            assert L.oracle_locscale_lpdf(d, [0.5], [1.0], [bad])["rc"] == _abi.BAD_OPERAND
    for bad_y in (-1.0, 2.0, np.nan):
        assert L.oracle_locscale_lpdf("beta", [bad_y], [1.0], [1.0])["rc"] == _abi.BAD_OPERAND
    for bad_y in (np.inf, np.nan):
        assert L.oracle_locscale_lpdf("weibull", [bad_y], [1.0], [1.0])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_locscale_lpdf("weibull", [1.0, -0.5], [1.0], [1.0])["rc"] == _abi.LOG_ZERO   # weibull_lpdf.hpp:56-61
    for bad in (0.0, -1.0, np.inf, np.nan):              # test/prob/inv_gamma/inv_gamma_test.hpp:30-55
        assert L.oracle_locscale_lpdf("inv_gamma", [0.5], [bad], [1.0])["rc"] == _abi.BAD_OPERAND
        assert L.oracle_locscale_lpdf("inv_gamma", [0.5], [1.0], [bad])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_locscale_lpdf("inv_gamma", [np.nan], [1.0], [1.0])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_locscale_lpdf("inv_gamma", [1.0, 0.0], [1.0], [1.0])["rc"] == _abi.LOG_ZERO      # inv_gamma_lpdf.hpp:58-63
    for bad_y in (-10.0, np.nan):                       # test/prob/exponential/exponential_test.hpp:31-36
        assert L.oracle_locscale_lpdf("exponential", [bad_y], [0.0], [1.0])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_locscale_lpdf("lognormal", [-1.0], [0.0], [1.0])["rc"] == _abi.BAD_OPERAND
    r = L.oracle_locscale_lpdf("lognormal", [1.0, 0.0], [0.0], [1.0])
    assert r["rc"] == _abi.LOG_ZERO and r["lp"] == -np.inf
    assert L.oracle_locscale_lpdf("cauchy", [np.inf], [0.0], [1.0])["rc"] == _abi.NONFINITE   # legal outcome, -inf density
    assert L.oracle_locscale_lpdf("normal", [1.0], [0.0], [np.inf])["rc"] in (0, _abi.NONFINITE)   # check_positive only


def synth(density, n, rng):
    """operands in the density's domain (gamma: mu holds the shape; exponential: mu is one zero)"""
    y = rng.standard_normal(n) * 2
    mu, sg = rng.standard_normal(n), rng.uniform(0.3, 3.0, n)
    if density == "lognormal":
        y = np.exp(y)
    elif density in ("gamma", "exponential", "inv_gamma"):
        y = rng.gamma(2.0, 1.5, n) + (1e-3 if density == "inv_gamma" else 0.0)
        mu = np.zeros(1) if density == "exponential" else rng.uniform(0.2, 8.0, n)
    elif density == "beta":
        y, mu, sg = rng.beta(2.0, 3.0, n).clip(1e-9, 1 - 1e-9), rng.uniform(0.3, 6.0, n), rng.uniform(0.3, 6.0, n)
    elif density == "weibull":
        y, mu = rng.weibull(1.5, n) * 2.0 + 1e-6, rng.uniform(0.4, 4.0, n)
    return y, mu, sg


def masks(density):
    return (0, 1, 4, 5) if density == "exponential" else range(8)


DENSITIES = ["normal", "cauchy", "lognormal", "gamma", "exponential", "beta", "weibull", "inv_gamma"]


@needs_ref
@pytest.mark.parametrize("density", DENSITIES)
def test_oracle_vs_compiled_reference(density):
    rng = np.random.default_rng(5)
    n = 3000
    y, mu, sg = synth(density, n, rng)
    for var_mask in masks(density):
        for propto in (False, True):
            for ops in ((y, mu, sg), (y, mu[:1], sg[:1]), (y[:1], mu, sg)):
                ref = L.ref_locscale_lpdf(density, *ops, propto=propto, var_mask=var_mask)
                got = L.oracle_locscale_lpdf(density, *ops, propto=propto, var_mask=var_mask)
                assert got["rc"] == 0
                assert_close(got["lp"], ref["lp"], "%s lp" % density, 1e-12)
                c = {"y": ops[0], "mu": ops[1], "sigma": ops[2]}
                for k, bit in KEYS:
                    if var_mask & bit and k == "dmu" and density in ("gamma", "inv_gamma"):
                        check_gamma_dalpha(got[k], ref[k], c, "gamma", 1e-12)
                    elif var_mask & bit and k != "dy" and density == "beta":
                        check_beta_shape_partial(got[k], ref[k], c, k, "beta", 1e-12)
                    elif var_mask & bit and density == "weibull":
                        err = np.abs(got[k] - ref[k]) / np.maximum(np.abs(ref[k]), weibull_scale(c, k, ref[k].size))
                        assert not (err > 1e-12).any(), ("weibull", k, err.max())
                    elif var_mask & bit:   # (a scalar operand's partial is a sum of n terms that may cancel: the parity bar)
                        assert_close(got[k], ref[k], "%s %s" % (density, k), 1e-12 if np.size(ref[k]) > 1 else 1e-10)


# ---- the CUDA path -------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    g.init(1)
    g.cache_clear()
    yield g
    g.cache_clear()


@pytest.mark.gpu
def test_device_matches_golden(glm, golden_ls):
    by_id = {_abi.LOCSCALE_CAUCHY: glm.cauchy_lpdf, _abi.LOCSCALE_LOGNORMAL: glm.lognormal_lpdf}
    for name in golden_ls["cases"]:
        c = ls_case(golden_ls, str(name))
        r = glm.locscale_eval(c["density"], c["y"], c["mu"], c["sigma"], glm.locscale_flags(c["density"], c["propto"], c["var_mask"]),
                              y_is_data=not (c["var_mask"] & 1))
        check_case(r, c, "cuda vs golden " + str(name))
        if "known_answer" in c:
            assert abs(r["lp"] - c["known_answer"]) <= 1e-12 * max(1.0, abs(c["known_answer"])), name
        if c["density"] not in by_id:
            continue
        lp, g = by_id[c["density"]](c["y"], c["mu"], c["sigma"], propto=c["propto"], var_mask=c["var_mask"])
        assert lp == r["lp"] or (c["propto"] and c["var_mask"] == 0 and lp == 0.0)
        for (key, bit), gname in zip(KEYS, ("y", "mu", "sigma")):
            if c["var_mask"] & bit:
                assert np.array_equal(g[gname], r[key])


@pytest.mark.gpu
@pytest.mark.parametrize("density", DENSITIES)
@pytest.mark.parametrize("n", [1, 2, 31, 1025, 100_003, 1_000_000])
def test_device_against_oracle(glm, density, n):
    rng = np.random.default_rng(n + len(density))
    y, mu, sg = synth(density, n, rng)
    expo = density == "exponential"
    for var_mask, propto in ((0b101 if expo else 0b111, False), (0b100 if expo else 0b110, True), (0b001, True), (0b000, False)):
        flags = glm.locscale_flags(density, propto, var_mask)
        ref = L.oracle_locscale_lpdf(density, y, mu, sg, flags=flags)
        got = glm.locscale_eval(density, y, mu, sg, flags, y_is_data=not (var_mask & 1))
        c = dict(ref, var_mask=var_mask, density=_abi.LOCSCALES[density], y=y, mu=mu, sigma=sg)
        check_case(got, c, "%s n=%d mask=%d propto=%d" % (density, n, var_mask, propto))
    full = 0b101 if expo else 0b111
    flags = glm.locscale_flags(density, False, full)
    first = glm.locscale_eval(density, y, mu, sg, flags, y_is_data=False)
    again = glm.locscale_eval(density, y, mu, sg, flags, y_is_data=False)
    assert first["lp"] == again["lp"] and all(np.array_equal(first[k], again[k]) for k, b in KEYS if full & b)   # fixed summation order
    # scalar operands: their partials are sums over the elements
    for ops in ((y, mu[:1], sg[:1]), (y[:1], mu, sg), (y, mu, sg[:1])):
        ref = L.oracle_locscale_lpdf(density, *ops, flags=flags)
        got = glm.locscale_eval(density, *ops, flags, y_is_data=False)
        check_case(got, dict(ref, var_mask=full, density=_abi.LOCSCALES[density], y=ops[0], mu=ops[1], sigma=ops[2]),
                   "%s scalar operands n=%d" % (density, n))


@pytest.mark.gpu
def test_device_error_contract(glm):
    """cauchy_lpdf.hpp:42-58, lognormal_lpdf.hpp:24-45 and the invalid values of the reference's fixtures"""
    n = 3000
    y, mu, sg = np.linspace(0.1, 5, n), np.linspace(-1, 1, n), np.linspace(0.5, 2, n)
    for d, f in (("cauchy", glm.cauchy_lpdf), ("lognormal", glm.lognormal_lpdf)):
        for k, bad_value in ((1, np.inf), (1, -np.inf), (2, 0.0), (2, -1.0), (2, np.inf), (0, np.nan)):
            ops = [y.copy(), mu.copy(), sg.copy()]
            ops[k][1234] = bad_value
            with pytest.raises(glm.DomainError):
                f(*ops)
            flags = glm.locscale_flags(d, False, 0b110)
            # the kernel's own element test (y travels per call: a cached data vector is trusted by address + samples)
            assert glm.locscale_eval(d, *ops, flags, y_is_data=False)["rc"] == _abi.BAD_OPERAND
        with pytest.raises(ValueError) as e:
            f(y[:10], mu[:9], sg[:10])
        assert not isinstance(e.value, glm.DomainError)
        assert f([], [], [])[0] == 0.0
        assert f(y, mu, sg, propto=True, var_mask=0)[0] == 0.0
    # lognormal: a negative outcome throws, an outcome of zero is LOG_ZERO without partials
    yn = y.copy()
    yn[7] = -0.5
    with pytest.raises(glm.DomainError):
        glm.lognormal_lpdf(yn, mu, sg)
    yz = y.copy()
    yz[7] = 0.0
    lp, g = glm.lognormal_lpdf(yz, mu, sg)
    assert lp == -np.inf and g == {}
    assert glm.lognormal_lpdf(yz, mu, sg, propto=True, var_mask=0)[0] == -np.inf
    with pytest.raises(Exception):
        glm.locscale_eval(9, y, mu, sg, glm.locscale_flags("cauchy"))


@pytest.mark.gpu
def test_device_gamma_and_exponential_wrappers(glm, golden_ls):
    """glm.gamma_lpdf / glm.exponential_lpdf against the golden vectors, and their error contract
    (gamma_lpdf.hpp:48-74, exponential_lpdf.hpp:48-58; the invalid values of the reference's fixtures)"""
    for name in golden_ls["cases"]:
        c = ls_case(golden_ls, str(name))
        if c["density"] == _abi.LOCSCALE_GAMMA:
            lp, g = glm.gamma_lpdf(c["y"], c["mu"], c["sigma"], propto=c["propto"], var_mask=c["var_mask"])
            keys = (("y", "dy", 1), ("alpha", "dmu", 2), ("beta", "dsigma", 4))
        elif c["density"] == _abi.LOCSCALE_EXPONENTIAL:
            vm2 = (c["var_mask"] & 1) | ((c["var_mask"] & 4) >> 1)
            lp, g = glm.exponential_lpdf(c["y"], c["sigma"], propto=c["propto"], var_mask=vm2)
            keys = (("y", "dy", 1), ("beta", "dsigma", 4))
        else:
            continue
        assert_close(lp, c["lp"], "wrapper vs golden %s lp" % name)
        for gname, key, bit in keys:
            if c["var_mask"] & bit:
                assert_close(g[gname], c[key], "wrapper vs golden %s %s" % (name, key))
    n = 3000
    y, al, be = np.linspace(0.1, 5, n), np.linspace(0.5, 4, n), np.linspace(0.5, 2, n)
    for bad_value in (0.0, -1.0, np.inf, np.nan):
        for k in (1, 2):
            ops = [y.copy(), al.copy(), be.copy()]
            ops[k][1234] = bad_value
            with pytest.raises(glm.DomainError):
                glm.gamma_lpdf(*ops)
            assert glm.locscale_eval("gamma", *ops, glm.locscale_flags("gamma"), y_is_data=False)["rc"] == _abi.BAD_OPERAND
        bb = be.copy()
        bb[7] = bad_value
        with pytest.raises(glm.DomainError):
            glm.exponential_lpdf(y, bb)
        assert glm.locscale_eval("exponential", y, [0.0], bb, glm.locscale_flags("exponential", False, 0b100),
                                 y_is_data=False)["rc"] == _abi.BAD_OPERAND
    yn = y.copy()
    yn[5] = -0.5
    lp, g = glm.gamma_lpdf(yn, al, be)                       # a negative outcome: LOG_ZERO, no partials (:69-74)
    assert lp == -np.inf and g == {}
    with pytest.raises(glm.DomainError):
        glm.exponential_lpdf(yn, be)
    ynan = y.copy()
    ynan[5] = np.nan
    with pytest.raises(glm.DomainError):
        glm.gamma_lpdf(ynan, al, be)
    with pytest.raises(ValueError) as e:
        glm.gamma_lpdf(y[:10], al[:9], be[:10])
    assert not isinstance(e.value, glm.DomainError)
    assert glm.gamma_lpdf([], [], [])[0] == 0.0 and glm.exponential_lpdf([], [])[0] == 0.0
    assert glm.gamma_lpdf(y, al, be, propto=True, var_mask=0)[0] == 0.0


@pytest.mark.gpu
def test_device_beta_and_weibull_wrappers(glm, golden_ls):
    """glm.beta_lpdf / glm.weibull_lpdf against the golden vectors (value), and their error contract
    (beta_lpdf.hpp:48-56, weibull_lpdf.hpp:38-61; the invalid values of the reference's fixtures)"""
    for name in golden_ls["cases"]:
        c = ls_case(golden_ls, str(name))
        f = {_abi.LOCSCALE_BETA: glm.beta_lpdf, _abi.LOCSCALE_WEIBULL: glm.weibull_lpdf,
             _abi.LOCSCALE_INV_GAMMA: glm.inv_gamma_lpdf}.get(c["density"])
        if f is None:
            continue
        lp, g = f(c["y"], c["mu"], c["sigma"], propto=c["propto"], var_mask=c["var_mask"])
        assert_close(lp, c["lp"], "wrapper vs golden %s lp" % name)
        assert sorted(g) == sorted(nm for nm, bit in (("y", 1), ("alpha", 2), ("sigma" if f is glm.weibull_lpdf else "beta", 4))
                                   if c["var_mask"] & bit)
    n = 3000
    y, a, b = np.linspace(0.05, 0.95, n), np.linspace(0.5, 4, n), np.linspace(0.5, 2, n)
    for f, d in ((glm.beta_lpdf, "beta"), (glm.weibull_lpdf, "weibull")):
        for bad_value in (0.0, -1.0, np.inf, np.nan):
            for k in (1, 2):
                ops = [y.copy(), a.copy(), b.copy()]
                ops[k][1234] = bad_value
                with pytest.raises(glm.DomainError):
                    f(*ops)
                assert glm.locscale_eval(d, *ops, glm.locscale_flags(d), y_is_data=False)["rc"] == _abi.BAD_OPERAND
        with pytest.raises(ValueError) as e:
            f(y[:10], a[:9], b[:10])
        assert not isinstance(e.value, glm.DomainError)
        assert f([], [], [])[0] == 0.0 and f(y, a, b, propto=True, var_mask=0)[0] == 0.0
    for bad_y in (-0.1, 1.5, np.nan):
        yy = y.copy()
        yy[9] = bad_y
        with pytest.raises(glm.DomainError):
            glm.beta_lpdf(yy, a, b)
        assert glm.locscale_eval("beta", yy, a, b, glm.locscale_flags("beta"), y_is_data=False)["rc"] == _abi.BAD_OPERAND
    yy = y.copy()
    yy[9] = -0.5
    lp, g = glm.weibull_lpdf(yy, a, b)          # a negative outcome: LOG_ZERO, no partials
    assert lp == -np.inf and g == {}
    yy[9] = np.inf
    with pytest.raises(glm.DomainError):
        glm.weibull_lpdf(yy, a, b)
    yy[9] = 0.0
    lp, g = glm.inv_gamma_lpdf(yy, a, b)        # an outcome that is not positive: LOG_ZERO
    assert lp == -np.inf and g == {}
    yy[9] = np.nan
    with pytest.raises(glm.DomainError):
        glm.inv_gamma_lpdf(yy, a, b)
