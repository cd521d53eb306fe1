"""CPU tests (no GPU): the plain-C oracle against the reference's golden vectors, and against
the compiled reference itself when oracle/_ref/libstanref.so is present."""
import numpy as np
import pytest

from oracle import loader as L
from tests.helpers import (assert_close, cat_case, check_cat_result, check_dens_result, check_glm_result,
                           check_ord_result, dens_case, glm_case, normal_case, ord_case, synth_categorical, synth_glm,
                           synth_ordered)
from math_b200 import _abi


def test_golden_file_has_cases(golden):
    assert len(golden["glm_cases"]) >= 30 and len(golden["normal_cases"]) >= 8


def test_oracle_glm_matches_golden(golden):
    for name in golden["glm_cases"]:
        c = glm_case(golden, str(name))
        r = L.oracle_glm(c["family"], c["x"], c["y"], c["alpha"], c["beta"], c["sigma"],
                         propto=c["propto"], need_dy=c["need_dy"])
        assert r["rc"] == 0, name
        check_glm_result(r, c, "oracle vs golden " + str(name), rtol=1e-11)


def test_oracle_ordered_logistic_matches_golden(golden):
    assert len(golden["ord_cases"]) >= 8
    for name in golden["ord_cases"]:
        c = ord_case(golden, str(name))
        r = L.oracle_ordered_logistic_glm(c["x"], c["y"], c["beta"], c["cuts"], propto=c["propto"])
        assert r["rc"] == 0, name
        check_ord_result(r, c, "oracle vs golden ord " + str(name), rtol=1e-11)


def test_oracle_categorical_logit_matches_golden(golden):
    assert len(golden["cat_cases"]) >= 7
    for name in golden["cat_cases"]:
        c = cat_case(golden, str(name))
        r = L.oracle_categorical_logit_glm(c["x"], c["y"], c["alpha"], c["beta"], propto=c["propto"])
        assert r["rc"] == 0, name
        check_cat_result(r, c, "oracle vs golden cat " + str(name), rtol=1e-11)


def test_oracle_normal_matches_golden_and_known_answers(golden):
    for name in golden["normal_cases"]:
        c = normal_case(golden, str(name))
        r = L.oracle_normal_lpdf(c["y"], c["mu"], c["sigma"], propto=c["propto"], need_dy=c["need_dy"])
        assert_close(r["lp"], c["lp"], str(name) + " lp", 1e-12)
        assert_close(r["dmu"], c["dmu"], str(name) + " dmu", 1e-12)
        assert_close(r["dsigma"], c["dsigma"], str(name) + " dsigma", 1e-11)
        if c["need_dy"]:
            assert_close(r["dy"], c["dy"], str(name) + " dy", 1e-12)
        if "known_answer" in c:  # literals of test/prob/normal/normal_test.hpp:11-40
            assert abs(r["lp"] - c["known_answer"]) < 1e-14 * max(1, abs(c["known_answer"]))
            assert abs(c["lp"] - c["known_answer"]) < 1e-14 * max(1, abs(c["known_answer"]))


def test_oracle_reference_quirks():
    """Behaviours of the reference a 'cleaner' implementation would miss."""
    x = np.array([[-12, 46], [-42, 24], [25, 27]], float) * 0.01
    # poisson<propto=true> with var arguments returns 0 but keeps its gradient
    r = L.oracle_glm("poisson_log", x, [15, 3, 5], 0.3, [0.3, 2.0], propto=True)
    assert r["lp"] == 0.0 and np.all(r["dbeta"] != 0)
    # bernoulli: for ytheta > 20 the derivative is -exp(-ytheta), without the sign factor
    xe = np.array([[30.0]])
    r = L.oracle_glm("bernoulli_logit", xe, [1], 0.0, [1.0])
    assert r["dalpha"][0] == -np.exp(-30.0)
    r = L.oracle_glm("bernoulli_logit", xe, [0], 0.0, [1.0])   # ytheta = -30 < -20: d = sign = -1
    assert r["dalpha"][0] == -1.0 and r["lp"] == -30.0


def test_oracle_nonfinite_status():
    x = np.array([[1.0, 2.0], [3.0, np.inf]])
    assert L.oracle_glm("poisson_log", x, [1, 2], 0.1, [0.5, 0.5])["rc"] == _abi.NONFINITE
    assert L.oracle_glm("normal_id", x, [1.0, 2.0], 0.1, [0.5, 0.5], [1.0])["rc"] == _abi.NONFINITE
    assert L.oracle_glm("bernoulli_logit", x * np.nan, [1, 0], 0.1, [0.5, 0.5])["rc"] == _abi.NONFINITE


needs_ref = pytest.mark.skipif(not L.have_ref(), reason="oracle/_ref/libstanref.so not built")


@needs_ref
@pytest.mark.parametrize("family", ["bernoulli_logit", "poisson_log", "normal_id"])
@pytest.mark.parametrize("n,k", [(1, 1), (7, 3), (500, 32), (1200, 130)])
def test_oracle_vs_compiled_reference(family, n, k):
    s = synth_glm(family, n, k, seed=n * 131 + k)
    for propto in (False, True):
        ref = L.ref_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"], propto=propto)
        got = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"], propto=propto)
        check_glm_result(got, ref, "%s %dx%d propto=%s" % (family, n, k, propto), rtol=1e-11)


@needs_ref
@pytest.mark.parametrize("n,k,c", [(1, 1, 2), (50, 3, 1), (500, 32, 5), (1200, 130, 12)])
def test_oracle_ordered_logistic_vs_compiled_reference(n, k, c):
    s = synth_ordered(n, k, c, seed=n + 7 * k + c)
    ref = L.ref_ordered_logistic_glm(s["x"], s["y"], s["beta"], s["cuts"])
    got = L.oracle_ordered_logistic_glm(s["x"], s["y"], s["beta"], s["cuts"])
    check_ord_result(got, ref, "ordered %dx%d C=%d" % (n, k, c), rtol=1e-11)


@needs_ref
@pytest.mark.parametrize("n,k,c", [(1, 1, 2), (50, 3, 1), (500, 32, 5), (900, 130, 12)])
def test_oracle_categorical_logit_vs_compiled_reference(n, k, c):
    s = synth_categorical(n, k, c, seed=n + 7 * k + c)
    ref = L.ref_categorical_logit_glm(s["x"], s["y"], s["alpha"], s["beta"])
    got = L.oracle_categorical_logit_glm(s["x"], s["y"], s["alpha"], s["beta"])
    check_cat_result(got, ref, "categorical %dx%d C=%d" % (n, k, c), rtol=1e-11)


def test_oracle_densities_match_golden_and_known_answers(golden):
    """bernoulli_logit_lpmf / poisson_log_lpmf: outputs of the compiled reference, and the literal expected values
    of the reference's own tests (test/prob/bernoulli/bernoulli_logit_test.hpp:11-48, test/prob/poisson/poisson_log_test.hpp:11-31)."""
    assert len(golden["dens_cases"]) >= 19
    known = 0
    for name in golden["dens_cases"]:
        c = dens_case(golden, str(name))
        r = L.oracle_density(c["density"], c["y"], c["theta"], propto=c["propto"])
        assert r["rc"] == 0, name
        check_dens_result(r, c, "oracle vs golden dens " + str(name), rtol=1e-12)
        if "known_answer" in c:
            known += 1
            assert abs(r["lp"] - c["known_answer"]) <= 1e-13 * max(1.0, abs(c["known_answer"])), name
    assert known == 10


def test_oracle_poisson_and_bernoulli_match_golden_and_known_answers(golden_d2):
    """poisson_lpmf / bernoulli_lpmf: outputs of the compiled reference, and the literal expected values of the
    reference's own tests (test/prob/poisson/poisson_test.hpp:11-30, test/prob/bernoulli/bernoulli_test.hpp:11-35)."""
    assert len(golden_d2["dens_cases"]) >= 18
    known = 0
    for name in golden_d2["dens_cases"]:
        c = dens_case(golden_d2, str(name))
        r = L.oracle_density(c["density"], c["y"], c["theta"], propto=c["propto"])
        assert r["rc"] == 0, name
        check_dens_result(r, c, "oracle vs golden dens " + str(name), rtol=1e-12)
        if "known_answer" in c:
            known += 1
            assert abs(r["lp"] - c["known_answer"]) <= 1e-13 * max(1.0, abs(c["known_answer"])), name
    assert known == 7


def test_oracle_poisson_and_bernoulli_codes():
    """What the product must report where the reference throws (poisson_lpmf.hpp:31-33, bernoulli_lpmf.hpp:41-43; the
    invalid values of test/prob/poisson/poisson_test.hpp:32-43, test/prob/bernoulli/bernoulli_test.hpp:37-51) or
    returns LOG_ZERO (poisson_lpmf.hpp:45-54)."""
    for bad_n, bad_t in (([-1], [1.0]), ([1], [-1e-5]), ([1], [-1.0]), ([1, 2], [1.0, np.nan])):
        assert L.oracle_density("poisson", bad_n, bad_t)["rc"] == _abi.BAD_OPERAND
    for bad_n, bad_t in (([-1], [0.5]), ([2], [0.5]), ([1], [-0.001]), ([1], [1.001]), ([1, 0], [0.5, np.nan]), ([1], [np.inf])):
        assert L.oracle_density("bernoulli", bad_n, bad_t)["rc"] == _abi.BAD_OPERAND
    r = L.oracle_density("poisson", [1, 2], [np.inf, 0.1])
    assert r["rc"] == _abi.LOG_ZERO and r["lp"] == -np.inf
    assert L.oracle_density("poisson", [1, 2], [0.0, 0.1])["rc"] == _abi.LOG_ZERO
    assert L.oracle_density("poisson", [0, 2], [0.0, 0.1])["rc"] == 0
    assert L.oracle_density("poisson", [], [])["lp"] == 0.0 and L.oracle_density("bernoulli", [], [])["lp"] == 0.0
    # the corners of the support are legal: log(0) = -inf where the outcome contradicts a certain probability
    assert L.oracle_density("bernoulli", [1, 0], [1.0, 0.0])["lp"] == 0.0
    assert L.oracle_density("bernoulli", [0], [1.0])["lp"] == -np.inf


@needs_ref
@pytest.mark.parametrize("density", ["poisson", "bernoulli"])
def test_oracle_poisson_and_bernoulli_vs_compiled_reference(density):
    rng = np.random.default_rng(4)
    n = 5000
    theta = rng.gamma(2.0, 1.5, n) if density == "poisson" else rng.uniform(0.02, 0.98, n)
    y = rng.poisson(2.0, n).astype(np.int32) if density == "poisson" else (rng.random(n) < 0.4).astype(np.int32)
    for propto in (False, True):
        for yy, th in ((y, theta), (y, theta[:1]), (y[:1], theta)):
            ref = L.ref_density(density, yy, th, propto=propto)
            got = L.oracle_density(density, yy, th, propto=propto)
            assert got["rc"] == 0
            assert_close(got["lp"], ref["lp"], "%s lp" % density, 1e-12)
            assert_close(got["dtheta"], ref["dtheta"], "%s dtheta" % density, 1e-12)
    if density == "poisson":   # a zero rate next to a zero count: the term is skipped (:59), the partial is 0/0 - 1
        ref = L.ref_density("poisson", [0, 2, 0], [0.0, 0.1, 0.4])
        got = L.oracle_density("poisson", [0, 2, 0], [0.0, 0.1, 0.4])
        assert got["lp"] == ref["lp"] and np.array_equal(got["dtheta"], ref["dtheta"], equal_nan=True)
        assert np.isnan(ref["dtheta"][0])


def test_oracle_density_codes():
    """What the product must report where the reference throws or returns LOG_ZERO."""
    assert L.oracle_density("bernoulli_logit", [2], [0.1])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_density("bernoulli_logit", [1, 0], [0.1, np.nan])["rc"] == _abi.BAD_OPERAND
    assert L.oracle_density("poisson_log", [-1], [0.1])["rc"] == _abi.BAD_OPERAND
    r = L.oracle_density("poisson_log", [1, 2], [-np.inf, 0.1])
    assert r["rc"] == _abi.LOG_ZERO and r["lp"] == -np.inf
    assert L.oracle_density("poisson_log", [0, 2], [np.inf, 0.1])["rc"] == _abi.LOG_ZERO
    assert L.oracle_density("poisson_log", [0, 2], [-np.inf, 0.1])["rc"] == 0
    assert L.oracle_density("poisson_log", [], [])["lp"] == 0.0


@needs_ref
@pytest.mark.parametrize("density", ["bernoulli_logit", "poisson_log"])
def test_oracle_density_vs_compiled_reference(density):
    rng = np.random.default_rng(3)
    n = 5000
    theta = rng.standard_normal(n) * (4.0 if density == "bernoulli_logit" else 1.5)
    y = (rng.random(n) < 0.4).astype(np.int32) if density == "bernoulli_logit" else rng.poisson(2.0, n).astype(np.int32)
    for propto in (False, True):
        for yy, th in ((y, theta), (y, theta[:1]), (y[:1], theta)):
            ref = L.ref_density(density, yy, th, propto=propto)
            got = L.oracle_density(density, yy, th, propto=propto)
            assert got["rc"] == 0
            assert_close(got["lp"], ref["lp"], "lp", 1e-13)
            assert_close(got["dtheta"], ref["dtheta"], "dtheta", 1e-13)
    for bad in (([2] if density == "bernoulli_logit" else [-1], [0.1]), ([1], [np.nan])):
        with pytest.raises(L.RefError) as e:
            L.ref_density(density, *bad)
        assert e.value.rc == 1


@needs_ref
def test_reference_error_contract():
    """What the reference throws (the boundary must reproduce it): sizes -> invalid_argument (2),
    domain violations -> domain_error (1)."""
    x = np.ones((3, 2))
    with pytest.raises(L.RefError) as e:
        L.ref_glm("bernoulli_logit", x, [1, 0, 23], 0.3, [0.1, 0.2])
    assert e.value.rc == 1
    with pytest.raises(L.RefError) as e:
        L.ref_glm("poisson_log", x, [1, 0, -1], 0.3, [0.1, 0.2])
    assert e.value.rc == 1
    with pytest.raises(L.RefError) as e:
        L.ref_glm("normal_id", x, [1.0, 0.0, 2.0], 0.3, [0.1, 0.2], [-1.0])
    assert e.value.rc == 1
    with pytest.raises(L.RefError) as e:
        L.ref_glm("poisson_log", x, [1, 0, 1], 0.3, [0.1, np.inf])
    assert e.value.rc == 1 and "Weight vector" in str(e.value)
    xi = x.copy()
    xi[0, 1] = np.nan  # row 0: the reference's lazy check only looks at the first row (see below)
    with pytest.raises(L.RefError) as e:
        L.ref_glm("bernoulli_logit", xi, [1, 0, 1], 0.3, [0.1, 0.2])
    assert e.value.rc == 1 and "Matrix of independent variables" in str(e.value)
    # quirk: check_finite on the Eigen::Array ytheta inspects ONE element, so a NaN in another row
    # comes back as a NaN log-density without any exception
    xi = x.copy()
    xi[1, 1] = np.nan
    r = L.ref_glm("bernoulli_logit", xi, [1, 0, 1], 0.3, [0.1, 0.2])
    assert np.isnan(r["lp"])


@needs_ref
def test_reference_map_rect_baseline_matches_plain_call():
    """The TBB map_rect row-sharded baseline computes the same value and gradient as one call."""
    # In a process of its own: the baseline starts the reference's TBB scheduler, whose worker threads and the library
    # that holds their code are torn down in no particular order when an interpreter exits (the crash at exit that
    # bench.py --impl reference avoids the same way) -- the test session must not inherit that.
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import json, os, sys\n"
        "sys.path.insert(0, %r)\n"
        "from oracle import loader as L\n"
        "from tests.helpers import synth_glm\n"
        "s = synth_glm('poisson_log', 4000, 16, seed=5)\n"
        "plain = L.ref_glm('poisson_log', s['x'], s['y'], s['alpha'], s['beta'])\n"
        "sh = L.RefShards('poisson_log', s['x'], s['y'], n_shards=8, n_threads=4)\n"
        "r = sh.eval(s['beta'], alpha=s['alpha'])\n"
        "sh.close()\n"
        "out = {k: [float(v) for v in __import__('numpy').atleast_1d(d[k])] for d, k in ((r, 'lp'), (r, 'dbeta'), (r, 'dalpha'))}\n"
        "ref = {k: [float(v) for v in __import__('numpy').atleast_1d(plain[k])] for k in ('lp', 'dbeta', 'dalpha')}\n"
        "print(json.dumps({'got': out, 'ref': ref}))\n"
        "sys.stdout.flush()\n"
        "os._exit(0)\n" % root)
    run = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, timeout=600)
    assert run.returncode == 0, run.stderr[-2000:]
    res = json.loads(run.stdout.strip().splitlines()[-1])
    assert_close(res["got"]["lp"], res["ref"]["lp"], "map_rect lp", 1e-12)
    assert_close(res["got"]["dbeta"], res["ref"]["dbeta"], "map_rect dbeta", 1e-11)
    assert_close(res["got"]["dalpha"], res["ref"]["dalpha"], "map_rect dalpha", 1e-11)
