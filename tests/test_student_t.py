"""student_t_lpdf (stan/math/prim/scal/prob/student_t_lpdf.hpp:55-201).

CPU: the oracle restatement against the golden vectors of the compiled reference and the literals of the reference's
own fixture (test/prob/student_t/student_t_test.hpp:13-60), and against the compiled reference itself on fresh
inputs when oracle/_ref is present.  GPU: the CUDA path through the C ABI against the same goldens, against the
oracle on seeded inputs, and at large n through properties that need no checker (shift / scale invariance,
additivity over a split of the rows, the normal limit for nu -> inf).
"""
import numpy as np
import pytest

from math_b200 import _abi
from oracle import loader as L
from tests.helpers import assert_close, check_student_t_result, student_t_case


def _cases(golden_st):
    return [str(n) for n in golden_st["cases"]]


# ------------------------------------------------------------------------------------------ CPU
def test_oracle_matches_golden_and_known_answers(golden_st):
    assert len(_cases(golden_st)) >= 70
    for name in _cases(golden_st):
        c = student_t_case(golden_st, name)
        r = L.oracle_student_t_lpdf(c["y"], c["nu"], c["mu"], c["sigma"], propto=c["propto"], var_mask=c["var_mask"])
        assert r["rc"] == 0, name
        check_student_t_result(r, c, "oracle vs golden " + name, rtol=1e-12)
        if "known_answer" in c:  # literals of test/prob/student_t/student_t_test.hpp:13-30
            assert abs(r["lp"] - c["known_answer"]) < 1e-14 * max(1, abs(c["known_answer"]))
            assert abs(c["lp"] - c["known_answer"]) < 1e-14 * max(1, abs(c["known_answer"]))


# test/prob/student_t/student_t_test.hpp:33-60: index of the operand, invalid value
INVALID = [(1, 0.0), (1, -1.0), (1, np.inf), (1, -np.inf), (2, np.inf), (2, -np.inf), (3, 0.0), (3, -1.0), (3, np.inf),
           (3, -np.inf), (0, np.nan), (1, np.nan), (2, np.nan), (3, np.nan)]


@pytest.mark.parametrize("index,value", INVALID)
def test_oracle_rejects_what_the_reference_rejects(index, value):
    ops = [[1.0, 0.5], [1.0, 3.0], [0.0, 0.2], [1.0, 2.0]]
    ops[index][1] = value
    assert L.oracle_student_t_lpdf(*ops, var_mask=15)["rc"] == _abi.BAD_OPERAND
    if L.have_ref():
        with pytest.raises(L.RefError):
            L.ref_student_t_lpdf(*ops, var_mask=15)


def test_oracle_infinite_outcome_is_not_an_error():
    """check_not_nan(y) only (:67): y = inf gives lp = -inf, not an exception."""
    r = L.oracle_student_t_lpdf([np.inf], [3.0], [0.0], [1.0], var_mask=14)
    assert r["rc"] == _abi.NONFINITE and r["lp"] == -np.inf
    if L.have_ref():
        assert L.ref_student_t_lpdf([np.inf], [3.0], [0.0], [1.0], var_mask=14)["lp"] == -np.inf


@pytest.mark.skipif(not L.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_matches_compiled_reference_on_fresh_inputs():
    rng = np.random.default_rng(77)
    for trial in range(12):
        n = int(rng.integers(1, 400))
        y, mu = 4 * rng.standard_normal(n), rng.standard_normal(n)
        nu, sg = 10.0 ** rng.uniform(-2, 4, n), 10.0 ** rng.uniform(-2, 2, n)
        vm, propto = int(rng.integers(0, 16)), bool(trial & 1)
        ops = [v if rng.random() < 0.6 else v[:1] for v in (y, nu, mu, sg)]
        ref = L.ref_student_t_lpdf(*ops, propto=propto, var_mask=vm)
        got = L.oracle_student_t_lpdf(*ops, propto=propto, var_mask=vm)
        c = dict(zip(("y", "nu", "mu", "sigma"), ops), **ref)
        check_student_t_result(got, c, "oracle vs reference trial %d" % trial, rtol=1e-12)


def test_empty_operand_is_zero():
    assert L.oracle_student_t_lpdf([], [1.0], [0.0], [1.0])["lp"] == 0.0


# ------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def glm():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from math_b200 import glm as g
    g.init(1)
    return g


def _device(glm, c_or_ops, propto, var_mask, **kw):
    ops = c_or_ops if isinstance(c_or_ops, (list, tuple)) else [c_or_ops[k] for k in ("y", "nu", "mu", "sigma")]
    return glm.student_t_eval(*ops, L.student_t_flags(propto, var_mask), **kw)


@pytest.mark.gpu
def test_device_matches_golden(glm, golden_st):
    failures = []
    for name in _cases(golden_st):
        c = student_t_case(golden_st, name)
        r = _device(glm, c, c["propto"], c["var_mask"])
        assert r["rc"] == 0, name
        try:
            check_student_t_result(r, c, "device vs golden " + name)
        except AssertionError as e:
            failures.append(str(e))
    assert not failures, "\n".join(failures)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 255, 256, 257, 4097, 100_003, 1_000_000])
def test_device_matches_oracle_on_seeded_inputs(glm, n):
    rng = np.random.default_rng(n)
    y, mu = 3 * rng.standard_normal(n), rng.standard_normal(n)
    nu, sg = rng.uniform(0.5, 50, n), rng.uniform(0.3, 3, n)
    for ops, vm, propto in (((y, nu, mu, sg), 15, False), ((y, nu[:1], mu, sg[:1]), 14, False),
                            ((y, nu[:1], mu[:1], sg[:1]), 14, True), ((y[:1], nu, mu, sg), 6, True)):
        ref = L.oracle_student_t_lpdf(*ops, propto=propto, var_mask=vm)
        got = _device(glm, ops, propto, vm, y_is_data=False)
        assert got["rc"] == ref["rc"] == 0
        c = dict(zip(("y", "nu", "mu", "sigma"), ops), **ref)
        check_student_t_result(got, c, "device vs oracle n=%d mask=%d" % (n, vm))


@pytest.mark.gpu
@pytest.mark.parametrize("index,value", INVALID)
def test_device_rejects_what_the_reference_rejects(glm, index, value):
    n = 1000
    ops = [np.linspace(-2, 2, n), np.full(n, 3.0), np.zeros(n), np.ones(n)]
    ops[index][n - 7] = value
    assert _device(glm, ops, False, 15, y_is_data=False)["rc"] == _abi.BAD_OPERAND
    with pytest.raises(glm.DomainError):
        glm.student_t_lpdf(*ops, y_var=True)
    sc = [np.linspace(-2, 2, n), [3.0], [0.0], [1.0]]
    if index > 0:   # the same value in a scalar operand
        sc[index] = [value]
        assert _device(glm, sc, False, 14, y_is_data=False)["rc"] == _abi.BAD_OPERAND


@pytest.mark.gpu
def test_host_call_sequence(glm):
    y = np.linspace(-1, 2, 10)   # not symmetric about mu: the sum of d/dmu must not cancel to rounding noise
    assert glm.student_t_lpdf([], [1.0], [0.0], [1.0]) == (0.0, {})
    with pytest.raises(ValueError, match="dimension"):
        glm.student_t_lpdf(y, [1.0, 2.0], [0.0], [1.0])
    with pytest.raises(glm.DomainError):   # domain error before the size error (:67-73)
        glm.student_t_lpdf(y, [-1.0, 2.0], [0.0], [1.0])
    lp, g = glm.student_t_lpdf(y, [3.0], [0.0], [1.0], propto=True, nu_var=False, mu_var=False, sigma_var=False)
    assert lp == 0.0 and g == {}
    lp, g = glm.student_t_lpdf(y, [3.0], [0.0], [1.0])
    ref = L.oracle_student_t_lpdf(y, [3.0], [0.0], [1.0])
    assert_close(lp, ref["lp"], "lp")
    assert set(g) == {"nu", "mu", "sigma"}
    assert_close(g["mu"], ref["dmu"], "dmu")
    lp_inf, _ = glm.student_t_lpdf([np.inf], [3.0], [0.0], [1.0])
    assert lp_inf == -np.inf


@pytest.mark.gpu
def test_cached_outcome_is_reused_and_results_repeat_bitwise(glm):
    rng = np.random.default_rng(5)
    n = 300_000
    y, mu = rng.standard_normal(n), 0.1 * rng.standard_normal(n)
    flags = L.student_t_flags(False, 14)
    glm.cache_clear()
    a = glm.student_t_eval(y, [4.0], mu, [1.2], flags)
    up0 = glm.stats()["bytes_uploaded"]
    b = glm.student_t_eval(y, [4.0], mu, [1.2], flags)
    up1 = glm.stats()["bytes_uploaded"]
    assert a["lp"] == b["lp"] and np.array_equal(a["dmu"], b["dmu"]) and a["dnu"][0] == b["dnu"][0]
    assert up1 - up0 < y.nbytes   # y not re-sent (mu is uploaded per call and not counted as cached data)


@pytest.mark.gpu
def test_full_size_properties(glm):
    """n = 10^7 (C1's row count): no checker needed.  (1) the density depends on (y - mu) / sigma only: shifting y and
    mu together changes nothing; (2) additivity over a split of the rows; (3) nu -> inf is normal_lpdf;
    (4) sum of d/dy = - sum of d/dmu."""
    rng = np.random.default_rng(9)
    n = 10_000_000
    y, mu = rng.standard_normal(n), 0.2 * rng.standard_normal(n)
    f = L.student_t_flags(False, 15)
    full = glm.student_t_eval(y, [5.0], mu, [1.5], f, y_is_data=False)
    assert full["rc"] == 0
    h = n // 2
    a = glm.student_t_eval(y[:h], [5.0], mu[:h], [1.5], f, y_is_data=False)
    b = glm.student_t_eval(y[h:], [5.0], mu[h:], [1.5], f, y_is_data=False)
    assert_close(a["lp"] + b["lp"], full["lp"], "additivity lp", 1e-12)
    assert_close(a["dnu"] + b["dnu"], full["dnu"], "additivity dnu", 1e-9)
    assert_close(a["dsigma"] + b["dsigma"], full["dsigma"], "additivity dsigma", 1e-11)
    assert np.array_equal(np.concatenate([a["dmu"], b["dmu"]]), full["dmu"])
    assert_close(-np.sum(full["dy"]), np.sum(full["dmu"]), "sum dy = -sum dmu", 1e-9)
    shifted = glm.student_t_eval(y + 8.0, [5.0], mu + 8.0, [1.5], f, y_is_data=False)
    assert_close(shifted["lp"], full["lp"], "shift invariance", 1e-9)
    # the limit holds term by term; the lgamma terms are left out (at nu = 1e12 they are a difference of two 1e13's)
    f_main = _abi.LP_SCALE | _abi.LP_MAIN | _abi.NEED_DALPHA
    big_nu = glm.student_t_eval(y, [1e12], mu, [1.5], f_main, y_is_data=False)
    normal = glm.normal_lpdf(y, mu, [1.5], propto=True, y_is_data=False)
    assert_close(big_nu["lp"], normal[0], "normal limit lp", 1e-8)
    assert_close(big_nu["dmu"], normal[1]["mu"], "normal limit dmu", 1e-8)
