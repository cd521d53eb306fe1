"""CPU tests (no GPU) of the host-side mirror of the reference interface (math_b200/glm.py): the call sequences that
end BEFORE any device work -- size_zero, the reference's order of domain and size errors, propto with nothing to
compute -- and, where the compiled reference is available, that it reacts the same way."""
import numpy as np
import pytest

from math_b200 import glm
from oracle import loader as L

needs_ref = pytest.mark.skipif(not L.have_ref(), reason="oracle/_ref/libstanref.so not built")


@pytest.mark.parametrize("f,kw", [(glm.bernoulli_logit_lpmf, "theta_var"), (glm.poisson_log_lpmf, "alpha_var"),
                                  (glm.poisson_lpmf, "lambda_var"), (glm.bernoulli_lpmf, "theta_var")])
def test_density_paths_that_need_no_device(f, kw):
    th = np.linspace(-1, 1, 10) if f in (glm.bernoulli_logit_lpmf, glm.poisson_log_lpmf) else np.linspace(0.1, 0.9, 10)
    ones = np.ones(10, dtype=np.int32)
    assert f([], [])[0] == 0.0                                           # size_zero(n, theta)
    assert f(ones, [])[0] == 0.0
    assert f(ones, th, propto=True, **{kw: False})[0] == 0.0            # include_summand<propto, double> is false
    with pytest.raises(ValueError) as e:                                # sizes: std::invalid_argument ...
        f(ones[:7], th)
    assert not isinstance(e.value, glm.DomainError)
    bad = ones[:7].copy()
    bad[3] = -1
    with pytest.raises(glm.DomainError):                                # ... but a domain error comes first
        f(bad, th)
    nan = th.copy()
    nan[2] = np.nan
    with pytest.raises(glm.DomainError):
        f(ones[:7], nan)
    with pytest.raises(glm.DomainError):                                # checks run even when nothing is computed
        f(bad, th[:7], propto=True, **{kw: False})
    if f in (glm.poisson_lpmf, glm.bernoulli_lpmf):                     # the operand's own range check (rate >= 0, probability in [0, 1])
        neg = th.copy()
        neg[4] = -0.5
        with pytest.raises(glm.DomainError):
            f(ones, neg, propto=True, **{kw: False})


@needs_ref
@pytest.mark.parametrize("density", ["bernoulli_logit", "poisson_log", "poisson", "bernoulli"])
def test_reference_reacts_the_same_way(density):
    th = np.linspace(-1, 1, 10) if density.endswith(("logit", "log")) else np.linspace(0.1, 0.9, 10)
    ones = np.ones(10, dtype=np.int32)
    with pytest.raises(L.RefError) as e:
        L.ref_density(density, ones[:7], th)
    assert e.value.rc == 2                                               # std::invalid_argument
    bad = ones[:7].copy()
    bad[3] = -1
    with pytest.raises(L.RefError) as e:
        L.ref_density(density, bad, th)
    assert e.value.rc == 1                                               # std::domain_error first
    assert L.ref_density(density, ones, th, propto=True, theta_var=False)["lp"] == 0.0


def test_glm_wrappers_size_and_zero_paths_need_no_device():
    x = np.ones((3, 2))
    for f in (glm.bernoulli_logit_glm_lpmf, glm.poisson_log_glm_lpmf):
        with pytest.raises(ValueError):
            f([1, 0], x, 0.3, [0.1, 0.2])                                # y has 2 entries, x 3 rows
        with pytest.raises(ValueError):
            f([1, 0, 1], x, 0.3, [0.1, 0.2, 0.3])                        # beta has 3 entries, x 2 columns
        with pytest.raises(ValueError):
            f([1, 0, 1], x, [0.3, 0.1], [0.1, 0.2])                      # vector alpha of the wrong length
