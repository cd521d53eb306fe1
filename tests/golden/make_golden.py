"""Generates tests/golden/glm_golden.npz from the COMPILED REFERENCE (oracle/_ref/libstanref.so).

Run in the dev container (needs /root/reference to have built the shim):
    python tests/golden/make_golden.py
Cases:
  * the fixed fixtures of the reference's own GLM unit tests
    (test/unit/math/rev/mat/prob/{bernoulli_logit,poisson_log,normal_id}_glm_*_test.cpp)
  * the extreme-value bernoulli fixture exercising the cutoff branches (bernoulli..._test.cpp:63-108)
  * vector-alpha / vector-sigma / var-y variants, propto on and off
  * a 153 x 71 random case per family (the size of the reference's OpenCL-vs-CPU "big" test,
    test/unit/math/opencl/rev/bernoulli_logit_glm_lpmf_test.cpp:255-297) and a 2000 x 128 one
  * normal_lpdf: the reference's known-answer triples (test/prob/normal/normal_test.hpp:11-40)
    plus a vectorised case
Every case stores inputs and the reference's outputs (value + all adjoints).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402

out = {}
names = []


def add_glm(name, fam, x, y, alpha, beta, sigma=None, propto=False, need_dy=False, xkey=None):
    r = L.ref_glm(fam, x, y, alpha, beta, sigma, propto=propto, need_dy=need_dy)
    p = "glm/%s/" % name
    xkey = xkey or name
    out[p + "xkey"] = np.array(xkey)   # design matrices are stored once and shared between cases
    out[p + "family"] = np.array(L._abi.FAMILIES[fam])
    out[p + "propto"] = np.array(int(propto))
    out[p + "need_dy"] = np.array(int(need_dy))
    out["x/" + xkey] = np.asarray(x, float)
    out[p + "y"] = np.asarray(y, float if fam == "normal_id" else np.int32)
    out[p + "alpha"] = np.atleast_1d(np.asarray(alpha, float))
    out[p + "beta"] = np.atleast_1d(np.asarray(beta, float))
    if sigma is not None:
        out[p + "sigma"] = np.atleast_1d(np.asarray(sigma, float))
    out[p + "lp"] = np.array(r["lp"])
    out[p + "dbeta"] = r["dbeta"]
    out[p + "dalpha"] = r["dalpha"]
    if "dsigma" in r:
        out[p + "dsigma"] = r["dsigma"]
    if need_dy:
        out[p + "dy"] = r["dy"]
    names.append("glm/" + name)


X = np.array([[-12, 46], [-42, 24], [25, 27]], float)
BETA, ALPHA = [0.3, 2.0], 0.3
for propto in (False, True):
    s = "_propto" if propto else ""
    add_glm("bern_fixture" + s, "bernoulli_logit", X, [1, 0, 1], ALPHA, BETA, propto=propto)
    add_glm("bern_fixture2" + s, "bernoulli_logit", X, [0, 1, 0], ALPHA, BETA, propto=propto)
    add_glm("pois_fixture" + s, "poisson_log", X * 0.01, [15, 3, 5], ALPHA, BETA, propto=propto)
    add_glm("pois_fixture2" + s, "poisson_log", X * 0.01, [14, 2, 5], ALPHA, BETA, propto=propto)
    add_glm("norm_fixture" + s, "normal_id", X, [51., 32., 12.], ALPHA, BETA, [10.0], propto=propto)
    add_glm("norm_fixture2" + s, "normal_id", X, [14., 32., 21.], ALPHA, BETA, [10.0], propto=propto)
# cutoff branches: |ytheta| > 20 on both sides
Xe = np.array([[-1234, 46], [-42, 24], [25, 27]], float)
add_glm("bern_extreme", "bernoulli_logit", Xe, [1, 0, 1], 0.3, [0.3, 2000.0])
add_glm("bern_extreme2", "bernoulli_logit", Xe, [0, 1, 0], 0.3, [0.3, 2000.0])

rng = np.random.default_rng(20191104)
for n, k, tag in ((153, 71, "big"), (600, 128, "wide"), (1000, 1, "k1"), (37, 16, "k16"), (300, 300, "k300")):
    x = rng.standard_normal((n, k))
    bt = rng.standard_normal(k) * (0.6 / np.sqrt(k))
    beta = bt + rng.standard_normal(k) * 0.05
    eta = 0.2 + x @ bt
    add_glm("bern_" + tag, "bernoulli_logit", x, (rng.random(n) < 1 / (1 + np.exp(-eta))).astype(np.int32), 0.3, beta, xkey=tag)
    add_glm("pois_" + tag, "poisson_log", x, rng.poisson(np.exp(eta)), 0.3, beta, xkey=tag)
    yn = eta + 0.7 * rng.standard_normal(n)
    add_glm("norm_" + tag, "normal_id", x, yn, 0.3, beta, [1.0], xkey=tag)
    if tag == "big":
        av = 0.3 + 0.1 * rng.standard_normal(n)
        sv = 0.5 + rng.random(n)
        add_glm("bern_big_valpha", "bernoulli_logit", x, (rng.random(n) < 0.5).astype(np.int32), av, beta, xkey=tag)
        add_glm("pois_big_valpha", "poisson_log", x, rng.poisson(np.exp(eta)), av, beta, xkey=tag)
        add_glm("norm_big_valpha_vsigma", "normal_id", x, yn, av, beta, sv, xkey=tag)
        add_glm("norm_big_vsigma_dy", "normal_id", x, yn, 0.3, beta, sv, need_dy=True, xkey=tag)
        add_glm("norm_big_dy_propto", "normal_id", x, yn, av, beta, [1.3], propto=True, need_dy=True, xkey=tag)

nl = []


def add_normal(name, y, mu, sigma, propto=False, need_dy=False, value_only=None):
    p = "normal/%s/" % name
    r = L.ref_normal_lpdf(y, mu, sigma, propto=propto, need_dy=need_dy)
    out[p + "y"], out[p + "mu"], out[p + "sigma"] = (np.atleast_1d(np.asarray(v, float)) for v in (y, mu, sigma))
    out[p + "propto"], out[p + "need_dy"] = np.array(int(propto)), np.array(int(need_dy))
    out[p + "lp"], out[p + "dmu"], out[p + "dsigma"] = np.array(r["lp"]), r["dmu"], r["dsigma"]
    if need_dy:
        out[p + "dy"] = r["dy"]
    if value_only is not None:
        out[p + "known_answer"] = np.array(value_only)  # literal from the reference's test file
    nl.append("normal/" + name)


# test/prob/normal/normal_test.hpp:11-40 (literal expected log-densities)
add_normal("ka0", [0.0], [0.0], [1.0], value_only=-0.918938533204672669541)
add_normal("ka1", [1.0], [0.0], [1.0], value_only=-1.418938533204672669541)
add_normal("ka2", [-2.0], [0.0], [1.0], value_only=-2.918938533204672669541)
add_normal("ka3", [-3.5], [1.9], [7.2], value_only=-3.174269559226682080322)
yv = rng.standard_normal(1000)
add_normal("vec_mu", yv, 0.1 * rng.standard_normal(1000), [1.3])
add_normal("vec_mu_propto", yv, 0.1 * rng.standard_normal(1000), [1.3], propto=True)
add_normal("vec_all_dy", yv, 0.1 * rng.standard_normal(1000), 0.5 + rng.random(1000), need_dy=True)
add_normal("scalar_mu", yv, [0.25], [0.8])

# neg_binomial_2_log_glm (added after everything above so the earlier draws are unchanged):
# the fixed fixture of test/unit/math/rev/mat/prob/neg_binomial_2_log_glm_lpmf_test.cpp:13-30
# (x = [-12 46; -42 24; 25 27], beta = (0.3, 2), alpha = 0.3, phi = 2, y = {1, 0, 1}) scaled like
# the poisson fixture, large counts (% 200 in the reference's random test), vector alpha / phi
rng2 = np.random.default_rng(20191105)
for propto in (False, True):
    s = "_propto" if propto else ""
    add_glm("nb_fixture" + s, "neg_binomial_2_log", X * 0.01, [1, 0, 1], ALPHA, BETA, [2.0], propto=propto)
    add_glm("nb_fixture2" + s, "neg_binomial_2_log", X * 0.01, [14, 2, 5], ALPHA, BETA, [2.0], propto=propto)
for n, k, tag in ((153, 71, "nbbig"), (600, 128, "nbwide"), (1000, 1, "nbk1")):
    x = rng2.standard_normal((n, k))
    bt = rng2.standard_normal(k) * (0.6 / np.sqrt(k))
    beta = bt + rng2.standard_normal(k) * 0.05
    eta = 0.2 + x @ bt
    yc = rng2.negative_binomial(3.0, 3.0 / (3.0 + np.exp(eta)))
    add_glm("nb_" + tag, "neg_binomial_2_log", x, yc, 0.3, beta, [2.5], xkey=tag)
    if tag == "nbbig":
        av = 0.3 + 0.1 * rng2.standard_normal(n)
        pv = 0.5 + 4 * rng2.random(n)
        add_glm("nb_big_valpha", "neg_binomial_2_log", x, yc, av, beta, [0.7], xkey=tag)
        add_glm("nb_big_vphi", "neg_binomial_2_log", x, yc, 0.3, beta, pv, xkey=tag)
        add_glm("nb_big_valpha_vphi_propto", "neg_binomial_2_log", x, yc, av, beta, pv, propto=True, xkey=tag)
        add_glm("nb_big_counts", "neg_binomial_2_log", x, rng2.integers(0, 200, n), 0.3, beta, [1.5], xkey=tag)

# ordered_logistic_glm (own key space "ord/"): the fixed fixture of
# test/unit/math/rev/mat/prob/ordered_logistic_glm_lpmf_test.cpp:33-47 (y = {1,4,3,3,2}, cuts = (0.9, 1.1, 7),
# beta = (1.1, 0.4), x = [1 2; 3 4; 5 6; 7 8; 9 0]), a single-cut case, and random cases up to 10 classes (:293-331)
ord_names = []


def add_ord(name, x, y, beta, cuts, propto=False):
    r = L.ref_ordered_logistic_glm(x, y, beta, cuts, propto=propto)
    p = "ord/%s/" % name
    out[p + "x"], out[p + "y"] = np.asarray(x, float), np.asarray(y, np.int32)
    out[p + "beta"], out[p + "cuts"] = np.atleast_1d(np.asarray(beta, float)), np.atleast_1d(np.asarray(cuts, float))
    out[p + "propto"] = np.array(int(propto))
    out[p + "lp"], out[p + "dbeta"], out[p + "dcuts"] = np.array(r["lp"]), r["dbeta"], r["dcuts"]
    ord_names.append(name)


rng3 = np.random.default_rng(20191106)
XO = np.array([[1, 2], [3, 4], [5, 6], [7, 8], [9, 0]], float)
add_ord("fixture", XO, [1, 4, 3, 3, 2], [1.1, 0.4], [0.9, 1.1, 7])
add_ord("fixture_propto", XO, [1, 4, 3, 3, 2], [1.1, 0.4], [0.9, 1.1, 7], propto=True)
add_ord("two_classes", XO, [1, 2, 2, 1, 2], [0.11, -0.4], [0.3])
for n, k, c, tag in ((153, 71, 4, "big"), (600, 128, 10, "wide"), (1000, 1, 3, "k1"), (777, 33, 40, "many")):
    x = rng3.standard_normal((n, k))
    beta = rng3.standard_normal(k) * (1.5 / np.sqrt(k))
    cuts = np.sort(rng3.standard_normal(c - 1) * 1.5)
    add_ord("rand_" + tag, x, rng3.integers(1, c + 1, n), beta, cuts)
# far tails: locations +-40 from the cut-points (the branches of :128-131 and :165-171)
x = rng3.standard_normal((64, 2)) * 30
add_ord("tails", x, rng3.integers(1, 4, 64), [1.0, -1.0], [-1.0, 2.0])

# categorical_logit_glm (key space "cat/"): the fixture of
# test/unit/math/rev/mat/prob/categorical_logit_glm_lpmf_test.cpp:551-580 (x 5 x 2, beta 2 x 3, alpha (0.5, -2, 4),
# y = {1, 3, 1, 2, 2}) and random cases up to 16 classes
cat_names = []


def add_cat(name, x, y, alpha, beta, propto=False):
    r = L.ref_categorical_logit_glm(x, y, alpha, beta, propto=propto)
    p = "cat/%s/" % name
    out[p + "x"], out[p + "y"] = np.asarray(x, float), np.asarray(y, np.int32)
    out[p + "alpha"], out[p + "beta"] = np.atleast_1d(np.asarray(alpha, float)), np.asarray(beta, float)
    out[p + "propto"] = np.array(int(propto))
    out[p + "lp"], out[p + "dalpha"], out[p + "dbeta"] = np.array(r["lp"]), r["dalpha"], r["dbeta"]
    cat_names.append(name)


rng4 = np.random.default_rng(20191107)
XC = np.array([[-12, 46], [-42, 24], [25, 27], [-14, -11], [5, 18]], float)
BC = np.array([[0.3, 2, 0.4], [-0.1, -1.3, 1]], float)
add_cat("fixture", XC, [1, 3, 1, 2, 2], [0.5, -2, 4], BC)
add_cat("fixture_propto", XC, [1, 3, 1, 2, 2], [0.5, -2, 4], BC, propto=True)
for n, k, c, tag in ((153, 71, 4, "big"), (400, 128, 10, "wide"), (1000, 1, 3, "k1"), (500, 33, 16, "many"), (300, 7, 2, "two")):
    x = rng4.standard_normal((n, k))
    beta = rng4.standard_normal((k, c)) * (1.5 / np.sqrt(k))
    add_cat("rand_" + tag, x, rng4.integers(1, c + 1, n), rng4.standard_normal(c), beta)

# vectorised elementwise densities (own key space "dens/", own generator: earlier draws unchanged):
# the reference's known-answer pairs (test/prob/bernoulli/bernoulli_logit_test.hpp:11-48,
# test/prob/poisson/poisson_log_test.hpp:11-31) with their literal expected values, the cutoff branches of
# bernoulli_logit_lpmf (|theta| > 20), vector / scalar operands, propto on and off
dens_names = []


def add_dens(name, density, y, theta, propto=False, value_only=None):
    r = L.ref_density(density, y, theta, propto=propto)
    p = "dens/%s/" % name
    out[p + "density"] = np.array(L._abi.DENSITIES[density])
    out[p + "y"] = np.atleast_1d(np.asarray(y, np.int32))
    out[p + "theta"] = np.atleast_1d(np.asarray(theta, float))
    out[p + "propto"] = np.array(int(propto))
    out[p + "lp"], out[p + "dtheta"] = np.array(r["lp"]), r["dtheta"]
    if value_only is not None:
        out[p + "known_answer"] = np.array(value_only)
    dens_names.append(name)


def _logit(q):
    return float(np.log(q / (1.0 - q)))


for i, (yy, th, ka) in enumerate([(1, _logit(0.25), np.log(0.25)), (0, _logit(0.25), np.log(0.75)),
                                  (1, _logit(0.01), np.log(0.01)), (0, _logit(0.01), np.log(0.99)), (0, 25.0, -25.0),
                                  (1, -25.0, -25.0), (0, -25.0, -np.exp(-25.0))]):
    add_dens("bern_ka%d" % i, "bernoulli_logit", [yy], [th], value_only=ka)
for i, (yy, al, ka) in enumerate([(17, np.log(13.0), -2.900934373290765311282), (192, np.log(42.0), -145.3546649655311853166),
                                  (0, np.log(3.0), -3.0)]):
    add_dens("pois_ka%d" % i, "poisson_log", [yy], [al], value_only=ka)
rng5 = np.random.default_rng(20191108)
th5 = np.concatenate([rng5.standard_normal(2000) * 3, [25.0, -25.0, 21.0, -21.0, 19.99, -19.99, 700.0, -700.0]])
y5 = (rng5.random(th5.size) < 0.5).astype(np.int32)
c5 = rng5.poisson(3.0, th5.size).astype(np.int32)
al5 = np.concatenate([rng5.standard_normal(2000) * 1.5, [5.0, -5.0, 20.0, -30.0, 0.0, 1e-3, -745.0, 6.5]])
for propto in (False, True):
    s = "_propto" if propto else ""
    add_dens("bern_vec" + s, "bernoulli_logit", y5, th5, propto=propto)
    add_dens("pois_vec" + s, "poisson_log", c5, al5, propto=propto)
add_dens("bern_scalar_theta", "bernoulli_logit", y5, [0.3])
add_dens("bern_scalar_n", "bernoulli_logit", [1], th5)
add_dens("pois_scalar_alpha", "poisson_log", c5, [0.7])
add_dens("pois_scalar_n", "poisson_log", [4], al5)
add_dens("pois_minus_inf_zero_count", "poisson_log", [0, 2, 0], [-np.inf, 0.1, 0.4])

out["glm_cases"] = np.array(names)
out["dens_cases"] = np.array(dens_names)
out["cat_cases"] = np.array(cat_names)
out["ord_cases"] = np.array(ord_names)
out["normal_cases"] = np.array(nl)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "glm_golden.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(names), "GLM cases,", len(nl), "normal cases,", os.path.getsize(path), "bytes")
