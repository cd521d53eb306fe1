"""Generates tests/golden/locscale_golden.npz from the UNMODIFIED reference (oracle/_ref/libstanref.so, compiled from
/root/reference by oracle/Makefile): cauchy_lpdf(y | mu, sigma), lognormal_lpdf(y | mu, sigma), gamma_lpdf(y | alpha, beta) and
exponential_lpdf(y | beta), value + reverse-mode gradient, for
  * the reference's known-answer triples (test/prob/cauchy/cauchy_test.hpp:10-31,
    test/prob/lognormal/lognormal_test.hpp:10-25; the literal expected log-densities stored beside the output),
  * every subset of {y, mu, sigma} as autodiff variables x propto on one seeded vector case (the summands the
    reference drops under propto depend on which operands are variables, cauchy_lpdf.hpp:87-95,
    lognormal_lpdf.hpp:92-118),
  * every mix of scalar / vector operands,
  * the corners: outcomes far in the tails, tiny and huge scales, lognormal outcomes from 1e-300 to 1e300.
Run here (the container with /root/reference):  python tests/golden/make_locscale_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402

out, names = {}, []


def add(name, density, y, mu, sigma, propto=False, var_mask=0b110, known_answer=None):
    p = "ls/%s/" % name
    r = L.ref_locscale_lpdf(density, y, mu, sigma, propto=propto, var_mask=var_mask)
    out[p + "density"] = np.array(L._abi.LOCSCALES[density])
    for key, v in (("y", y), ("mu", mu), ("sigma", sigma)):
        out[p + key] = np.asarray(v, float).reshape(-1)
    out[p + "propto"], out[p + "var_mask"], out[p + "lp"] = np.array(int(propto)), np.array(int(var_mask)), np.array(r["lp"])
    for key in ("dy", "dmu", "dsigma"):
        if key in r:
            out[p + key] = np.asarray(r[key], float)
    if known_answer is not None:
        out[p + "known_answer"] = np.array(known_answer)
    names.append(name)


for i, (y, mu, sg, ka) in enumerate([(1.0, 0.0, 1.0, -1.837877066409345339082), (-1.5, 0.0, 1.0, -2.323384882191046330036),
                                     (-1.5, -1.0, 1.0, -1.367873437163609873224)]):
    add("cauchy_ka%d" % i, "cauchy", [y], [mu], [sg], var_mask=0b111, known_answer=ka)
for i, (y, mu, sg, ka) in enumerate([(1.2, 0.3, 1.5, -1.509802579439102343528), (12.0, 3.0, 0.9, -3.462263160811773765602)]):
    add("lognormal_ka%d" % i, "lognormal", [y], [mu], [sg], var_mask=0b111, known_answer=ka)

rng = np.random.default_rng(20191110)
n = 400
z = np.concatenate([rng.standard_normal(n) * 2.0, [1e3, -1e3, 1e8, -1e-8, 0.0, 300.0]])
mu = np.concatenate([rng.standard_normal(n) * 0.7, [0.5, -0.5, 1.0, 0.0, 0.0, -300.0]])
sg = np.concatenate([rng.uniform(0.3, 3.0, n), [1e-3, 1e3, 2.0, 1e-8, 1e8, 0.7]])
yl = np.concatenate([np.exp(rng.standard_normal(n) * 1.5), [1e-300, 1e300, 1.0, 1e-8, 1e8, 5.0]])
mul = np.concatenate([rng.standard_normal(n) * 0.7, [-600.0, 600.0, 0.0, -10.0, 10.0, 1.0]])
for density, yy, mm in (("cauchy", z, mu), ("lognormal", yl, mul)):
    for var_mask in range(8):
        for propto in (False, True):
            add("%s_vec_m%d%s" % (density, var_mask, "_propto" if propto else ""), density, yy, mm, sg, propto=propto,
                var_mask=var_mask)
    # (scalar operands against the seeded part only: next to the corner entries they overflow the partials)
    add(density + "_scalar_mu_sigma", density, yy[:n], [0.2], [1.3], var_mask=0b111)
    add(density + "_scalar_y", density, yy[:1] if density == "cauchy" else [2.5], mm[:n], sg[:n], var_mask=0b111)
    add(density + "_scalar_sigma", density, yy[:n], mm[:n], [0.8], var_mask=0b110)
    add(density + "_scalar_mu", density, yy[:n], [-0.4], sg[:n], var_mask=0b011)

# gamma_lpdf(y | alpha, beta) (alpha in the mu slot, beta in the sigma slot) and exponential_lpdf(y | beta) (beta in the
# sigma slot, the mu slot holds one zero): known answers of test/prob/gamma/gamma_test.hpp:14-31 and
# test/prob/exponential/exponential_test.hpp:14-28, every subset of variables x propto, scalar / vector operands, corners
for i, (y, al, be, ka) in enumerate([(1.0, 2.0, 2.0, -0.6137056388801093165242), (2.0, 0.25, 0.75, -3.379803428230981676705),
                                     (1.0, 1.0, 1.0, -1.0)]):
    add("gamma_ka%d" % i, "gamma", [y], [al], [be], var_mask=0b111, known_answer=ka)
for i, (y, be, ka) in enumerate([(2.0, 1.5, -2.594534891891835393096), (15.0, 3.9, -57.13902344686439249699),
                                 (1e-08, 3.9, 1.3609765141356007)]):
    add("exponential_ka%d" % i, "exponential", [y], [0.0], [be], var_mask=0b101, known_answer=ka)
yg = np.concatenate([rng.gamma(2.0, 1.5, n), [1e-300, 1e-8, 1e8, 0.5, 3.0, 700.0]])
al = np.concatenate([rng.uniform(0.2, 8.0, n), [1e-3, 0.5, 150.0, 1.0, 1.4616321449683623, 30.0]])
be = np.concatenate([rng.uniform(0.3, 3.0, n), [1e-3, 1e3, 1e-6, 2.0, 1.0, 0.05]])
for var_mask in range(8):
    for propto in (False, True):
        add("gamma_vec_m%d%s" % (var_mask, "_propto" if propto else ""), "gamma", yg, al, be, propto=propto, var_mask=var_mask)
        if not var_mask & 2:
            add("exponential_vec_m%d%s" % (var_mask, "_propto" if propto else ""), "exponential", yg, [0.0], be, propto=propto,
                var_mask=var_mask)
add("gamma_scalar_alpha_beta", "gamma", yg[:n], [2.2], [1.3], var_mask=0b111)
add("gamma_scalar_y", "gamma", [2.5], al[:n], be[:n], var_mask=0b111)
add("gamma_scalar_beta", "gamma", yg[:n], al[:n], [0.8], var_mask=0b110)
add("gamma_zero_outcome", "gamma", [0.0, 1.0, 2.0], [2.0, 0.5, 1.0], [1.0], var_mask=0b110)   # log_y stays 0 (:86-90)
add("exponential_scalar_beta", "exponential", yg[:n], [0.0], [1.3], var_mask=0b101)
add("exponential_scalar_y", "exponential", [2.5], [0.0], be[:n], var_mask=0b101)

# beta_lpdf(y | alpha, beta) and weibull_lpdf(y | alpha, sigma) (alpha in the mu slot): known answers of
# test/prob/beta/beta_test.hpp:14-25 and test/prob/weibull/weibull_test.hpp:14-31, every subset of variables x propto,
# scalar / vector operands, corners
for i, (y, a, b, ka) in enumerate([(0.4, 0.5, 0.5, -0.4311717080293276382896), (0.5, 2.0, 5.0, -0.06453852113757105324332)]):
    add("beta_ka%d" % i, "beta", [y], [a], [b], var_mask=0b111, known_answer=ka)
for i, (y, a, sg, ka) in enumerate([(2.0, 1.0, 1.0, -2.0), (0.25, 2.9, 1.8, -3.277093769205283724233),
                                    (3.9, 1.7, 0.25, -102.8962074392704266756)]):
    add("weibull_ka%d" % i, "weibull", [y], [a], [sg], var_mask=0b111, known_answer=ka)
yb = np.concatenate([rng.beta(2.0, 3.0, n), [1e-12, 1 - 1e-12, 0.5, 1e-3, 0.999, 0.25]])
ab = np.concatenate([rng.uniform(0.3, 6.0, n), [0.01, 50.0, 1.0, 2.0, 300.0, 0.5]])
bb = np.concatenate([rng.uniform(0.3, 6.0, n), [50.0, 0.01, 1.0, 0.5, 200.0, 2.0]])
yw = np.concatenate([rng.weibull(1.5, n) * 2.0, [1e-8, 30.0, 1.0, 0.5, 1e-3, 5.0]])
aw = np.concatenate([rng.uniform(0.4, 4.0, n), [0.1, 2.0, 1.0, 8.0, 0.5, 1.5]])
sw = np.concatenate([rng.uniform(0.5, 3.0, n), [1.0, 10.0, 1.0, 0.6, 100.0, 2.0]])
for density, yy, aa, ss in (("beta", yb, ab, bb), ("weibull", yw, aw, sw)):
    for var_mask in range(8):
        for propto in (False, True):
            add("%s_vec_m%d%s" % (density, var_mask, "_propto" if propto else ""), density, yy, aa, ss, propto=propto,
                var_mask=var_mask)
    add(density + "_scalar_shapes", density, yy[:n], [1.7], [2.3], var_mask=0b111)
    add(density + "_scalar_y", density, [0.3], aa[:n], ss[:n], var_mask=0b111)
    add(density + "_scalar_second", density, yy[:n], aa[:n], [0.8], var_mask=0b110)

# inv_gamma_lpdf(y | alpha, beta): known answers of test/prob/inv_gamma/inv_gamma_test.hpp:14-25, every subset of variables x
# propto, scalar / vector operands
for i, (y, a, b, ka) in enumerate([(1.0, 1.0, 1.0, -1.0), (0.5, 2.9, 3.1, -0.8185294827413338580868)]):
    add("inv_gamma_ka%d" % i, "inv_gamma", [y], [a], [b], var_mask=0b111, known_answer=ka)
# (the gamma corners 1e-300 / 1e8 overflow 1 / y^2 in the reference itself: milder ones here)
yi = np.concatenate([yg[:n], [1e-3, 1e3, 0.5, 3.0]])
ai = np.concatenate([al[:n], [0.5, 150.0, 1.4616321449683623, 30.0]])
bi = np.concatenate([be[:n], [1e-3, 1e3, 2.0, 0.05]])
for var_mask in range(8):
    for propto in (False, True):
        add("inv_gamma_vec_m%d%s" % (var_mask, "_propto" if propto else ""), "inv_gamma", yi, ai, bi, propto=propto, var_mask=var_mask)
add("inv_gamma_scalar_alpha_beta", "inv_gamma", yg[:n], [2.2], [1.3], var_mask=0b111)
add("inv_gamma_scalar_y", "inv_gamma", [2.5], al[:n], be[:n], var_mask=0b111)

out["cases"] = np.array(names)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "locscale_golden.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(names), "cases,", os.path.getsize(path), "bytes")
