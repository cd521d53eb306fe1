"""Generates tests/golden/student_t_golden.npz from the UNMODIFIED reference (oracle/_ref/libstanref.so, compiled from
/root/reference by oracle/Makefile): student_t_lpdf(y | nu, mu, sigma) value + reverse-mode gradient for
  * the reference's known-answer quadruples (test/prob/student_t/student_t_test.hpp:13-30, literal expected
    log-densities stored beside the reference's output),
  * every subset of {y, nu, mu, sigma} as autodiff variables x propto on one seeded vector case (the summands the
    reference drops under propto depend on which operands are variables, student_t_lpdf.hpp:155-166),
  * every mix of scalar / vector operands,
  * the corners of the domain: tiny and huge degrees of freedom, outcomes far in the tails, tiny and huge scales.
Run here (the container with /root/reference):  python tests/golden/make_student_t_golden.py
"""
import itertools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402

out, names = {}, []


def add(name, y, nu, mu, sigma, propto=False, var_mask=0b1110, known_answer=None):
    p = "st/%s/" % name
    r = L.ref_student_t_lpdf(y, nu, mu, sigma, propto=propto, var_mask=var_mask)
    for key, v in (("y", y), ("nu", nu), ("mu", mu), ("sigma", sigma)):
        out[p + key] = np.asarray(v, float).reshape(-1)
    out[p + "propto"] = np.array(int(propto))
    out[p + "var_mask"] = np.array(int(var_mask))
    out[p + "lp"] = np.array(r["lp"])
    for key in ("dy", "dnu", "dmu", "dsigma"):
        if key in r:
            out[p + key] = np.asarray(r[key], float)
    if known_answer is not None:
        out[p + "known_answer"] = np.array(known_answer)
    names.append(name)


# test/prob/student_t/student_t_test.hpp:13-30
add("ka0", [1.0], [1.0], [0.0], [1.0], var_mask=15, known_answer=-1.837877066409345339082)
add("ka1", [-3.0], [2.0], [0.0], [1.0], var_mask=15, known_answer=-3.596842909197555560041)

rng = np.random.default_rng(20240611)
n = 257
y = 3.0 * rng.standard_normal(n)
nu = rng.uniform(0.3, 40.0, n)
mu = rng.standard_normal(n)
sg = rng.uniform(0.2, 4.0, n)
for vm, propto in itertools.product(range(16), (False, True)):
    add("vec_m%d%s" % (vm, "_propto" if propto else ""), y, nu, mu, sg, propto=propto, var_mask=vm)

# scalar / vector mixes (bit q set: operand q is a vector), all four operands variables
for shape in range(16):
    ops = [v if shape >> q & 1 else v[:1] for q, v in enumerate((y, nu, mu, sg))]
    add("shape_%d" % shape, *ops, var_mask=15)
    add("shape_%d_propto_params" % shape, *ops, propto=True, var_mask=0b1110)

# corners of the domain
add("tiny_nu", y, [1e-3], mu, sg, var_mask=15)
add("huge_nu", y, [1e8], mu, sg, var_mask=15)
add("mixed_nu", y, 10.0 ** rng.uniform(-3, 8, n), mu, sg, var_mask=15)
add("far_tails", 1e6 * rng.standard_normal(n), nu, mu, sg, var_mask=15)
add("tiny_sigma", y, nu, mu, [1e-8], var_mask=15)
add("huge_sigma", y, nu, mu, [1e12], var_mask=15)
add("y_equals_mu", mu, nu, mu, sg, var_mask=15)
add("one_element_vectors", [0.5], [4.0], [-0.25], [2.0], var_mask=15)

big = 6007
add("n6007", 2.0 * rng.standard_normal(big), [4.5], 0.3 * rng.standard_normal(big), [1.7], var_mask=0b1110)
add("n6007_vec_nu", 2.0 * rng.standard_normal(big), rng.uniform(1, 30, big), [0.1], rng.uniform(0.5, 2, big), var_mask=15)

out["cases"] = np.array(names)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "student_t_golden.npz"), **out)
print("wrote %d cases" % len(names))
