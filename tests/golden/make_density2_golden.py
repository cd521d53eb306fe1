"""Generates tests/golden/density2_golden.npz from the UNMODIFIED reference (oracle/_ref/libstanref.so, compiled from
/root/reference by oracle/Makefile): poisson_lpmf(n | lambda) and bernoulli_lpmf(n | theta), value + reverse-mode
gradient, for
  * the reference's known-answer pairs (test/prob/poisson/poisson_test.hpp:11-30,
    test/prob/bernoulli/bernoulli_test.hpp:11-35; the literal expected log-densities stored beside the output),
  * seeded vector cases with propto on and off,
  * a scalar operand against a vector of outcomes (bernoulli_lpmf's closed form from the number of successes,
    bernoulli_lpmf.hpp:58-91, with its all-ones and all-zeros branches) and a scalar outcome against a vector operand,
  * the corners: rates from 1e-300 to 1e6, probabilities at 1e-300 and 1 - 1e-16.  (A zero rate next to a zero count --
    poisson_lpmf.hpp:59 skips the term and the partial is 0/0 - 1 = NaN -- is compared NaN-aware against the compiled
    reference in tests/test_oracle.py instead of being stored.)
Same layout as the density cases of glm_golden.npz (tests/helpers.dens_case).
Run here (the container with /root/reference):  python tests/golden/make_density2_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402

out, names = {}, []


def add(name, density, y, theta, propto=False, known_answer=None):
    r = L.ref_density(density, y, theta, propto=propto)
    p = "dens/%s/" % name
    out[p + "density"] = np.array(L._abi.DENSITIES[density])
    out[p + "y"] = np.atleast_1d(np.asarray(y, np.int32))
    out[p + "theta"] = np.atleast_1d(np.asarray(theta, float))
    out[p + "propto"] = np.array(int(propto))
    out[p + "lp"], out[p + "dtheta"] = np.array(r["lp"]), r["dtheta"]
    if known_answer is not None:
        out[p + "known_answer"] = np.array(known_answer)
    names.append(name)


for i, (n, lam, ka) in enumerate([(17, 13.0, -2.900934373290765311282), (192, 42.0, -145.3546649655311853166), (0, 3.0, -3.0)]):
    add("pois_ka%d" % i, "poisson", [n], [lam], known_answer=ka)
for i, (n, th, ka) in enumerate([(1, 0.25, np.log(0.25)), (0, 0.25, np.log(0.75)), (1, 0.01, np.log(0.01)), (0, 0.01, np.log(0.99))]):
    add("bern_ka%d" % i, "bernoulli", [n], [th], known_answer=ka)

rng = np.random.default_rng(20191109)
lam = np.concatenate([rng.gamma(2.0, 2.0, 2000), [1e-300, 1e-12, 1e-3, 1.0, 700.0, 1e6, 0.5, 3.0]])
cnt = np.concatenate([rng.poisson(3.0, 2000), [0, 0, 1, 2, 650, 1000123, 0, 40]]).astype(np.int32)
th = np.concatenate([rng.uniform(0.01, 0.99, 2000), [1e-300, 1e-12, 0.5, 1 - 1e-16, 1 - 1e-9, 1e-9, 0.999, 0.001]])
yb = np.concatenate([(rng.random(2000) < 0.5), [1, 0, 1, 0, 1, 0, 0, 1]]).astype(np.int32)
for propto in (False, True):
    s = "_propto" if propto else ""
    add("pois_vec" + s, "poisson", cnt, lam, propto=propto)
    add("bern_vec" + s, "bernoulli", yb, th, propto=propto)
add("pois_scalar_lambda", "poisson", cnt[:2000], [2.7])
add("pois_scalar_n", "poisson", [4], lam)
add("bern_scalar_theta", "bernoulli", yb, [0.3])
add("bern_scalar_theta_all_ones", "bernoulli", np.ones(500, np.int32), [0.3])
add("bern_scalar_theta_all_zeros", "bernoulli", np.zeros(500, np.int32), [0.3])
add("bern_scalar_n", "bernoulli", [1], th)
add("bern_scalar_n0", "bernoulli", [0], th)

out["dens_cases"] = np.array(names)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "density2_golden.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(names), "cases,", os.path.getsize(path), "bytes")
