"""Generates tests/golden/dense_golden.npz from the UNMODIFIED reference (oracle/_ref/libstanref.so, compiled from
/root/reference by oracle/Makefile): stan::math::multiply(A, B) for double matrices (prim/mat/fun/multiply.hpp:66-83)
and, through multiply_mat_vari (rev/mat/fun/multiply.hpp), the product with var operands plus the adjoints
dA = W B', dB = A' W of f = sum(W * AB), for every mix of var / double operands.  The reference holds no literal
golden values for multiply (its tests compare OpenCL against CPU and autodiff against finite differences); the
fixtures of test/unit/math/prim/mat/fun/multiply_test.cpp:50-90 are included with the reference's output.
Run here (the container with /root/reference):  python tests/golden/make_dense_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402

out, names = {}, []


def add(name, a, b, w=None):
    a, b = np.asarray(a, float), np.asarray(b, float)
    rng = np.random.default_rng(len(names) + 11)
    w = rng.uniform(-1, 1, (a.shape[0], b.shape[1])) if w is None else np.asarray(w, float)
    p = "mm/%s/" % name
    c, _ = L.ref_multiply(a, b)
    out[p + "a"], out[p + "b"], out[p + "w"], out[p + "c"] = a, b, w, c
    for mask in (1, 2, 3):  # the adjoint of an operand does not depend on whether the other one is a variable
        r = L.ref_multiply_rev(a, b, w, mask)
        assert np.array_equal(r["c"], c)
        for key, bit in (("da", 1), ("db", 2)):
            if mask & bit:
                assert p + key not in out or np.array_equal(out[p + key], r[key])
                out[p + key] = r[key]
    names.append(name)


# test/unit/math/prim/mat/fun/multiply_test.cpp: the 2x3 * 3x2 and 3x2 * 2x3 fixtures
add("fixture_2x3_3x2", [[1, 2, 3], [4, 5, 6]], [[1, 2], [3, 4], [5, 6]])
add("fixture_3x2_2x3", [[1, 2], [3, 4], [5, 6]], [[1, 2, 3], [4, 5, 6]])
rng = np.random.default_rng(99)
for (m, k, n) in ((1, 1, 1), (17, 19, 23), (64, 64, 64), (65, 63, 66), (129, 9, 127), (8, 300, 8)):
    add("rand_%dx%dx%d" % (m, k, n), rng.uniform(-5, 5, (m, k)), rng.uniform(-5, 5, (k, n)))
# cancellation: rows of A nearly orthogonal to columns of B
a = rng.standard_normal((40, 60))
b = np.linalg.qr(rng.standard_normal((60, 60)))[0][:, :30]
add("cancelling", a - a @ b @ b.T, b)
# wide dynamic range
add("wide_range", rng.standard_normal((30, 50)) * 10.0 ** rng.uniform(-8, 8, (30, 50)), rng.standard_normal((50, 20)))

out["cases"] = np.array(names)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "dense_golden.npz"), **out)
print("wrote %d cases" % len(names))
