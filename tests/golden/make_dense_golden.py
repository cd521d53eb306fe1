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

# ---- cholesky_decompose: factor and reverse sweep (rev/mat/fun/cholesky_decompose.hpp: cholesky_scalar up to n = 35,
# cholesky_block above) for f = sum(tril(W) * L)
chol_names = []


def add_chol(name, a, w=None):
    a = np.asarray(a, float)
    n = a.shape[0]
    rng = np.random.default_rng(1000 + len(chol_names))
    w = np.tril(rng.uniform(-1, 1, (n, n))) if w is None else np.tril(np.asarray(w, float))
    p = "chol/%s/" % name
    l, _ = L.ref_cholesky(a)
    r = L.ref_cholesky_rev(a, w)
    assert np.array_equal(np.tril(r["l"]), np.tril(l))
    out[p + "a"], out[p + "w"], out[p + "l"], out[p + "a_adj"] = a, w, l, r["a_adj"]
    chol_names.append(name)


# test/unit/math/opencl/cholesky_decompose_test.cpp:13-17 (the 3 x 3 factor is [[5,0,0],[3,3,0],[-1,1,3]])
add_chol("fixture_3x3", [[25, 15, -5], [15, 18, 0], [-5, 0, 11]])
add_chol("fixture_4x4", [[18, 22, 54, 42], [22, 70, 86, 62], [54, 86, 174, 134], [42, 62, 134, 106]])
for n in (1, 2, 10, 35, 36, 64, 65, 100):
    x = rng.uniform(-1, 1, (n, n))
    a = x @ x.T + n * np.eye(n)
    add_chol("spd_%d" % n, 0.5 * (a + a.T))
# a GP covariance with small jitter: condition number ~1e6
t = np.linspace(0, 3, 60)
add_chol("gp_60", 0.75 ** 2 * np.exp(-0.5 * (t[:, None] - t[None, :]) ** 2 / 0.5 ** 2) + 1e-4 * np.eye(60))
out["chol_cases"] = np.array(chol_names)

# ---- mdivide_left_tri: solve and reverse sweep (rev/mat/fun/mdivide_left_tri.hpp) for f = sum(W * X)
tri_names = []


def add_tri(name, a, b, lower):
    a, b = np.asarray(a, float), np.asarray(b, float)
    rng = np.random.default_rng(2000 + len(tri_names))
    w = rng.uniform(-1, 1, b.shape)
    p = "tri/%s/" % name
    x, _ = L.ref_mdivide_left_tri(a, b, lower)
    r = L.ref_mdivide_left_tri_rev(a, b, w, lower, 3)
    assert np.array_equal(r["x"], x)
    for mask in (1, 2):   # the adjoint of an operand does not depend on whether the other one is a variable (the
        q = L.ref_mdivide_left_tri_rev(a, b, w, lower, mask)   # three nodes of the reference differ in rounding only)
        got, want = (q["da"], r["da"]) if mask == 1 else (q["db"], r["db"])
        assert np.max(np.abs(got - want)) <= 1e-13 * np.max(np.abs(want))
    out[p + "a"], out[p + "b"], out[p + "w"], out[p + "lower"] = a, b, w, np.array(int(lower))
    out[p + "x"], out[p + "da"], out[p + "db"] = x, r["da"], r["db"]
    tri_names.append(name)


for (n, m) in ((1, 1), (3, 2), (63, 9), (65, 17), (130, 40)):
    a = rng.uniform(-5, 5, (n, n))   # the reference's fixture (rev/mat/fun/mdivide_left_tri_test.cpp:131-140): diagonal 20
    a[np.diag_indices(n)] = 20.0
    b = rng.uniform(-5, 5, (n, m))
    add_tri("lower_%dx%d" % (n, m), a, b, True)
    add_tri("upper_%dx%d" % (n, m), a, b, False)
out["tri_cases"] = np.array(tri_names)

out["cases"] = np.array(names)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "dense_golden.npz"), **out)
print("wrote %d multiply, %d cholesky, %d triangular-solve cases" % (len(names), len(chol_names), len(tri_names)))
