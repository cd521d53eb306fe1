"""Dense FP64 product on the device (the `multiply` hooks, stan/math/prim/mat/fun/multiply.hpp:66-83,
stan/math/rev/mat/fun/multiply.hpp:35-150, and the matrix_cl surface they use).

CPU: the oracle's triple loop against the golden vectors of the compiled reference (value and the two adjoint
products) and against the compiled reference on fresh inputs.  GPU: stan_cuda_multiply and the matrix handle API
through the C ABI against the goldens and the oracle on shapes around every tile edge, and at 4096^3 through
properties that need no checker.

The gate for a product is componentwise: |got - ref| <= 1e-10 * (|A| |B|)_ij, the scale to which ANY summation order
of the same products is accurate (an entry of A B can be arbitrarily small against its own terms; the reference's
tests for its OpenCL product use an absolute 1e-10 / 1e-12 on O(1000) values for the same reason,
test/unit/math/rev/mat/fun/multiply_test.cpp:39-40).
"""
import ctypes

import numpy as np
import pytest

from math_b200 import _abi
from oracle import loader as L


def product_error(got, ref, a, b):
    bound = np.abs(a) @ np.abs(b)
    bound = np.where(bound > 0, bound, 1.0)
    return float(np.max(np.abs(np.asarray(got) - np.asarray(ref)) / bound)) if ref.size else 0.0


@pytest.fixture(scope="session")
def golden_mm():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    return np.load(os.path.join(root, "tests", "golden", "dense_golden.npz"), allow_pickle=False)


def _case(g, name):
    p = "mm/%s/" % name
    return {k: g[p + k] for k in ("a", "b", "w", "c", "da", "db")}


# ------------------------------------------------------------------------------------------ CPU
def test_struct_layout():
    assert ctypes.sizeof(_abi.MultiplyArgs) == 8 * 3 + 8 + 8 + 4 + 4 + 8 + 8 + 8 + 8


def test_oracle_matches_golden(golden_mm):
    assert len(golden_mm["cases"]) >= 10
    for name in golden_mm["cases"]:
        c = _case(golden_mm, str(name))
        a, b, w = c["a"], c["b"], c["w"]
        assert product_error(L.oracle_multiply(a, b), c["c"], a, b) <= 1e-13, name
        assert product_error(L.oracle_multiply(w, b, trans_b=True), c["da"], w, b.T) <= 1e-13, name   # dA = W B'
        assert product_error(L.oracle_multiply(a, w, trans_a=True), c["db"], a.T, w) <= 1e-13, name   # dB = A' W
    fx = _case(golden_mm, "fixture_2x3_3x2")
    assert np.array_equal(fx["c"], np.array([[22.0, 28.0], [49.0, 64.0]]))


@pytest.mark.skipif(not L.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_matches_compiled_reference_on_fresh_inputs():
    rng = np.random.default_rng(3)
    for _ in range(6):
        m, k, n = (int(v) for v in rng.integers(1, 150, 3))
        a, b, w = rng.standard_normal((m, k)), rng.standard_normal((k, n)), rng.standard_normal((m, n))
        ref = L.ref_multiply_rev(a, b, w, 3)
        assert product_error(L.oracle_multiply(a, b), ref["c"], a, b) <= 1e-13
        assert product_error(L.oracle_multiply(w, b, trans_b=True), ref["da"], w, b.T) <= 1e-13
        assert product_error(L.oracle_multiply(a, w, trans_a=True), ref["db"], a.T, w) <= 1e-13


def _chol_case(g, name):
    p = "chol/%s/" % name
    return {k: g[p + k] for k in ("a", "w", "l", "a_adj")}


def lower_rel_err(got, ref):
    """the reference's own metric for its OpenCL factor (test/unit/math/opencl/cholesky_decompose_test.cpp:52-60):
    worst relative error over the lower triangle, with the parity gate's floor of 1e-6 * max|ref|"""
    got, ref = np.tril(np.asarray(got)), np.tril(np.asarray(ref))
    if ref.size == 0:
        return 0.0
    scale = np.maximum(np.abs(ref), 1e-6 * np.max(np.abs(ref)))
    scale = np.where(scale == 0, 1.0, scale)
    return float(np.max(np.abs(got - ref) / scale))


def test_oracle_cholesky_matches_golden(golden_mm):
    assert len(golden_mm["chol_cases"]) >= 10
    for name in golden_mm["chol_cases"]:
        c = _chol_case(golden_mm, str(name))
        rc, l = L.oracle_cholesky(c["a"])
        # the GP case has condition number ~1e6: two backward-stable algorithms agree to cond * eps, not to eps
        tol = 1e-9 if str(name).startswith("gp") else 1e-12
        assert rc == 0 and lower_rel_err(l, c["l"]) <= tol, name
        assert not np.triu(l, 1).any()
        assert lower_rel_err(L.oracle_cholesky_adjoint(c["l"], c["w"]), c["a_adj"]) <= tol, name
    fx = _chol_case(golden_mm, "fixture_3x3")
    assert np.allclose(fx["l"], [[5, 0, 0], [3, 3, 0], [-1, 1, 3]], atol=1e-14)


def test_oracle_cholesky_rejects_what_the_reference_rejects():
    for bad in (-np.eye(3), np.array([[1.0, 2.0], [2.0, 1.0]]), np.array([[1.0, np.nan], [np.nan, 1.0]]), np.zeros((2, 2))):
        assert L.oracle_cholesky(bad)[0] == _abi.NOT_POS_DEF
        if L.have_ref():
            with pytest.raises(L.RefError):
                L.ref_cholesky(bad)


def _tri_case(g, name):
    p = "tri/%s/" % name
    c = {k: g[p + k] for k in ("a", "b", "w", "x", "da", "db")}
    c["lower"] = bool(g[p + "lower"])
    return c


def _check_tri(solve, c, tol, what):
    """x = A^-1 b, adjB = A^-T w, adjA = -(adjB x') on the operand triangle (rev/mat/fun/mdivide_left_tri.hpp:98-119)"""
    lower, a = c["lower"], c["a"]
    x = solve(a, c["b"], lower, False)
    assert np.max(np.abs(x - c["x"])) <= tol * np.max(np.abs(c["x"])), what
    adj_b = solve(a, c["w"], lower, True)
    assert np.max(np.abs(adj_b - c["db"])) <= tol * np.max(np.abs(c["db"])), what
    bound = np.abs(c["db"]) @ np.abs(c["x"]).T
    adj_a = -(adj_b @ x.T)
    adj_a = np.tril(adj_a) if lower else np.triu(adj_a)
    assert np.max(np.abs(adj_a - c["da"]) / np.where(bound > 0, bound, 1.0)) <= tol, what


def test_oracle_tri_solve_matches_golden(golden_mm):
    assert len(golden_mm["tri_cases"]) >= 10
    for name in golden_mm["tri_cases"]:
        _check_tri(lambda a, b, lower, trans: L.oracle_tri_solve(a, b, lower, trans), _tri_case(golden_mm, str(name)), 1e-12, name)


def test_oracle_gp_exp_quad_cov_known_values():
    """the fixture of the reference's test (test/unit/math/prim/mat/fun/gp_exp_quad_cov_test.cpp: x = (-2, -1, -0.5),
    sigma = 0.2, l = 5: cov(i, j) = sigma^2 exp(-(x_i - x_j)^2 / (2 l^2))) and the two-sample / per-dimension forms
    against the closed formula; the C++ binaries pin the same function against the compiled reference"""
    x = np.array([-2.0, -1.0, -0.5])
    cov = L.oracle_gp_exp_quad_cov(x, 0.2, 5.0)
    want = 0.04 * np.exp(-0.5 * (x[:, None] - x[None, :]) ** 2 / 25.0)
    assert np.max(np.abs(cov - want)) <= 1e-17 and np.all(np.diag(cov) == 0.2 * 0.2)
    rng = np.random.default_rng(1)
    p1, p2, ls = rng.standard_normal((7, 3)), rng.standard_normal((5, 3)), np.array([0.7, 1.3, 2.0])
    cross = L.oracle_gp_exp_quad_cov(p1, 1.5, ls, p2)
    d2 = (((p1[:, None, :] - p2[None, :, :]) / ls) ** 2).sum(-1)
    assert np.max(np.abs(cross - 2.25 * np.exp(-0.5 * d2))) <= 1e-15


# ------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def dense():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from math_b200 import dense as d, glm
    glm.init(1)
    return d


@pytest.mark.gpu
def test_device_matches_golden(dense, golden_mm):
    for name in golden_mm["cases"]:
        c = _case(golden_mm, str(name))
        a, b, w = c["a"], c["b"], c["w"]
        assert product_error(dense.multiply_host(a, b), c["c"], a, b) <= 1e-10, name
        assert product_error(dense.multiply_host(w, b, trans_b=True), c["da"], w, b.T) <= 1e-10, name
        assert product_error(dense.multiply_host(a, w, trans_a=True), c["db"], a.T, w) <= 1e-10, name


SHAPES = [(1, 1, 1), (1, 7, 1), (2, 3, 4), (63, 8, 65), (64, 7, 64), (64, 9, 64), (65, 16, 63), (127, 33, 129), (128, 128, 128),
          (129, 130, 131), (255, 17, 257), (300, 1, 300), (1, 300, 1), (5, 1000, 3), (1000, 5, 3), (3, 5, 1000), (513, 64, 700),
          (2048, 24, 1500), (64, 5000, 130), (1, 3000, 1), (100, 1024, 100)]   # the last three: split-K


@pytest.mark.gpu
@pytest.mark.parametrize("m,k,n", SHAPES)
def test_device_matches_oracle_on_every_tile_edge(dense, m, k, n):
    """shapes below, at and above the 64- and 128-wide tiles and the 8-deep K step; the last one is large enough for
    the 128 x 128 kernel (>= 148 tiles), the others run the 64 x 64 kernel"""
    rng = np.random.default_rng(m * 1000003 + k * 1009 + n)
    a, b = rng.uniform(-5, 5, (m, k)), rng.uniform(-5, 5, (k, n))
    ref = L.oracle_multiply(a, b)
    assert product_error(dense.multiply_host(a, b), ref, a, b) <= 1e-10
    at, bt = np.asfortranarray(a.T), np.asfortranarray(b.T)
    assert product_error(dense.multiply_host(at, b, trans_a=True), ref, a, b) <= 1e-10
    assert product_error(dense.multiply_host(a, bt, trans_b=True), ref, a, b) <= 1e-10
    assert product_error(dense.multiply_host(at, bt, trans_a=True, trans_b=True), ref, a, b) <= 1e-10


@pytest.mark.gpu
def test_handle_api(dense):
    rng = np.random.default_rng(8)
    a, b = rng.standard_normal((150, 70)), rng.standard_normal((70, 90))
    da, db = dense.MatrixCuda(a), dense.MatrixCuda(b)
    assert da.shape == (150, 70) and da.T.shape == (70, 150)
    assert np.array_equal(da.to_host(), a) and np.array_equal(da.T.to_host(), a.T)
    ref = L.oracle_multiply(a, b)
    c = da @ db
    assert c.shape == (150, 90) and product_error(c.to_host(), ref, a, b) <= 1e-10
    # the products of the reverse sweep from resident operands: adjA = W B', adjB = A' W
    w = rng.standard_normal((150, 90))
    dw = dense.MatrixCuda(w)
    assert product_error((dw @ db.T).to_host(), L.oracle_multiply(w, b, trans_b=True), w, b.T) <= 1e-10
    assert product_error((da.T @ dw).to_host(), L.oracle_multiply(a, w, trans_a=True), a.T, w) <= 1e-10
    with pytest.raises(ValueError, match="must match in size"):
        da @ da
    with pytest.raises(Exception):   # the ABI refuses a result of the wrong shape / an aliased result
        from math_b200._lib import check, lib
        check(lib().stan_cuda_matrix_multiply(da._h, 0, db._h, 0, da._h))
    empty = dense.MatrixCuda(np.zeros((0, 4)))
    assert empty.to_host().shape == (0, 4)
    z = dense.MatrixCuda(np.zeros((5, 0))) @ dense.MatrixCuda(np.zeros((0, 6)))   # empty inner dimension: zeros
    assert z.shape == (5, 6) and not z.to_host().any()


@pytest.mark.gpu
def test_results_repeat_bitwise_and_do_not_depend_on_the_tiling(dense):
    """every entry is one FMA chain over k ascending, whichever kernel and tile owns it: a column block of the product
    equals the product with that column block of B bit for bit (here the full product runs the 128-wide kernel and
    the block the 64-wide one), and two calls agree bit for bit"""
    rng = np.random.default_rng(21)
    a, b = rng.standard_normal((2000, 300)), rng.standard_normal((300, 1800))
    full = dense.multiply_host(a, b)
    assert np.array_equal(full, dense.multiply_host(a, b))
    assert np.array_equal(full[:, 100:260], dense.multiply_host(a, b[:, 100:260]))
    assert np.array_equal(full[37:100, :], dense.multiply_host(a[37:100, :], b))


@pytest.mark.gpu
def test_full_size_properties(dense):
    """4096^3 (137 GFLOP: minutes for the oracle's triple loop): (A B) x = A (B x) for random x, A I = A exactly,
    A (B + B2) = A B + A B2 to rounding, and a sampled block against the oracle."""
    rng = np.random.default_rng(5)
    n = 4096
    a, b = rng.standard_normal((n, n)), rng.standard_normal((n, n))
    da, db = dense.MatrixCuda(a), dense.MatrixCuda(b)
    c = (da @ db).to_host()
    x = rng.standard_normal(n)
    lhs, rhs = c @ x, a @ (b @ x)
    scale = np.abs(a) @ (np.abs(b) @ np.abs(x))
    assert np.max(np.abs(lhs - rhs) / scale) <= 1e-10
    eye = dense.MatrixCuda(np.eye(n))
    assert np.array_equal((da @ eye).to_host(), a)
    rows, cols = rng.integers(0, n, 24), rng.integers(0, n, 24)
    block_ref = L.oracle_multiply(a[rows, :], b[:, cols])
    assert product_error(c[np.ix_(rows, cols)], block_ref, a[rows, :], b[:, cols]) <= 1e-10
    b2 = rng.standard_normal((n, n))
    c2 = (da @ dense.MatrixCuda(b2)).to_host()
    csum = (da @ dense.MatrixCuda(b + b2)).to_host()
    assert product_error(csum, c + c2, a, np.abs(b) + np.abs(b2)) <= 1e-10


# ---- cholesky_decompose ------------------------------------------------------------------------------------
def _spd(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (n, n))
    a = x @ x.T + n * np.eye(n)
    return 0.5 * (a + a.T)


@pytest.mark.gpu
def test_device_cholesky_matches_golden(dense, golden_mm):
    for name in golden_mm["chol_cases"]:
        c = _chol_case(golden_mm, str(name))
        l = dense.cholesky_decompose(c["a"])
        tol = 1e-8 if str(name).startswith("gp") else 1e-10   # cond(A) ~ 1e6 for the GP case
        assert lower_rel_err(l, c["l"]) <= tol, name
        assert not np.triu(l, 1).any()
        assert lower_rel_err(dense.cholesky_adjoint(c["l"], c["w"]), c["a_adj"]) <= tol, name


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 3, 31, 63, 64, 65, 127, 128, 129, 200, 321, 640, 1000])
def test_device_cholesky_matches_oracle(dense, n):
    a = _spd(n, n)
    rc, ref = L.oracle_cholesky(a)
    assert rc == 0
    l = dense.cholesky_decompose(a)
    assert lower_rel_err(l, ref) <= 1e-10
    w = np.tril(np.random.default_rng(n + 1).uniform(-1, 1, (n, n)))
    if n <= 400:   # the oracle's unblocked sweep is O(n^3) scalar loops
        assert lower_rel_err(dense.cholesky_adjoint(ref, w), L.oracle_cholesky_adjoint(ref, w)) <= 1e-10


@pytest.mark.gpu
def test_device_cholesky_rejects_what_the_reference_rejects(dense):
    for bad in (-np.eye(3), np.array([[1.0, 2.0], [2.0, 1.0]]), np.array([[np.nan, 0.5], [0.5, 1.0]]), np.zeros((2, 2))):
        with pytest.raises(dense.NotPositiveDefinite):
            dense.cholesky_decompose(bad)
    big = _spd(300, 7)
    big[250, 250] = -1.0    # fails in the fourth block: the flag survives the remaining steps
    with pytest.raises(dense.NotPositiveDefinite):
        dense.cholesky_decompose(big)
    asym = _spd(200, 8)
    asym[150, 3] += 1e-6      # beyond CONSTRAINT_TOLERANCE = 1e-8: the device check reports it before any pivot
    with pytest.raises(dense.NotSymmetric):
        dense.cholesky_decompose(asym)
    asym[150, 3] = asym[3, 150] + 1e-10   # inside the tolerance: accepted, the lower triangle is what is factorised
    assert lower_rel_err(dense.cholesky_decompose(asym), L.oracle_cholesky(asym)[1]) <= 1e-10
    nan_off = _spd(100, 9)
    nan_off[40, 7] = np.nan   # NaN fails the symmetry test as in the reference (check_symmetric.hpp:43-44)
    with pytest.raises(dense.NotSymmetric):
        dense.cholesky_decompose(nan_off)
    assert np.array_equal(dense.cholesky_decompose(asym, symmetry_tolerance=None), dense.cholesky_decompose(asym))
    with pytest.raises(ValueError, match="square"):
        dense.cholesky_decompose(np.zeros((2, 3)))
    assert dense.cholesky_decompose(np.zeros((0, 0))).shape == (0, 0)


@pytest.mark.gpu
def test_cholesky_full_size_properties(dense):
    """n = 4096: L L' reproduces A to the backward-error bound, L is lower triangular with a positive diagonal, the
    factor of a block-diagonal matrix is block diagonal, and the reverse sweep satisfies the identity
    <L_adj, dL> = <A_adj, dA> for a random symmetric perturbation (directional derivative by finite differences)."""
    n = 4096
    a = _spd(n, 5)
    l = dense.cholesky_decompose(a)
    assert not np.triu(l, 1).any() and np.all(np.diag(l) > 0)
    resid = dense.multiply_host(l, l, trans_b=True) - a
    assert np.max(np.abs(resid)) <= 1e-12 * np.max(np.abs(a)) * n ** 0.5
    rng = np.random.default_rng(6)
    w = np.tril(rng.uniform(-1, 1, (n, n)))
    a_adj = dense.cholesky_adjoint(l, w)
    e = rng.uniform(-1, 1, (n, n))
    e = 0.5 * (e + e.T)
    h = 1e-6
    lp, lm = dense.cholesky_decompose(a + h * e), dense.cholesky_decompose(a - h * e)
    fd = np.sum(w * (lp - lm)) / (2 * h)
    # the operands are the lower triangle of A only: a symmetric perturbation moves (i, j) and (j, i) together
    analytic = np.sum(a_adj * np.tril(e))
    assert abs(fd - analytic) <= 1e-6 * max(abs(analytic), 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [50, 300, 1100])
def test_cholesky_adjoint_reads_lower_triangles_only(dense, n):
    """the strict upper triangles of L and of its adjoint are never read (the C++ node hands over work matrices whose
    upper triangles hold whatever the previous evaluation left there): poisoned with NaN, same result bit for bit.
    n = 50 travels whole, the others as triangles (staged_copy's tri mode)."""
    a = _spd(n, 21)
    l = dense.cholesky_decompose(a)
    w = np.tril(np.random.default_rng(n).uniform(-1, 1, (n, n)))
    want = dense.cholesky_adjoint(l, w)
    upper = np.triu(np.ones((n, n), bool), 1)
    lp, wp = l.copy(), w.copy()
    lp[upper] = np.nan
    wp[upper] = np.nan
    got = dense.cholesky_adjoint(np.asfortranarray(lp), np.asfortranarray(wp))
    assert np.array_equal(got, want)
    assert not np.triu(got, 1).any()


# ---- mdivide_left_tri --------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_device_tri_solve_matches_golden(dense, golden_mm):
    for name in golden_mm["tri_cases"]:
        _check_tri(lambda a, b, lower, trans: dense.tri_solve(a, b, lower, trans), _tri_case(golden_mm, str(name)), 1e-10, name)


@pytest.mark.gpu
@pytest.mark.parametrize("n,m", [(1, 1), (2, 5), (63, 64), (64, 1), (65, 130), (128, 128), (129, 7), (300, 300), (1000, 33)])
def test_device_tri_solve_matches_oracle(dense, n, m):
    rng = np.random.default_rng(n * 1009 + m)
    a = rng.uniform(-5, 5, (n, n))
    a[np.diag_indices(n)] = max(20.0, 0.6 * n)
    b = rng.uniform(-5, 5, (n, m))
    for lower in (True, False):
        for trans in (False, True):
            ref = L.oracle_tri_solve(a, b, lower, trans)
            got = dense.tri_solve(a, b, lower, trans)
            assert np.max(np.abs(got - ref)) <= 1e-10 * np.max(np.abs(ref)), (lower, trans)
    v = b[:, 0].copy()
    assert np.max(np.abs(dense.tri_solve(a, v) - L.oracle_tri_solve(a, v)[:, 0])) <= 1e-10 * np.max(np.abs(v))


@pytest.mark.gpu
def test_tri_solve_full_size_properties(dense):
    """n = 4096, 512 right-hand sides: the residual op(A) x - b is at rounding level, the other triangle is never read
    (poisoned with NaN), and solving against the identity gives the inverse (A A^-1 = I)."""
    rng = np.random.default_rng(4)
    n, m = 4096, 512
    a = rng.uniform(-1, 1, (n, n))
    a[np.diag_indices(n)] = 0.05 * n
    b = rng.uniform(-1, 1, (n, m))
    for lower in (True, False):
        tri = np.tril(a) if lower else np.triu(a)
        poisoned = np.where(tri != 0, tri, np.nan)
        for trans in (False, True):
            x = dense.tri_solve(poisoned, b, lower, trans)
            op = tri.T if trans else tri
            assert np.max(np.abs(op @ x - b)) <= 1e-12 * n ** 0.5 * np.max(np.abs(b)) * 10
    inv = dense.tri_solve(a, np.eye(n), True, False)
    assert np.max(np.abs(np.tril(a) @ inv - np.eye(n))) <= 1e-11
    with pytest.raises(ValueError, match="must match in size"):
        dense.tri_solve(a, np.zeros((5, 2)))


# ---- gp_exp_quad_cov ---------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n1,n2,dim", [(1, 1, 1), (3, 3, 1), (31, 33, 2), (300, 300, 1), (257, 129, 3), (1000, 40, 5)])
def test_device_gp_exp_quad_cov_matches_oracle(dense, n1, n2, dim):
    rng = np.random.default_rng(n1 * 31 + n2 * 7 + dim)
    x1, x2 = rng.uniform(-2, 2, (n1, dim)), rng.uniform(-2, 2, (n2, dim))
    if dim == 1:
        x1, x2 = x1[:, 0], x2[:, 0]
    ls = rng.uniform(0.5, 2.0, dim)
    for length_scale in (0.8, ls) if dim > 1 else (0.8,):
        ref = L.oracle_gp_exp_quad_cov(x1, 1.3, length_scale)
        got = dense.gp_exp_quad_cov(x1, 1.3, length_scale)
        assert np.max(np.abs(got - ref)) <= 1e-10 * np.max(np.abs(ref))
        assert np.array_equal(got, got.T) and np.all(np.diag(got) == 1.3 * 1.3)
        ref = L.oracle_gp_exp_quad_cov(x1, 1.3, length_scale, x2)
        got = dense.gp_exp_quad_cov(x1, 1.3, length_scale, x2)
        assert np.max(np.abs(got - ref)) <= 1e-10 * np.max(np.abs(ref))


@pytest.mark.gpu
def test_gp_exp_quad_cov_feeds_cholesky_at_full_size(dense):
    """n = 4096 points on a line: the covariance is symmetric with sigma^2 on the diagonal, entries decay with distance,
    and with jitter it is positive definite for the device Cholesky (the GP pipeline the hooks exist for)."""
    n = 4096
    x = np.linspace(0.0, 40.0, n)
    cov = dense.gp_exp_quad_cov(x, 0.75, 0.5)
    assert np.array_equal(cov, cov.T) and np.all(np.diag(cov) == 0.75 * 0.75)
    assert abs(cov[0, 1] - 0.5625 * np.exp(-0.5 * (x[1] - x[0]) ** 2 / 0.25)) <= 1e-15
    l = dense.cholesky_decompose(cov + 1e-6 * np.eye(n))
    assert np.max(np.abs(dense.multiply_host(l, l, trans_b=True) - cov - 1e-6 * np.eye(n))) <= 1e-12


@pytest.mark.gpu
def test_concurrent_host_threads(dense):
    """four host threads call the dense entry points at once (per-thread streams and staging buffers, one shared pool
    of copy helpers): every result equals the single-threaded one bit for bit"""
    import threading
    rng = np.random.default_rng(11)
    a = np.asfortranarray(rng.standard_normal((1500, 1100)))
    b = np.asfortranarray(rng.standard_normal((1100, 1300)))
    spd = _spd(900, 12)
    want_c, want_l = dense.multiply_host(a, b), dense.cholesky_decompose(spd)
    errors = []

    def work(k):
        try:
            for _ in range(3):
                if k % 2 == 0:
                    assert np.array_equal(dense.multiply_host(a, b), want_c)
                else:
                    assert np.array_equal(dense.cholesky_decompose(spd), want_l)
        except Exception as e:   # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
