"""Dense FP64 matrices on the device and their product: the ctypes mirror of what the reference's
OpenCL backend gives its `multiply` hooks -- matrix_cl (stan/math/opencl/matrix_cl.hpp:32-454), to_matrix_cl /
from_matrix_cl (opencl/copy.hpp:37-93), transpose and operator* (opencl/multiply.hpp) -- used by the tests and the
benchmark.  Column-major like Eigen; numpy arrays are converted with order="F"."""
import ctypes as C

import numpy as np

from . import _abi
from ._lib import check, lib


def _fmat(a):
    a = np.asarray(a, dtype=np.float64)
    if a.ndim != 2:
        raise ValueError("expected a matrix")
    return np.asfortranarray(a)


class MatrixCuda:
    """matrix_cl<double>: a rows x cols matrix in device memory.  `MatrixCuda(A)` uploads (to_matrix_cl),
    `.to_host()` downloads (from_matrix_cl), `a @ b` / `multiply(a, b)` is the product, `.T` a transposed view."""

    def __init__(self, a=None, rows=None, cols=None):
        self._h = C.c_void_p()
        self._trans = False
        self._owner = None
        if a is not None:
            a = _fmat(a)
            rows, cols = a.shape
        check(lib().stan_cuda_matrix_create(int(rows), int(cols), C.byref(self._h)))
        if a is not None and a.size:
            check(lib().stan_cuda_matrix_upload(self._h, a.ctypes.data, max(a.shape[0], 1)))

    @property
    def shape(self):
        r, c = lib().stan_cuda_matrix_rows(self._h), lib().stan_cuda_matrix_cols(self._h)
        return (c, r) if self._trans else (r, c)

    @property
    def T(self):
        v = MatrixCuda.__new__(MatrixCuda)
        v._h, v._trans, v._owner = self._h, not self._trans, self._owner or self
        return v

    def to_host(self, out=None):
        """from_matrix_cl; `out` (optional): an r x c Fortran-ordered array to fill instead of a fresh one"""
        r, c = lib().stan_cuda_matrix_rows(self._h), lib().stan_cuda_matrix_cols(self._h)
        if out is None:
            out = np.zeros((r, c), order="F")
        assert out.shape == (r, c) and out.flags.f_contiguous and out.dtype == np.float64
        if out.size:
            check(lib().stan_cuda_matrix_download(self._h, out.ctypes.data, max(r, 1)))
        return out.T if self._trans else out

    def __matmul__(self, other):
        return multiply(self, other)

    def __del__(self):
        if getattr(self, "_owner", None) is None and getattr(self, "_h", None):
            try:
                lib().stan_cuda_matrix_free(self._h)
            except Exception:
                pass
            self._h = None


def multiply(a, b):
    """Product of two MatrixCuda (views): a new MatrixCuda.  Raises ValueError where the reference's
    check_multiplicable throws std::invalid_argument (prim/mat/fun/multiply.hpp:69)."""
    (m, ka), (kb, n) = a.shape, b.shape
    if ka != kb:
        raise ValueError("multiply: Columns of m1 (%d) and Rows of m2 (%d) must match in size" % (ka, kb))
    c = MatrixCuda(rows=m, cols=n)
    check(lib().stan_cuda_matrix_multiply(a._h, int(a._trans), b._h, int(b._trans), c._h))
    return c


def multiply_host(a, b, trans_a=False, trans_b=False):
    """op(a) @ op(b) for host matrices through stan_cuda_multiply (upload, product, download in one call)."""
    a, b = _fmat(a), _fmat(b)
    m, ka = (a.shape[1], a.shape[0]) if trans_a else a.shape
    kb, n = (b.shape[1], b.shape[0]) if trans_b else b.shape
    if ka != kb:
        raise ValueError("multiply: Columns of m1 (%d) and Rows of m2 (%d) must match in size" % (ka, kb))
    c = np.zeros((m, n), order="F")
    g = _abi.MultiplyArgs()
    g.m, g.n, g.k = m, n, ka
    g.a, g.lda, g.trans_a = a.ctypes.data, max(a.shape[0], 1), int(trans_a)
    g.b, g.ldb, g.trans_b = b.ctypes.data, max(b.shape[0], 1), int(trans_b)
    g.c, g.ldc = c.ctypes.data, max(m, 1)
    check(lib().stan_cuda_multiply(C.byref(g)))
    return c


class NotPositiveDefinite(ValueError):
    """cholesky_decompose: the reference throws std::domain_error ("Matrix m is not positive definite")."""


class NotSymmetric(ValueError):
    """cholesky_decompose: the reference throws std::domain_error ("m is not symmetric")."""


def cholesky_decompose(a, symmetry_tolerance=1e-8):
    """Lower Cholesky factor of a symmetric positive-definite matrix (stan_cuda_cholesky_decompose); the symmetry check
    of the reference (CONSTRAINT_TOLERANCE = 1e-8) runs on the device; None skips it."""
    a = _fmat(a)
    if a.shape[0] != a.shape[1]:
        raise ValueError("cholesky_decompose: Expecting a square matrix; rows of A (%d) and columns of A (%d) must match in size" % a.shape)
    n = a.shape[0]
    l = np.zeros((n, n), order="F")
    rc = check(lib().stan_cuda_cholesky_decompose(n, a.ctypes.data, max(n, 1), l.ctypes.data, max(n, 1),
                                                  -1.0 if symmetry_tolerance is None else float(symmetry_tolerance)))
    if rc == _abi.NOT_SYMMETRIC:
        raise NotSymmetric("cholesky_decompose: m is not symmetric")
    if rc == _abi.NOT_POS_DEF:
        raise NotPositiveDefinite("cholesky_decompose: Matrix m is not positive definite")
    return l


def cholesky_adjoint(l, l_adj):
    """Reverse sweep of L = cholesky_decompose(A): the adjoint of the lower triangle of A (zeros above)."""
    l, l_adj = _fmat(l), _fmat(l_adj)
    n = l.shape[0]
    out = np.zeros((n, n), order="F")
    check(lib().stan_cuda_cholesky_adjoint(n, l.ctypes.data, max(n, 1), l_adj.ctypes.data, max(n, 1), out.ctypes.data, max(n, 1)))
    return out


def tri_solve(a, b, lower=True, trans=False):
    """op(A)^-1 b for a triangular A (stan_cuda_tri_solve): mdivide_left_tri<Lower | Upper>(A, b) and, with trans, the
    A^-T adj product of its reverse sweep."""
    a = _fmat(a)
    b = np.asarray(b, dtype=np.float64)
    vec = b.ndim == 1
    b = _fmat(b.reshape(-1, 1) if vec else b)
    if a.shape[0] != a.shape[1]:
        raise ValueError("mdivide_left_tri: Expecting a square matrix; rows of A (%d) and columns of A (%d) must match in size" % a.shape)
    if a.shape[1] != b.shape[0]:
        raise ValueError("mdivide_left_tri: Columns of A (%d) and Rows of b (%d) must match in size" % (a.shape[1], b.shape[0]))
    n, m = a.shape[0], b.shape[1]
    x = np.zeros((n, m), order="F")
    check(lib().stan_cuda_tri_solve(int(lower), int(trans), n, m, a.ctypes.data, max(n, 1), b.ctypes.data, max(n, 1),
                                    x.ctypes.data, max(n, 1)))
    return x[:, 0] if vec else x


def gp_exp_quad_cov(x1, sigma, length_scale, x2=None):
    """Squared-exponential covariance sigma^2 exp(-|x1_i - x2_j|^2 / (2 l^2)) (stan_cuda_gp_exp_quad_cov).  x: n scalars or
    an n x dim array of points; length_scale a scalar or one value per dimension (the points are then divided by it
    first, as the reference's divide_columns does).  x2 = None: the one-sample form (exact sigma^2 on the diagonal)."""
    def pack(x):
        x = np.asarray(x, dtype=np.float64)
        pts = x.reshape(-1, 1) if x.ndim == 1 else x
        if np.ndim(length_scale) > 0:
            pts = pts / np.asarray(length_scale, dtype=np.float64)[None, :]
        return np.ascontiguousarray(pts.T)   # dimension-major: [dim][n]
    p1 = pack(x1)
    p2 = None if x2 is None else pack(x2)
    dim, n1 = p1.shape
    n2 = n1 if p2 is None else p2.shape[1]
    factor = -0.5 if np.ndim(length_scale) > 0 else -0.5 / (float(length_scale) * float(length_scale))
    cov = np.zeros((n1, n2), order="F")
    check(lib().stan_cuda_gp_exp_quad_cov(n1, n2, dim, p1.ctypes.data, None if p2 is None else p2.ctypes.data,
                                          float(sigma) * float(sigma), factor, cov.ctypes.data, max(n1, 1)))
    return cov
