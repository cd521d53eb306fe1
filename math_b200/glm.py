"""Python mirrors of the reference's entry points over the C ABI (ctypes).

Same names, argument order and error behaviour as the reference functions
  bernoulli_logit_glm_lpmf(y, x, alpha, beta)     stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:46-49
  poisson_log_glm_lpmf(y, x, alpha, beta)         stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp:44-49
  normal_id_glm_lpdf(y, x, alpha, beta, sigma)    stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp:49-53
  neg_binomial_2_log_glm_lpmf(y, x, alpha, beta, phi)  stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp:60-63
  normal_lpdf(y, mu, sigma)                       stan/math/prim/scal/prob/normal_lpdf.hpp:36-39
  bernoulli_logit_lpmf(n, theta)                  stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:28-29
  poisson_log_lpmf(n, alpha)                      stan/math/prim/scal/prob/poisson_log_lpmf.hpp:22-23
Each returns (lp, grads) where grads holds what reverse mode would leave in the adjoints of the
var arguments (alpha, beta, sigma, mu; y on request).  Size mismatches raise ValueError
(std::invalid_argument), domain violations raise DomainError (std::domain_error).
"""
import contextlib
import ctypes as C

import numpy as np

from . import _abi
from ._lib import StanCudaError, check, lib  # noqa: F401


class DomainError(ValueError):
    """Counterpart of std::domain_error thrown by the reference's check_* functions."""


def init(n_devices=0):
    check(lib().stan_cuda_init(n_devices))


def init_device(ordinal):
    check(lib().stan_cuda_init_device(ordinal))


def describe():
    return lib().stan_cuda_describe().decode()


def stats():
    s = _abi.Stats()
    check(lib().stan_cuda_stats(C.byref(s)))
    return {f: getattr(s, f) for f, _ in s._fields_}


def cache_clear():
    lib().stan_cuda_cache_clear()


def cache_invalidate(arr):
    return lib().stan_cuda_cache_invalidate(arr.ctypes.data)


@contextlib.contextmanager
def shard_scope(call_id, job, device):
    """map_rect shard scope: evaluations inside run on `device`, data keyed by (call_id, job)."""
    lib().stan_cuda_shard_scope_enter(call_id, job, device)
    try:
        yield
    finally:
        lib().stan_cuda_shard_scope_exit()


def _f64(a):
    return np.ascontiguousarray(np.atleast_1d(np.asarray(a, dtype=np.float64)))


def _prep_x(x):
    x = np.asarray(x, dtype=np.float64)
    if x.ndim != 2:
        raise ValueError("x must be a matrix")
    if x.flags.f_contiguous:
        return x, 0, max(x.shape[0], 1)
    if x.flags.c_contiguous:
        return x, 1, max(x.shape[1], 1)
    x = np.asfortranarray(x)
    return x, 0, max(x.shape[0], 1)


def _size_check(function, name, v, n, other="Vector of dependent variables"):
    # check_consistent_size(s): scalars are consistent with anything
    if v.size != 1 and v.size != n:
        raise ValueError("%s: %s has dimension = %d, expecting dimension = %d (%s)"
                         % (function, name, v.size, n, other))


def glm_flags(family, propto, alpha_var=True, beta_var=True, sigma_var=True, y_var=False):
    """The include_summand<propto, ...> decisions of the three reference bodies, for the given
    constness of the arguments (x is always data on this path)."""
    any_var = alpha_var or beta_var or (family == _abi.NORMAL_ID and (sigma_var or y_var))
    if family == _abi.NEG_BINOMIAL_2_LOG:      # neg_binomial_2_log_glm_lpmf.hpp:130-153
        f = (_abi.NEED_DBETA if beta_var else 0) | (_abi.NEED_DALPHA if alpha_var else 0) \
            | (_abi.NEED_DSIGMA if sigma_var else 0)
        if not propto:
            f |= _abi.LP_CONST
        if not propto or sigma_var:
            f |= _abi.LP_SCALE
        if not propto or alpha_var or beta_var or sigma_var:
            f |= _abi.LP_MAIN
        if not propto or alpha_var or beta_var:
            f |= _abi.LP_LINEAR
        return f
    f = 0
    if beta_var:
        f |= _abi.NEED_DBETA
    if alpha_var:
        f |= _abi.NEED_DALPHA
    if family == _abi.NORMAL_ID and sigma_var:
        f |= _abi.NEED_DSIGMA
    if family == _abi.NORMAL_ID and y_var:
        f |= _abi.NEED_DY
    if family == _abi.BERNOULLI_LOGIT:
        f |= _abi.LP_MAIN                      # bernoulli_logit_glm_lpmf.hpp:101-104 (no summand test)
    elif family == _abi.POISSON_LOG:
        if not propto:                         # poisson_log_glm_lpmf.hpp:102,109: both tests reduce to !propto
            f |= _abi.LP_CONST | _abi.LP_MAIN
    else:
        if not propto:
            f |= _abi.LP_CONST                 # normal_id_glm_lpdf.hpp:160
        if not propto or sigma_var:
            f |= _abi.LP_SCALE                 # :163
        if not propto or any_var:
            f |= _abi.LP_MAIN                  # :170
    return f


def glm_eval(family, y, x, alpha, beta, sigma=None, flags=None, propto=False, need_dy=False, y_per_call=False,
             need_dx=False, x_per_call=None):
    """Raw C-ABI call (no argument checks beyond shapes).  Returns dict with rc, lp and partials.
    y_per_call: y is not immutable data (a parameter vector): cache keyed on x alone, y uploaded every call.
    need_dx: x is a parameter matrix: "dx" (n, k) = d * beta^T comes back, and (unless x_per_call=False) x is
    uploaded on every call instead of being looked up in the cache."""
    fam = _abi.FAMILIES[family] if isinstance(family, str) else family
    x, row_major, ldx = _prep_x(x)
    n, k = x.shape
    if fam == _abi.NORMAL_ID:
        y = _f64(y)
    else:
        y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    if y.size == 1 and n != 1:
        y = np.full(n, y[0], dtype=y.dtype)
    alpha, beta = _f64(alpha), _f64(beta)
    sigma = _f64(sigma) if sigma is not None else None
    if flags is None:
        flags = glm_flags(fam, propto, y_var=need_dy)
    a = _abi.GlmArgs()
    lp, dbeta, dalpha = np.zeros(1), np.zeros(max(k, 1)), np.zeros(max(alpha.size, 1))
    dsigma = np.zeros(max(sigma.size, 1)) if sigma is not None else np.zeros(1)
    dy = np.zeros(max(n, 1)) if (flags & _abi.NEED_DY) else None
    aux = np.zeros(4)
    a.family, a.flags, a.n, a.k = fam, flags, n, k
    a.x, a.ldx, a.x_row_major = x.ctypes.data, ldx, row_major
    a.y, a.y_is_double = y.ctypes.data, int(fam == _abi.NORMAL_ID)
    a.y_per_call = int(bool(y_per_call))
    dx = None
    if need_dx:
        dx = np.zeros((max(n, 1), max(k, 1)), order="F")
        a.flags = flags = flags | _abi.NEED_DX
        a.dx, a.lddx = dx.ctypes.data, max(n, 1)
    a.x_per_call = int(need_dx if x_per_call is None else bool(x_per_call))
    a.alpha, a.alpha_len, a.beta = alpha.ctypes.data, alpha.size, beta.ctypes.data
    if sigma is not None:
        a.sigma, a.sigma_len = sigma.ctypes.data, sigma.size
    a.lp, a.dbeta, a.dalpha, a.dsigma = lp.ctypes.data, dbeta.ctypes.data, dalpha.ctypes.data, dsigma.ctypes.data
    a.dy = dy.ctypes.data if dy is not None else None
    a.aux = aux.ctypes.data
    rc = check(lib().stan_cuda_glm_eval(C.byref(a)))
    out = {"rc": rc, "lp": float(lp[0]), "dbeta": dbeta[:k].copy(), "dalpha": dalpha[:alpha.size].copy(),
           "aux": aux, "_keep": (x, y)}
    if fam in _abi.HAS_SCALE:
        out["dsigma"] = dsigma[:sigma.size].copy()
    if dy is not None:
        out["dy"] = dy[:n].copy()
    if dx is not None:
        out["dx"] = dx[:n, :k].copy()
    return out


def _lazy_finite_checks(function, fam, rc, x, y, alpha, beta, sigma):
    """The reference's lazy check_finite sequence, run only when the result is not finite
    (bernoulli_logit_glm_lpmf.hpp:106-110, poisson_log_glm_lpmf.hpp:97-101,
    normal_id_glm_lpdf.hpp:150-156).  Quirk kept on purpose: for bernoulli and poisson the last
    check is handed an Eigen::Array, which check_finite treats as ONE element (is_vector<Array> is
    false), so only row 0 of the linear predictor is looked at; if that row is finite the
    reference returns the non-finite value without throwing -- and so does this wrapper."""
    if rc != _abi.NONFINITE:
        return
    if fam == _abi.NORMAL_ID and not np.all(np.isfinite(y)):
        raise DomainError("%s: Vector of dependent variables is not finite" % function)
    if not np.all(np.isfinite(beta)):
        raise DomainError("%s: Weight vector is not finite" % function)
    if not np.all(np.isfinite(alpha)):
        raise DomainError("%s: Intercept is not finite" % function)
    if fam in _abi.HAS_SCALE:   # normal_id :150-156; neg_binomial_2_log :120 (whole theta vector checked)
        raise DomainError("%s: Matrix of independent variables is not finite" % function)
    with np.errstate(all="ignore"):
        theta0 = float(np.dot(x[0, :], np.broadcast_to(beta, (x.shape[1],))) + alpha[0])
    if not np.isfinite(theta0):
        raise DomainError("%s: Matrix of independent variables[1] is not finite" % function)


def _glm(function, fam, y, x, alpha, beta, sigma, propto, need_dy=False):
    xa = np.asarray(x, dtype=np.float64)
    if xa.ndim != 2:
        raise ValueError("%s: x must be a matrix" % function)
    n, k = xa.shape
    ya = np.atleast_1d(np.asarray(y))
    al, be = _f64(alpha), _f64(beta)
    sg = _f64(sigma) if sigma is not None else None
    if fam == _abi.NORMAL_ID:
        if not (np.all(np.isfinite(sg)) and np.all(sg > 0)):       # check_positive_finite, :70
            raise DomainError("%s: Scale vector must be positive finite" % function)
    if fam == _abi.NEG_BINOMIAL_2_LOG:   # eager parameter checks, neg_binomial_2_log_glm_lpmf.hpp:90-92
        if not np.all(np.isfinite(be)):
            raise DomainError("%s: Weight vector is not finite" % function)
        if not np.all(np.isfinite(al)):
            raise DomainError("%s: Intercept is not finite" % function)
        if not (np.all(np.isfinite(sg)) and np.all(sg > 0)):
            raise DomainError("%s: Precision parameter must be positive finite" % function)
    _size_check(function, "Vector of dependent variables", ya, n)
    if be.size != k and not (be.size == 1 and k == 1):
        raise ValueError("%s: Weight vector has dimension = %d, expecting dimension = %d" % (function, be.size, k))
    if sg is not None:
        _size_check(function, "Vector of scale parameters", sg, n)
    _size_check(function, "Vector of intercepts", al, n)
    if n == 0 or (fam in _abi.HAS_SCALE and sg.size == 0):
        return 0.0, {}
    r = glm_eval(fam, ya, xa, al, be, sg, propto=propto, need_dy=need_dy)
    if r["rc"] == _abi.BAD_OUTCOME:
        # check_bounded / check_nonnegative (:70): answered by the device from the range of y computed
        # once at cache fill; the host loop runs only here, to word the exception
        if fam == _abi.BERNOULLI_LOGIT:
            raise DomainError("%s: Vector of dependent variables must be in the interval [0, 1]" % function)
        if fam == _abi.NEG_BINOMIAL_2_LOG:
            raise DomainError("%s: Failures variables must be >= 0" % function)
        raise DomainError("%s: Vector of dependent variables must be >= 0" % function)
    _lazy_finite_checks(function, fam, r["rc"], xa, ya, al, be, sg)
    grads = {"beta": r["dbeta"], "alpha": r["dalpha"]}
    if fam == _abi.NORMAL_ID:
        grads["sigma"] = r["dsigma"]
    if fam == _abi.NEG_BINOMIAL_2_LOG:
        grads["phi"] = r["dsigma"]
    if need_dy:
        grads["y"] = r["dy"]
    return r["lp"], grads


def bernoulli_logit_glm_lpmf(y, x, alpha, beta, propto=False):
    return _glm("bernoulli_logit_glm_lpmf", _abi.BERNOULLI_LOGIT, y, x, alpha, beta, None, propto)


def poisson_log_glm_lpmf(y, x, alpha, beta, propto=False):
    return _glm("poisson_log_glm_lpmf", _abi.POISSON_LOG, y, x, alpha, beta, None, propto)


def normal_id_glm_lpdf(y, x, alpha, beta, sigma, propto=False, need_dy=False):
    return _glm("normal_id_glm_lpdf", _abi.NORMAL_ID, y, x, alpha, beta, sigma, propto, need_dy)


def neg_binomial_2_log_glm_lpmf(y, x, alpha, beta, phi, propto=False):
    """stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp:60-63; grads: beta, alpha, phi."""
    return _glm("neg_binomial_2_log_glm_lpmf", _abi.NEG_BINOMIAL_2_LOG, y, x, alpha, beta, phi, propto)


def ordered_logistic_glm_eval(y, x, beta, cuts, flags):
    """Raw C-ABI call for family ORDERED_LOGISTIC: dict with rc, lp, dbeta, dcuts."""
    x, row_major, ldx = _prep_x(x)
    n, k = x.shape
    y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    beta = _f64(beta)
    cuts = _f64(cuts) if np.size(cuts) else np.zeros(0)
    a = _abi.GlmArgs()
    lp, dbeta, dcuts, aux = np.zeros(1), np.zeros(max(k, 1)), np.zeros(max(cuts.size, 1)), np.zeros(4)
    a.family, a.flags, a.n, a.k = _abi.ORDERED_LOGISTIC, flags, n, k
    a.x, a.ldx, a.x_row_major = x.ctypes.data, ldx, row_major
    a.y, a.y_is_double = y.ctypes.data, 0
    a.alpha, a.alpha_len, a.beta = None, 1, beta.ctypes.data
    a.cuts, a.cuts_len, a.dcuts = cuts.ctypes.data, cuts.size, dcuts.ctypes.data
    a.lp, a.dbeta, a.aux = lp.ctypes.data, dbeta.ctypes.data, aux.ctypes.data
    rc = check(lib().stan_cuda_glm_eval(C.byref(a)))
    return {"rc": rc, "lp": float(lp[0]), "dbeta": dbeta[:k].copy(), "dcuts": dcuts[:cuts.size].copy(), "aux": aux,
            "_keep": (x, y)}


def ordered_logistic_glm_lpmf(y, x, beta, cuts, propto=False, beta_var=True, cuts_var=True):
    """stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:46-52; grads: beta, cuts.
    Checks in the reference's order (:71-86); check_bounded(y, 1, N_classes) is answered by the
    device from the range of y cached at fill."""
    function = "ordered_logistic_glm_lpmf"
    xa = np.asarray(x, dtype=np.float64)
    if xa.ndim != 2:
        raise ValueError("%s: x must be a matrix" % function)
    n, k = xa.shape
    ya = np.atleast_1d(np.asarray(y))
    be = _f64(beta)
    cu = _f64(cuts) if np.size(cuts) else np.zeros(0)
    n_classes = cu.size + 1
    if ya.size != n:
        raise ValueError("%s: Vector of dependent variables has dimension = %d, expecting dimension = %d"
                         % (function, ya.size, n))
    if be.size != k:
        raise ValueError("%s: Weight vector has dimension = %d, expecting dimension = %d" % (function, be.size, k))

    def y_error():
        return DomainError("%s: Vector of dependent variables must be in the interval [1, %d]" % (function, n_classes))

    def cuts_error():
        if cu.size > 1 and not np.all(cu[1:] > cu[:-1]):   # check_ordered: strictly ascending, NaN fails
            return DomainError("%s: Cut-points is not a valid ordered vector" % function)
        if cu.size and not (np.isfinite(cu[-1]) and np.isfinite(cu[0])):
            return DomainError("%s: cut-point is not finite" % function)
        return None

    ce = cuts_error()
    nothing_to_do = n == 0 or cu.size == 0 or (propto and not (beta_var or cuts_var))
    if ce is not None or nothing_to_do:
        # error path or early return (:88-92): nothing goes to the device, the outcome check runs here
        if n and (ya.min() < 1 or ya.max() > n_classes):
            raise y_error()
        if ce is not None:
            raise ce
        return 0.0, {}
    flags = _abi.LP_MAIN | (_abi.NEED_DBETA if beta_var else 0) | (_abi.NEED_DCUTS if cuts_var else 0)
    r = ordered_logistic_glm_eval(ya, xa, be, cu, flags)
    if r["rc"] == _abi.BAD_OUTCOME:
        raise y_error()
    if r["rc"] == _abi.NONFINITE:      # :119-122
        if not np.all(np.isfinite(be)):
            raise DomainError("%s: Weight vector is not finite" % function)
        if not np.all(np.isfinite(xa)):
            raise DomainError("%s: Matrix of independent variables is not finite" % function)
    grads = {}
    if beta_var:
        grads["beta"] = r["dbeta"]
    if cuts_var:
        grads["cuts"] = r["dcuts"]
    return r["lp"], grads


def categorical_logit_glm_eval(y, x, alpha, beta, flags):
    """Raw C-ABI call for family CATEGORICAL_LOGIT; beta is (k, C).  dict with rc, lp, dalpha (C), dbeta (k, C)."""
    x, row_major, ldx = _prep_x(x)
    n, k = x.shape
    y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    alpha = _f64(alpha)
    beta = np.asfortranarray(np.asarray(beta, dtype=np.float64).reshape(k, -1))
    ncls = beta.shape[1]
    a = _abi.GlmArgs()
    lp, dalpha, aux = np.zeros(1), np.zeros(max(ncls, 1)), np.zeros(4)
    dbeta = np.zeros((max(k, 1), max(ncls, 1)), order="F")
    a.family, a.flags, a.n, a.k, a.n_classes = _abi.CATEGORICAL_LOGIT, flags, n, k, ncls
    a.x, a.ldx, a.x_row_major = x.ctypes.data, ldx, row_major
    a.y, a.y_is_double = y.ctypes.data, 0
    a.alpha, a.alpha_len, a.beta = alpha.ctypes.data, alpha.size, beta.ctypes.data
    a.lp, a.dbeta, a.dalpha, a.aux = lp.ctypes.data, dbeta.ctypes.data, dalpha.ctypes.data, aux.ctypes.data
    rc = check(lib().stan_cuda_glm_eval(C.byref(a)))
    return {"rc": rc, "lp": float(lp[0]), "dalpha": dalpha[:ncls].copy(),
            "dbeta": dbeta.reshape(-1, order="F")[:k * ncls].reshape((k, ncls), order="F").copy(), "_keep": (x, y)}


def categorical_logit_glm_lpmf(y, x, alpha, beta, propto=False, alpha_var=True, beta_var=True):
    """stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:38-42; beta is (k, C); grads: alpha, beta.
    Checks in the reference's order (:56-64); the outcome range comes from the device."""
    function = "categorical_logit_glm_lpmf"
    xa = np.asarray(x, dtype=np.float64)
    be = np.asarray(beta, dtype=np.float64)
    if xa.ndim != 2 or be.ndim != 2:
        raise ValueError("%s: x and beta must be matrices" % function)
    n, k = xa.shape
    ncls = be.shape[1]
    ya = np.atleast_1d(np.asarray(y))
    al = _f64(alpha) if np.size(alpha) else np.zeros(0)
    if ya.size != n:
        raise ValueError("%s: Vector of dependent variables has dimension = %d, expecting dimension = %d"
                         % (function, ya.size, n))
    if al.size != ncls:
        raise ValueError("%s: Intercept vector has dimension = %d, expecting dimension = %d" % (function, al.size, ncls))
    if be.shape[0] != k:
        raise ValueError("%s: x.cols() (%d) and beta.rows() (%d) must match in size" % (function, k, be.shape[0]))
    y_error = DomainError("%s: categorical outcome out of support must be in the interval [1, %d]" % (function, ncls))
    if n == 0 or ncls == 1 or (propto and not (alpha_var or beta_var)):
        if n and (ya.min() < 1 or ya.max() > ncls):     # nothing goes to the device: the outcome check runs here
            raise y_error
        return 0.0, {}
    flags = _abi.LP_MAIN | (_abi.NEED_DBETA if beta_var else 0) | (_abi.NEED_DALPHA if alpha_var else 0)
    r = categorical_logit_glm_eval(ya, xa, al, be, flags)
    if r["rc"] == _abi.BAD_OUTCOME:
        raise y_error
    if r["rc"] == _abi.NONFINITE:       # :109-113
        if not np.all(np.isfinite(be)):
            raise DomainError("%s: Weight vector is not finite" % function)
        if not np.all(np.isfinite(al)):
            raise DomainError("%s: Intercept is not finite" % function)
        if not np.all(np.isfinite(xa)):
            raise DomainError("%s: Matrix of independent variables is not finite" % function)
    grads = {}
    if alpha_var:
        grads["alpha"] = r["dalpha"]
    if beta_var:
        grads["beta"] = r["dbeta"]
    return r["lp"], grads


def normal_lpdf(y, mu, sigma, propto=False, need_dy=False, y_is_data=True):
    """normal_lpdf(y | mu, sigma) with mu and sigma as the differentiated operands."""
    function = "normal_lpdf"
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    if min(y.size, mu.size, sigma.size) == 0:
        return 0.0, {}
    if np.any(np.isnan(y)):
        raise DomainError("%s: Random variable is nan" % function)          # check_not_nan, :51
    if not np.all(np.isfinite(mu)):
        raise DomainError("%s: Location parameter must be finite" % function)  # check_finite, :52
    if not np.all(sigma > 0):
        raise DomainError("%s: Scale parameter must be > 0" % function)      # check_positive, :53
    n = max(y.size, mu.size, sigma.size)
    for name, v in (("Random variable", y), ("Location parameter", mu), ("Scale parameter", sigma)):
        if v.size != 1 and v.size != n:
            raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (function, name, v.size, n))
    flags = _abi.NEED_DALPHA | _abi.NEED_DSIGMA | (_abi.NEED_DY if need_dy else 0)
    flags |= (_abi.LP_SCALE | _abi.LP_MAIN) if propto else _abi.LP_ALL
    a = _abi.NormalArgs()
    lp, dmu, dsigma, dy = np.zeros(1), np.zeros(mu.size), np.zeros(sigma.size), np.zeros(y.size)
    a.flags, a.y_is_data, a.n = flags, int(y_is_data), n
    a.y, a.y_len, a.mu, a.mu_len = y.ctypes.data, y.size, mu.ctypes.data, mu.size
    a.sigma, a.sigma_len = sigma.ctypes.data, sigma.size
    a.lp, a.dmu, a.dsigma, a.dy = lp.ctypes.data, dmu.ctypes.data, dsigma.ctypes.data, dy.ctypes.data
    rc = check(lib().stan_cuda_normal_lpdf_eval(C.byref(a)))
    if rc == _abi.BAD_OPERAND:   # unreachable after the checks above; kept for symmetry with the C++ wrapper
        raise DomainError("%s: operand outside the domain" % function)
    grads = {"mu": dmu, "sigma": dsigma}
    if need_dy:
        grads["y"] = dy
    return float(lp[0]), grads


def locscale_eval(density, y, mu, sigma, flags, y_is_data=True):
    """One stan_cuda_locscale_lpdf_eval call (density: "normal" | "cauchy" | "lognormal"); returns {"rc", "lp", "dy",
    "dmu", "dsigma"} (partials for the NEED_* bits in flags, each the length of its operand)."""
    d = _abi.LOCSCALES[density] if isinstance(density, str) else density
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    n = max(y.size, mu.size, sigma.size) if min(y.size, mu.size, sigma.size) > 0 else 0
    a = _abi.NormalArgs()
    lp = np.zeros(1)
    dy, dmu, dsigma = np.zeros(max(y.size, 1)), np.zeros(max(mu.size, 1)), np.zeros(max(sigma.size, 1))
    a.flags, a.y_is_data, a.n = flags, int(y_is_data), n
    a.y, a.y_len, a.mu, a.mu_len = y.ctypes.data, y.size, mu.ctypes.data, mu.size
    a.sigma, a.sigma_len = sigma.ctypes.data, sigma.size
    a.lp, a.dmu, a.dsigma, a.dy = lp.ctypes.data, dmu.ctypes.data, dsigma.ctypes.data, dy.ctypes.data
    rc = check(lib().stan_cuda_locscale_lpdf_eval(d, C.byref(a)))
    res = {"rc": rc, "lp": float(lp[0])}
    for key, bit, g, op in (("dy", _abi.NEED_DY, dy, y), ("dmu", _abi.NEED_DALPHA, dmu, mu), ("dsigma", _abi.NEED_DSIGMA, dsigma, sigma)):
        if flags & bit:
            res[key] = g[:op.size].copy()
    return res


def _locscale_lpdf(function, density, y, mu, sigma, propto, var_mask, y_is_data):
    """Call sequence of cauchy_lpdf.hpp:42-58 / lognormal_lpdf.hpp:24-45: domain errors in the reference's order, then the
    size error, nothing to compute -> 0 (lognormal: LOG_ZERO comes first), partials only for the variables."""
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    lognormal = density == "lognormal"
    if min(y.size, mu.size, sigma.size) == 0 and not lognormal:
        return 0.0, {}
    if np.any(np.isnan(y)):
        raise DomainError("%s: Random variable is nan" % function)
    if lognormal and np.any(y < 0):
        raise DomainError("%s: Random variable must be >= 0" % function)
    if not np.all(np.isfinite(mu)):
        raise DomainError("%s: Location parameter must be finite" % function)
    if not np.all((sigma > 0) & np.isfinite(sigma)):
        raise DomainError("%s: Scale parameter must be positive finite" % function)
    if min(y.size, mu.size, sigma.size) == 0:
        return 0.0, {}
    n = max(y.size, mu.size, sigma.size)
    for name, v in (("Random variable", y), ("Location parameter", mu), ("Scale parameter", sigma)):
        if v.size != 1 and v.size != n:
            raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (function, name, v.size, n))
    names = (("y", "dy", 1), ("mu", "dmu", 2), ("sigma", "dsigma", 4))
    if propto and var_mask == 0:
        if lognormal and np.any(y <= 0):
            return -np.inf, {}
        return 0.0, {}
    flags = locscale_flags(density, propto, var_mask)
    r = locscale_eval(density, y, mu, sigma, flags, y_is_data=y_is_data and not (var_mask & 1))
    if r["rc"] == _abi.BAD_OPERAND:   # unreachable after the checks above; kept for symmetry with the C++ wrapper
        raise DomainError("%s: operand outside the domain" % function)
    if r["rc"] == _abi.LOG_ZERO:      # lognormal_lpdf.hpp:41-45: an outcome equal to zero, no partials
        return -np.inf, {}
    return r["lp"], {name: r[key] for name, key, bit in names if var_mask & bit}


def locscale_flags(density, propto=False, var_mask=0b110):
    """Which summands the reference keeps under include_summand<propto, ...> (normal_lpdf.hpp:89-97,
    cauchy_lpdf.hpp:87-95, lognormal_lpdf.hpp:92-118) for the operands that are variables (var_mask bit 0: y, 1: mu,
    2: sigma), and which partials they need."""
    d = _abi.LOCSCALES[density] if isinstance(density, str) else density
    vy, vm, vs = bool(var_mask & 1), bool(var_mask & 2), bool(var_mask & 4)
    f = (_abi.NEED_DY if vy else 0) | (_abi.NEED_DALPHA if vm else 0) | (_abi.NEED_DSIGMA if vs else 0)
    if d in (_abi.LOCSCALE_GAMMA, _abi.LOCSCALE_INV_GAMMA):   # gamma_lpdf.hpp:118-129, inv_gamma_lpdf.hpp:115-126 (alpha = mu, beta = sigma)
        f |= (_abi.LP_DOF if (not propto or vm) else 0) | (_abi.LP_SCALE if (not propto or vm or vs) else 0)
        f |= (_abi.LP_Y if (not propto or vy or vm) else 0) | (_abi.LP_MAIN if (not propto or vy or vs) else 0)
        return f
    if d == _abi.LOCSCALE_BETA:         # beta_lpdf.hpp:150-164 (alpha = the mu operand, beta = the sigma operand)
        f |= (_abi.LP_SCALE if (not propto or vm or vs) else 0) | (_abi.LP_DOF if (not propto or vm) else 0)
        f |= (_abi.LP_LINEAR if (not propto or vs) else 0) | (_abi.LP_Y if (not propto or vy or vm) else 0)
        return f | (_abi.LP_MAIN if (not propto or vy or vs) else 0)
    if d == _abi.LOCSCALE_WEIBULL:      # weibull_lpdf.hpp:111-123 (alpha = the mu operand)
        f |= (_abi.LP_DOF if (not propto or vm) else 0) | (_abi.LP_Y if (not propto or vy or vm) else 0)
        f |= (_abi.LP_SCALE if (not propto or vm or vs) else 0)
        return f | (_abi.LP_MAIN if (not propto or vy or vm or vs) else 0)
    if d == _abi.LOCSCALE_EXPONENTIAL:  # exponential_lpdf.hpp:76-81 (beta = the sigma operand)
        f &= ~_abi.NEED_DALPHA
        return f | (_abi.LP_SCALE if (not propto or vs) else 0) | (_abi.LP_MAIN if (not propto or vy or vs) else 0)
    if not propto:
        f |= _abi.LP_CONST
    if not propto or vs:
        f |= _abi.LP_SCALE
    if not propto or vy or vm or vs:
        f |= _abi.LP_MAIN
    if d == _abi.LOCSCALE_LOGNORMAL and (not propto or vy):
        f |= _abi.LP_Y
    return f


def cauchy_lpdf(y, mu, sigma, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/cauchy_lpdf.hpp:35-37; var_mask: which of y (bit 0), mu (1), sigma (2) are variables;
    grads: the partials of those."""
    return _locscale_lpdf("cauchy_lpdf", "cauchy", y, mu, sigma, propto, var_mask, y_is_data)


def lognormal_lpdf(y, mu, sigma, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/lognormal_lpdf.hpp:19-21; var_mask and grads as for cauchy_lpdf."""
    return _locscale_lpdf("lognormal_lpdf", "lognormal", y, mu, sigma, propto, var_mask, y_is_data)

def gamma_lpdf(y, alpha, beta, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/gamma_lpdf.hpp:41-43; var_mask: which of y (bit 0), alpha (1), beta (2) are variables;
    grads: the partials of those.  Call sequence :48-74: size_zero, domain errors, size error, nothing to compute -> 0,
    a negative outcome -> LOG_ZERO."""
    f = "gamma_lpdf"
    y, alpha, beta = _f64(y), _f64(alpha), _f64(beta)
    if min(y.size, alpha.size, beta.size) == 0:
        return 0.0, {}
    if np.any(np.isnan(y)):
        raise DomainError("%s: Random variable is nan" % f)
    if not np.all((alpha > 0) & np.isfinite(alpha)):
        raise DomainError("%s: Shape parameter must be positive finite" % f)
    if not np.all((beta > 0) & np.isfinite(beta)):
        raise DomainError("%s: Inverse scale parameter must be positive finite" % f)
    n = max(y.size, alpha.size, beta.size)
    for name, v in (("Random variable", y), ("Shape parameter", alpha), ("Inverse scale parameter", beta)):
        if v.size != 1 and v.size != n:
            raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (f, name, v.size, n))
    if propto and var_mask == 0:
        return 0.0, {}
    r = locscale_eval("gamma", y, alpha, beta, locscale_flags("gamma", propto, var_mask), y_is_data=y_is_data and not (var_mask & 1))
    if r["rc"] == _abi.BAD_OPERAND:
        raise DomainError("%s: operand outside the domain" % f)
    if r["rc"] == _abi.LOG_ZERO:
        return -np.inf, {}
    return r["lp"], {name: r[key] for name, key, bit in (("y", "dy", 1), ("alpha", "dmu", 2), ("beta", "dsigma", 4)) if var_mask & bit}


def _shape_density(f, density, y, alpha, second, propto, var_mask, y_is_data, y_check, second_name, names, log_zero=None):
    """beta_lpdf.hpp:48-60 / weibull_lpdf.hpp:38-49: domain errors, size error, nothing to compute -> 0."""
    y, alpha, second = _f64(y), _f64(alpha), _f64(second)
    if not np.all((alpha > 0) & np.isfinite(alpha)):
        raise DomainError("%s: %s must be positive finite" % (f, "First shape parameter" if density == "beta" else "Shape parameter"))
    if not np.all((second > 0) & np.isfinite(second)):
        raise DomainError("%s: %s must be positive finite" % (f, second_name))
    y_check(y)
    if min(y.size, alpha.size, second.size) == 0:
        return 0.0, {}
    n = max(y.size, alpha.size, second.size)
    for name, v in (("Random variable", y), ("Shape parameter", alpha), (second_name, second)):
        if v.size != 1 and v.size != n:
            raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (f, name, v.size, n))
    if propto and var_mask == 0:
        return 0.0, {}
    r = locscale_eval(density, y, alpha, second, locscale_flags(density, propto, var_mask), y_is_data=y_is_data and not (var_mask & 1))
    if r["rc"] == _abi.BAD_OPERAND:
        raise DomainError("%s: operand outside the domain" % f)
    if r["rc"] == _abi.LOG_ZERO:
        return -np.inf, {}
    return r["lp"], {name: r[key] for name, key, bit in names if var_mask & bit}


def beta_lpdf(y, alpha, beta, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/beta_lpdf.hpp:40-43; var_mask: which of y (bit 0), alpha (1), beta (2) are variables."""
    def y_check(y):
        if np.any(np.isnan(y)):
            raise DomainError("beta_lpdf: Random variable is nan")
        if not np.all((y >= 0) & (y <= 1)):
            raise DomainError("beta_lpdf: Random variable must be in [0, 1]")
    return _shape_density("beta_lpdf", "beta", y, alpha, beta, propto, var_mask, y_is_data, y_check, "Second shape parameter",
                          (("y", "dy", 1), ("alpha", "dmu", 2), ("beta", "dsigma", 4)))


def weibull_lpdf(y, alpha, sigma, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/weibull_lpdf.hpp:29-31; var_mask: which of y (bit 0), alpha (1), sigma (2) are variables;
    a negative outcome is LOG_ZERO (:56-61)."""
    def y_check(y):
        if not np.all(np.isfinite(y)):
            raise DomainError("weibull_lpdf: Random variable must be finite")
    return _shape_density("weibull_lpdf", "weibull", y, alpha, sigma, propto, var_mask, y_is_data, y_check, "Scale parameter",
                          (("y", "dy", 1), ("alpha", "dmu", 2), ("sigma", "dsigma", 4)))


def inv_gamma_lpdf(y, alpha, beta, propto=False, var_mask=0b110, y_is_data=True):
    """stan/math/prim/scal/prob/inv_gamma_lpdf.hpp:33-35; var_mask: which of y (bit 0), alpha (1), beta (2) are variables;
    an outcome that is not positive is LOG_ZERO (:58-63)."""
    def y_check(y):
        if np.any(np.isnan(y)):
            raise DomainError("inv_gamma_lpdf: Random variable is nan")
    # (the reference tests y first, :41-43; with the operands all valid or one invalid the order is not observable here)
    return _shape_density("inv_gamma_lpdf", "inv_gamma", y, alpha, beta, propto, var_mask, y_is_data, y_check, "Scale parameter",
                          (("y", "dy", 1), ("alpha", "dmu", 2), ("beta", "dsigma", 4)))


def exponential_lpdf(y, beta, propto=False, var_mask=0b10, y_is_data=True):
    """stan/math/prim/scal/prob/exponential_lpdf.hpp:42-44; var_mask: which of y (bit 0), beta (bit 1) are variables."""
    f = "exponential_lpdf"
    y, beta = _f64(y), _f64(beta)
    if min(y.size, beta.size) == 0:
        return 0.0, {}
    if not np.all(y >= 0):
        raise DomainError("%s: Random variable must be >= 0" % f)
    if not np.all((beta > 0) & np.isfinite(beta)):
        raise DomainError("%s: Inverse scale parameter must be positive finite" % f)
    n = max(y.size, beta.size)
    for name, v in (("Random variable", y), ("Inverse scale parameter", beta)):
        if v.size != 1 and v.size != n:
            raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (f, name, v.size, n))
    mask3 = (var_mask & 1) | ((var_mask & 2) << 1)   # y -> bit 0, beta -> the sigma operand's bit 2
    if propto and var_mask == 0:
        return 0.0, {}
    r = locscale_eval("exponential", y, [0.0], beta, locscale_flags("exponential", propto, mask3),
                      y_is_data=y_is_data and not (var_mask & 1))
    if r["rc"] == _abi.BAD_OPERAND:
        raise DomainError("%s: operand outside the domain" % f)
    return r["lp"], {name: r[key] for name, key, bit in (("y", "dy", 1), ("beta", "dsigma", 4)) if mask3 & bit}


def student_t_eval(y, nu, mu, sigma, flags, y_is_data=True):
    """One stan_cuda_student_t_lpdf_eval call; returns {"rc", "lp", "dy", "dnu", "dmu", "dsigma"} (partials for the
    NEED_* bits in flags, each the length of its operand)."""
    ops = [_f64(v) for v in (y, nu, mu, sigma)]
    n = max(o.size for o in ops) if min(o.size for o in ops) > 0 else 0
    lp = np.zeros(1)
    grads = [np.zeros(max(o.size, 1)) for o in ops]
    a = _abi.StudentTArgs()
    a.flags, a.y_is_data, a.n = flags, int(y_is_data), n
    a.y, a.y_len, a.nu, a.nu_len = ops[0].ctypes.data, ops[0].size, ops[1].ctypes.data, ops[1].size
    a.mu, a.mu_len, a.sigma, a.sigma_len = ops[2].ctypes.data, ops[2].size, ops[3].ctypes.data, ops[3].size
    a.lp, a.dy, a.dnu, a.dmu, a.dsigma = (lp.ctypes.data, grads[0].ctypes.data, grads[1].ctypes.data,
                                          grads[2].ctypes.data, grads[3].ctypes.data)
    rc = check(lib().stan_cuda_student_t_lpdf_eval(C.byref(a)))
    res = {"rc": rc, "lp": float(lp[0])}
    for name, bit, g, o in zip(("dy", "dnu", "dmu", "dsigma"),
                               (_abi.NEED_DY, _abi.NEED_DNU, _abi.NEED_DALPHA, _abi.NEED_DSIGMA), grads, ops):
        if flags & bit:
            res[name] = g[:o.size]
    return res


def student_t_lpdf(y, nu, mu, sigma, propto=False, y_var=False, nu_var=True, mu_var=True, sigma_var=True,
                   y_is_data=True):
    """student_t_lpdf(y | nu, mu, sigma), stan/math/prim/scal/prob/student_t_lpdf.hpp:55-201; grads for the operands
    marked *_var.  Call sequence of the reference: size_zero -> 0 (:57-59), the four domain checks (:67-70), the
    size check (:71-73), nothing to compute under propto with constant operands -> 0 (:75-77)."""
    function = "student_t_lpdf"
    ops = [_f64(v) for v in (y, nu, mu, sigma)]
    names = ("Random variable", "Degrees of freedom parameter", "Location parameter", "Scale parameter")
    if min(o.size for o in ops) == 0:
        return 0.0, {}

    def domain_checks():
        if np.any(np.isnan(ops[0])):
            raise DomainError("%s: %s is nan, but must not be nan!" % (function, names[0]))
        for q in (1, 3):
            if not np.all((ops[q] > 0) & np.isfinite(ops[q])):
                raise DomainError("%s: %s must be positive finite!" % (function, names[q]))
            if q == 1 and not np.all(np.isfinite(ops[2])):
                raise DomainError("%s: %s must be finite!" % (function, names[2]))

    n = max(o.size for o in ops)
    sizes_ok = all(o.size in (1, n) for o in ops)
    var_mask = (1 if y_var else 0) | (2 if nu_var else 0) | (4 if mu_var else 0) | (8 if sigma_var else 0)
    if not sizes_ok or (propto and var_mask == 0):
        domain_checks()   # no device work: the reference's eager checks, in its order
        for name, o in zip(names, ops):
            if o.size not in (1, n):
                raise ValueError("%s: %s has dimension = %d, expecting dimension = %d" % (function, name, o.size, n))
        return 0.0, {}
    need = ((_abi.NEED_DY if y_var else 0) | (_abi.NEED_DNU if nu_var else 0) | (_abi.NEED_DALPHA if mu_var else 0)
            | (_abi.NEED_DSIGMA if sigma_var else 0))
    if propto:
        lpf = (_abi.LP_DOF if nu_var else 0) | (_abi.LP_SCALE if sigma_var else 0) | _abi.LP_MAIN
    else:
        lpf = _abi.LP_CONST | _abi.LP_DOF | _abi.LP_SCALE | _abi.LP_MAIN
    r = student_t_eval(ops[0], ops[1], ops[2], ops[3], need | lpf, y_is_data=y_is_data and not y_var)
    if r["rc"] == _abi.BAD_OPERAND:   # the kernel saw a bad element: the reference's checks name it
        domain_checks()
        raise DomainError("%s: operand outside the domain" % function)
    grads = {}
    for key, name in (("dy", "y"), ("dnu", "nu"), ("dmu", "mu"), ("dsigma", "sigma")):
        if key in r:
            grads[name] = r[key]
    return r["lp"], grads


class GlmBatch:
    """`count` independent GLMs of one family and one K evaluated with ONE kernel launch per device
    (stan_cuda_glm_eval_batch): the jobs of a map_rect call whose functor is a single GLM call.
    The argument blocks are built once; `eval(alphas, betas, sigmas)` only rewrites the parameters."""

    def __init__(self, family, xs, ys, call_id=0, flags=None):
        fam = _abi.FAMILIES[family] if isinstance(family, str) else family
        self.fam, self.count, self.call_id = fam, len(xs), call_id
        self.args = (_abi.GlmArgs * self.count)()
        self.keep, self.k = [], None
        if flags is None:
            flags = glm_flags(fam, False)
        for j, (x, y) in enumerate(zip(xs, ys)):
            x, row_major, ldx = _prep_x(x)
            n, k = x.shape
            self.k = k if self.k is None else self.k
            y = _f64(y) if fam == _abi.NORMAL_ID else np.ascontiguousarray(np.asarray(y, dtype=np.int32))
            bufs = {"alpha": np.zeros(1), "beta": np.zeros(k), "sigma": np.ones(1), "lp": np.zeros(1), "dbeta": np.zeros(k),
                    "dalpha": np.zeros(1), "dsigma": np.zeros(1)}
            a = self.args[j]
            a.family, a.flags, a.n, a.k = fam, flags, n, k
            a.x, a.ldx, a.x_row_major = x.ctypes.data, ldx, row_major
            a.y, a.y_is_double = y.ctypes.data, int(fam == _abi.NORMAL_ID)
            a.alpha, a.alpha_len, a.beta = bufs["alpha"].ctypes.data, 1, bufs["beta"].ctypes.data
            if fam in _abi.HAS_SCALE:
                a.sigma, a.sigma_len = bufs["sigma"].ctypes.data, 1
            a.lp, a.dbeta, a.dalpha, a.dsigma = (bufs["lp"].ctypes.data, bufs["dbeta"].ctypes.data,
                                                 bufs["dalpha"].ctypes.data, bufs["dsigma"].ctypes.data)
            self.keep.append((x, y, bufs))
        self.item_rc = (C.c_int * self.count)()

    def eval(self, alphas, betas, sigmas=None):
        """alphas: (count,), betas: (count, k) or (k,) shared, sigmas: (count,) -> list of result dicts."""
        betas = np.asarray(betas, float)
        for j, (_, _, b) in enumerate(self.keep):
            b["alpha"][0] = alphas[j]
            b["beta"][:] = betas if betas.ndim == 1 else betas[j]
            if sigmas is not None:
                b["sigma"][0] = sigmas[j]
        rc = self.call()
        return rc, self.results()

    def call(self):
        """The C-ABI call alone, with the parameters last written by eval()."""
        return check(lib().stan_cuda_glm_eval_batch(self.args, self.count, self.call_id, self.item_rc))

    def results(self):
        return [{"rc": self.item_rc[j], "lp": float(b["lp"][0]), "dbeta": b["dbeta"].copy(), "dalpha": b["dalpha"].copy(),
                     "dsigma": b["dsigma"].copy()} for j, (_, _, b) in enumerate(self.keep)]


def density_eval(density, n, theta, flags):
    """Raw C-ABI call stan_cuda_density_eval; dict with rc, lp, dtheta (theta's shape)."""
    d = _abi.DENSITIES[density] if isinstance(density, str) else density
    n = np.ascontiguousarray(np.atleast_1d(np.asarray(n, dtype=np.int32)))
    theta = _f64(theta)
    size = max(n.size, theta.size) if min(n.size, theta.size) > 0 else 0
    lp, dth = np.zeros(1), np.zeros(max(theta.size, 1))
    a = _abi.DensityArgs()
    a.density, a.flags, a.n = d, flags, size
    a.y, a.y_len, a.theta, a.theta_len = n.ctypes.data, n.size, theta.ctypes.data, theta.size
    a.lp, a.dtheta = lp.ctypes.data, dth.ctypes.data
    rc = check(lib().stan_cuda_density_eval(C.byref(a)))
    return {"rc": rc, "lp": float(lp[0]), "dtheta": dth[:theta.size].copy()}


def _density(function, density, n, theta, propto, theta_var, n_check, n_error, theta_name, theta_check=None,
             theta_error=None):
    """Call sequence of the reference (bernoulli_logit_lpmf.hpp:36-49, poisson_log_lpmf.hpp:32-49): size_zero -> 0,
    domain errors before the size error, nothing to compute under propto with a constant operand -> 0."""
    nn = np.atleast_1d(np.asarray(n))
    th = _f64(theta)
    if nn.size == 0 or th.size == 0:
        return 0.0, {"theta": np.zeros(th.size)}
    size = max(nn.size, th.size)
    sizes_ok = nn.size in (1, size) and th.size in (1, size)
    if not sizes_ok or (propto and not theta_var):
        # no device work: the reference's eager checks, in its order
        if not n_check(nn):
            raise n_error
        if np.any(np.isnan(th)):
            raise DomainError("%s: %s is nan, but must not be nan!" % (function, theta_name))
        if theta_check is not None and not theta_check(th):
            raise theta_error
        if not sizes_ok:
            raise ValueError("%s: Random variable has dimension = %d, expecting dimension = %d" % (function, nn.size, size))
        return 0.0, {"theta": np.zeros(th.size)}
    flags = (_abi.LP_MAIN | _abi.NEED_DALPHA) if theta_var else _abi.LP_MAIN
    if not propto:
        flags |= _abi.LP_CONST
    r = density_eval(density, nn, th, flags)
    if r["rc"] == _abi.BAD_OPERAND:  # the kernel saw a bad element: the reference's checks name it
        if not n_check(nn):
            raise n_error
        if theta_check is not None and not np.any(np.isnan(th)):
            raise theta_error
        raise DomainError("%s: %s is nan, but must not be nan!" % (function, theta_name))
    if r["rc"] == _abi.LOG_ZERO:     # poisson_log_lpmf.hpp:51-62: LOG_ZERO, no partials
        return -np.inf, {"theta": np.zeros(th.size)}
    return r["lp"], {"theta": r["dtheta"]}


def bernoulli_logit_lpmf(n, theta, propto=False, theta_var=True):
    """stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:28-29; grads: theta."""
    f = "bernoulli_logit_lpmf"
    return _density(f, "bernoulli_logit", n, theta, propto, theta_var, lambda v: bool(np.all((v >= 0) & (v <= 1))),
                    DomainError("%s: n must be in the interval [0, 1]" % f), "Logit transformed probability parameter")


def poisson_log_lpmf(n, alpha, propto=False, alpha_var=True):
    """stan/math/prim/scal/prob/poisson_log_lpmf.hpp:22-23; grads: theta (the log rate alpha)."""
    f = "poisson_log_lpmf"
    return _density(f, "poisson_log", n, alpha, propto, alpha_var, lambda v: bool(np.all(v >= 0)),
                    DomainError("%s: Random variable must be >= 0!" % f), "Log rate parameter")


def poisson_lpmf(n, lam, propto=False, lambda_var=True):
    """stan/math/prim/scal/prob/poisson_lpmf.hpp:20-21; grads: theta (the rate lambda)."""
    f = "poisson_lpmf"
    return _density(f, "poisson", n, lam, propto, lambda_var, lambda v: bool(np.all(v >= 0)),
                    DomainError("%s: Random variable must be >= 0!" % f), "Rate parameter",
                    lambda t: bool(np.all(t >= 0)), DomainError("%s: Rate parameter must be >= 0!" % f))


def bernoulli_lpmf(n, theta, propto=False, theta_var=True):
    """stan/math/prim/scal/prob/bernoulli_lpmf.hpp:28-29; grads: theta (the chance of success)."""
    f = "bernoulli_lpmf"
    return _density(f, "bernoulli", n, theta, propto, theta_var, lambda v: bool(np.all((v >= 0) & (v <= 1))),
                    DomainError("%s: n must be in the interval [0, 1]" % f), "Probability parameter",
                    lambda t: bool(np.all(np.isfinite(t) & (t >= 0) & (t <= 1))),
                    DomainError("%s: Probability parameter must be finite and in the interval [0, 1]" % f))


class GlmCall:
    """One GLM evaluated many times at changing parameters through the host-level C-ABI call stan_cuda_glm_eval with
    HOST buffers: the argument block and the input / output arrays are built once, `eval(alpha, beta, sigma)` copies
    the parameters in and makes the call -- what a compiled caller (the C++ wrappers) does per evaluation, without
    the per-call array allocations of `glm_eval`.  Families bernoulli_logit / poisson_log / normal_id /
    neg_binomial_2_log with scalar intercept and scale."""

    def __init__(self, family, y, x, flags=None):
        fam = _abi.FAMILIES[family] if isinstance(family, str) else family
        x, row_major, ldx = _prep_x(x)
        n, k = x.shape
        y = _f64(y) if fam == _abi.NORMAL_ID else np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
        self.fam, self.n, self.k, self.x, self.y = fam, n, k, x, y
        self.alpha, self.beta, self.sigma = np.zeros(1), np.zeros(max(k, 1)), np.ones(1)
        self.lp, self.dbeta, self.dalpha, self.dsigma, self.aux = np.zeros(1), np.zeros(max(k, 1)), np.zeros(1), np.zeros(1), np.zeros(4)
        a = self.args = _abi.GlmArgs()
        a.family, a.flags, a.n, a.k = fam, (glm_flags(fam, False) if flags is None else flags), n, k
        a.x, a.ldx, a.x_row_major = x.ctypes.data, ldx, row_major
        a.y, a.y_is_double = y.ctypes.data, int(fam == _abi.NORMAL_ID)
        a.alpha, a.alpha_len, a.beta = self.alpha.ctypes.data, 1, self.beta.ctypes.data
        if fam in _abi.HAS_SCALE:
            a.sigma, a.sigma_len = self.sigma.ctypes.data, 1
        a.lp, a.dbeta, a.dalpha, a.dsigma = (self.lp.ctypes.data, self.dbeta.ctypes.data, self.dalpha.ctypes.data,
                                             self.dsigma.ctypes.data)
        a.aux = self.aux.ctypes.data
        self._ref = C.byref(a)
        self._fn = lib().stan_cuda_glm_eval

    def eval(self, alpha, beta, sigma=None):
        """Returns the C-ABI return code; results are in self.lp[0], self.dbeta[:k], self.dalpha[0], self.dsigma[0]."""
        self.alpha[0] = alpha
        self.beta[:self.k] = beta
        if sigma is not None:
            self.sigma[0] = sigma
        return check(self._fn(self._ref))


class GlmData:
    """Device-resident (x, y) of one GLM: the explicit form of the cache fill, for callers that
    own streams and device buffers themselves (bench.py via torch)."""

    def __init__(self, family, x, y):
        fam = _abi.FAMILIES[family] if isinstance(family, str) else family
        x, row_major, ldx = _prep_x(x)
        if fam == _abi.NORMAL_ID:
            y = _f64(y)
        else:
            y = np.ascontiguousarray(np.asarray(y, dtype=np.int32))
        n, k = x.shape
        h = C.c_void_p()
        check(lib().stan_cuda_data_create(fam, n, k, x.ctypes.data, ldx, row_major, y.ctypes.data,
                                          int(fam == _abi.NORMAL_ID), C.byref(h)))
        self.h, self.family, self.n, self.k = h, fam, n, k
        self.kpad = lib().stan_cuda_data_kpad(h)
        self.params_len = lib().stan_cuda_data_params_len(h)
        self.result_len = lib().stan_cuda_data_result_len(h)
        self.algorithmic_bytes = lib().stan_cuda_data_algorithmic_bytes(h)
        self.launches_per_eval = lib().stan_cuda_data_launches_per_eval(h)

    def launch(self, flags, d_params, d_result, stream, d_alpha_vec=None, d_sigma_vec=None,
               d_row_dalpha=None, d_row_dsigma=None):
        """Enqueue one fused evaluation; all pointers are device addresses (ints)."""
        check(lib().stan_cuda_glm_launch(self.h, flags, d_params, d_alpha_vec, d_sigma_vec, d_result,
                                         d_row_dalpha, d_row_dsigma, stream))

    def launch_ordered(self, flags, d_params, ncuts, d_result, stream):
        """ordered_logistic: d_params = [beta(kpad), 0, 0, cuts], d_result = [4 + kpad + ncuts]."""
        check(lib().stan_cuda_glm_launch_ordered(self.h, flags, d_params, ncuts, d_result, stream))

    def launch_categorical(self, flags, d_params, n_classes, d_result, stream):
        """categorical_logit: d_params = [beta: C rows of kpad, alpha(C)], d_result = [4 + C + C * kpad]."""
        check(lib().stan_cuda_glm_launch_categorical(self.h, flags, d_params, n_classes, d_result, stream))

    def close(self):
        if self.h:
            lib().stan_cuda_data_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def comm_init_from_torch(dist):
    """Build the library's NCCL communicator over an initialised torch.distributed group: rank 0
    creates the unique id, torch carries the 128 bytes to the other ranks (plumbing only)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if world == 1:
        return
    buf = C.create_string_buffer(128)
    if rank == 0:
        check(lib().stan_cuda_comm_unique_id(buf))
    box = [bytes(buf.raw)]
    dist.broadcast_object_list(box, src=0)
    check(lib().stan_cuda_comm_init_rank(world, rank, box[0]))
