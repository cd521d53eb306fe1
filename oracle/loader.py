"""oracle/loader.py -- TEST INFRASTRUCTURE, not product code.

ctypes front-ends of the two CPU checkers:

* ``oracle_*``  -> oracle/liboracle.so    (plain-C restatement, oracle/glm_oracle.c)
* ``ref_*``     -> oracle/_ref/libstanref.so (the compiled, unmodified reference behind a C shim,
                   oracle/ref_shim.cpp; present only if it was built where /root/reference exists)

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from math_b200 import _abi

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libstanref.so")

_oracle = None
_ref = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)


def oracle_lib():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
                os.path.join(HERE, "glm_oracle.c")):
            build_oracle()
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_glm_eval.restype = C.c_int
        lib.oracle_glm_eval.argtypes = [C.POINTER(_abi.GlmArgs)]
        lib.oracle_normal_lpdf_eval.restype = C.c_int
        lib.oracle_normal_lpdf_eval.argtypes = [C.POINTER(_abi.NormalArgs)]
        lib.oracle_locscale_lpdf_eval.restype = C.c_int
        lib.oracle_locscale_lpdf_eval.argtypes = [C.c_int, C.POINTER(_abi.NormalArgs)]
        lib.oracle_student_t_lpdf_eval.restype = C.c_int
        lib.oracle_student_t_lpdf_eval.argtypes = [C.POINTER(_abi.StudentTArgs)]
        i64, vp_ = C.c_int64, C.c_void_p
        lib.oracle_cholesky_decompose.restype = C.c_int
        lib.oracle_cholesky_decompose.argtypes = [i64, vp_, i64, vp_, i64]
        lib.oracle_cholesky_adjoint.restype = C.c_int
        lib.oracle_cholesky_adjoint.argtypes = [i64, vp_, i64, vp_, i64, vp_, i64]
        lib.oracle_tri_solve.restype = C.c_int
        lib.oracle_tri_solve.argtypes = [C.c_int, C.c_int, i64, i64, vp_, i64, vp_, i64, vp_, i64]
        lib.oracle_gp_exp_quad_cov.restype = C.c_int
        lib.oracle_gp_exp_quad_cov.argtypes = [i64, i64, i64, vp_, vp_, C.c_double, C.c_double, vp_, i64]
        lib.oracle_multiply.restype = C.c_int
        lib.oracle_multiply.argtypes = [C.POINTER(_abi.MultiplyArgs)]
        lib.oracle_density_eval.restype = C.c_int
        lib.oracle_density_eval.argtypes = [C.POINTER(_abi.DensityArgs)]
        _oracle = lib
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        dp, vp, i, l = C.c_void_p, C.c_void_p, C.c_int, C.c_long
        lib.stanref_glm.restype = i
        lib.stanref_glm.argtypes = [i, i, l, l, dp, vp, dp, l, dp, dp, l, i, dp, dp, dp, dp, dp, dp,
                                    C.c_char_p, i]
        if hasattr(lib, "stanref_ordered_logistic_glm"):
            lib.stanref_ordered_logistic_glm.restype = i
            lib.stanref_ordered_logistic_glm.argtypes = [i, l, l, dp, vp, dp, dp, l, dp, dp, dp, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_categorical_logit_glm"):
            lib.stanref_categorical_logit_glm.restype = i
            lib.stanref_categorical_logit_glm.argtypes = [i, l, l, l, dp, vp, dp, dp, dp, dp, dp, dp, C.c_char_p, i]
        lib.stanref_glm_value.restype = i
        lib.stanref_glm_value.argtypes = [i, i, l, l, dp, vp, dp, l, dp, dp, l, dp, C.c_char_p, i]
        lib.stanref_normal_lpdf.restype = i
        lib.stanref_normal_lpdf.argtypes = [i, l, dp, dp, l, dp, l, i, dp, dp, dp, dp, dp,
                                            C.c_char_p, i]
        lib.stanref_normal_lpdf_value.restype = i
        lib.stanref_normal_lpdf_value.argtypes = [i, l, dp, dp, l, dp, l, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_density"):
            lib.stanref_density.restype = i
            lib.stanref_density.argtypes = [i, i, l, vp, l, dp, l, i, dp, dp, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_locscale_lpdf"):
            lib.stanref_locscale_lpdf.restype = i
            lib.stanref_locscale_lpdf.argtypes = [i, i, i, dp, l, dp, l, dp, l, dp, dp, dp, dp, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_student_t_lpdf"):
            lib.stanref_student_t_lpdf.restype = i
            lib.stanref_student_t_lpdf.argtypes = [i, i, dp, l, dp, l, dp, l, dp, l, dp, dp, dp, dp, dp, dp,
                                                   C.c_char_p, i]
        if hasattr(lib, "stanref_multiply"):
            lib.stanref_multiply.restype = i
            lib.stanref_multiply.argtypes = [l, l, l, dp, dp, dp, dp, C.c_char_p, i]
            lib.stanref_multiply_rev.restype = i
            lib.stanref_multiply_rev.argtypes = [l, l, l, dp, dp, dp, i, dp, dp, dp, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_cholesky"):
            lib.stanref_cholesky.restype = i
            lib.stanref_cholesky.argtypes = [l, dp, dp, dp, C.c_char_p, i]
            lib.stanref_cholesky_rev.restype = i
            lib.stanref_cholesky_rev.argtypes = [l, dp, dp, dp, dp, dp, C.c_char_p, i]
        if hasattr(lib, "stanref_mdivide_left_tri"):
            lib.stanref_mdivide_left_tri.restype = i
            lib.stanref_mdivide_left_tri.argtypes = [i, l, l, dp, dp, dp, dp, C.c_char_p, i]
            lib.stanref_mdivide_left_tri_rev.restype = i
            lib.stanref_mdivide_left_tri_rev.argtypes = [i, i, l, l, dp, dp, dp, dp, dp, dp, dp, C.c_char_p, i]
        lib.stanref_init_threads.restype = i
        lib.stanref_init_threads.argtypes = [i]
        lib.stanref_shards_create.restype = C.c_void_p
        lib.stanref_shards_create.argtypes = [i, l, l, dp, vp, i]
        lib.stanref_shards_free.restype = None
        lib.stanref_shards_free.argtypes = [C.c_void_p]
        lib.stanref_shards_eval.restype = i
        lib.stanref_shards_eval.argtypes = [C.c_void_p, dp, dp, dp, dp, dp, dp, dp, dp, dp,
                                            C.c_char_p, i]
        lib.stanref_version.restype = C.c_char_p
        _ref = lib
    return _ref


# ------------------------------------------------------------------------------------------
def _f64(a):
    return np.ascontiguousarray(np.atleast_1d(np.asarray(a, dtype=np.float64)))


def _ptr(a):
    return None if a is None else a.ctypes.data


def prep_glm(family, x, y, alpha, beta, sigma=None):
    """Normalise GLM inputs: x -> column-major f64 (n,k); y -> int32 or f64; vectors -> f64."""
    fam = _abi.FAMILIES[family] if isinstance(family, str) else family
    x = np.asfortranarray(np.asarray(x, dtype=np.float64))
    if x.ndim != 2:
        raise ValueError("x must be 2-D")
    if fam == _abi.NORMAL_ID:
        y = _f64(y)
    else:
        y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    alpha, beta = _f64(alpha), _f64(beta)
    sigma = _f64(sigma) if sigma is not None else None
    return fam, x, y, alpha, beta, sigma


def glm_args(fam, x, y, alpha, beta, sigma, flags, need_dy=False, cuts=None):
    """Build a GlmArgs block + its output arrays (kept alive in the returned dict)."""
    n, k = x.shape
    out = {
        "lp": np.zeros(1), "dbeta": np.zeros(max(k, 1)), "dalpha": np.zeros(max(alpha.size, 1)),
        "dsigma": np.zeros(max(sigma.size if sigma is not None else 1, 1)),
        "dy": np.zeros(max(n, 1)) if need_dy else None, "aux": np.zeros(4),
    }
    a = _abi.GlmArgs()
    a.family, a.flags, a.n, a.k = fam, flags, n, k
    a.x, a.ldx, a.x_row_major = x.ctypes.data, max(n, 1), 0
    a.y, a.y_is_double = y.ctypes.data, int(fam == _abi.NORMAL_ID)
    a.alpha, a.alpha_len = alpha.ctypes.data, alpha.size
    a.beta = beta.ctypes.data
    a.sigma, a.sigma_len = (_ptr(sigma), sigma.size) if sigma is not None else (None, 0)
    a.lp, a.dbeta, a.dalpha = out["lp"].ctypes.data, out["dbeta"].ctypes.data, out["dalpha"].ctypes.data
    a.dsigma, a.dy, a.aux = out["dsigma"].ctypes.data, _ptr(out["dy"]), out["aux"].ctypes.data
    if cuts is not None:
        out["dcuts"] = np.zeros(max(cuts.size, 1))
        a.cuts, a.cuts_len, a.dcuts = cuts.ctypes.data, cuts.size, out["dcuts"].ctypes.data
    out["_keep"] = (x, y, alpha, beta, sigma, cuts)
    return a, out


def _finish_glm(rc, out, fam, n, k, alpha, sigma):
    res = {"rc": rc, "lp": float(out["lp"][0]), "dbeta": out["dbeta"][:k].copy(),
           "dalpha": out["dalpha"][:alpha.size].copy(), "aux": out["aux"].copy()}
    if fam in _abi.HAS_SCALE and sigma is not None:
        res["dsigma"] = out["dsigma"][:sigma.size].copy()
    if out["dy"] is not None:
        res["dy"] = out["dy"][:n].copy()
    return res


def flags_for(fam, propto=False, need_dy=False):
    """lp summands the reference keeps when every parameter is a var (see ref_shim.cpp):
    bernoulli: everything is parameter-dependent; poisson<propto=true> drops BOTH the lgamma term
    and (include_summand<propto, double>) the main term, poisson_log_glm_lpmf.hpp:102-112;
    normal_id<propto=true> drops N*NEG_LOG_SQRT_TWO_PI, normal_id_glm_lpdf.hpp:160-172."""
    f = _abi.NEED_DBETA | _abi.NEED_DALPHA | _abi.NEED_DSIGMA | (_abi.NEED_DY if need_dy else 0)
    if fam == _abi.NEG_BINOMIAL_2_LOG:
        # with every parameter a var only the lgamma(y + 1) term depends on propto
        # (neg_binomial_2_log_glm_lpmf.hpp:130-153)
        return f | _abi.LP_SCALE | _abi.LP_MAIN | _abi.LP_LINEAR | (0 if propto else _abi.LP_CONST)
    if not propto:
        return f | _abi.LP_ALL
    if fam == _abi.BERNOULLI_LOGIT:
        return f | _abi.LP_MAIN
    if fam == _abi.POISSON_LOG:
        return f
    return f | _abi.LP_SCALE | _abi.LP_MAIN


def oracle_glm(family, x, y, alpha, beta, sigma=None, propto=False, need_dy=False, flags=None):
    fam, x, y, alpha, beta, sigma = prep_glm(family, x, y, alpha, beta, sigma)
    if flags is None:
        flags = flags_for(fam, propto, need_dy)
    a, out = glm_args(fam, x, y, alpha, beta, sigma, flags, need_dy)
    rc = oracle_lib().oracle_glm_eval(C.byref(a))
    return _finish_glm(rc, out, fam, x.shape[0], x.shape[1], alpha, sigma)


def ordered_flags(propto=False):
    """every summand depends on beta / cuts: propto changes nothing when they are vars
    (ordered_logistic_glm_lpmf.hpp:93-94)"""
    return _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS


def oracle_ordered_logistic_glm(x, y, beta, cuts, propto=False, flags=None):
    fam, x, y, alpha, beta, _ = prep_glm(_abi.ORDERED_LOGISTIC, x, y, 0.0, beta, None)
    cuts = _f64(cuts) if np.size(cuts) else np.zeros(0)
    a, out = glm_args(fam, x, y, alpha, beta, None, ordered_flags(propto) if flags is None else flags, cuts=cuts)
    rc = oracle_lib().oracle_glm_eval(C.byref(a))
    return {"rc": rc, "lp": float(out["lp"][0]), "dbeta": out["dbeta"][:x.shape[1]].copy(),
            "dcuts": out["dcuts"][:cuts.size].copy(), "aux": out["aux"].copy()}


def _prep_categorical(x, y, alpha, beta):
    x = np.asfortranarray(np.asarray(x, dtype=np.float64))
    y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    alpha = _f64(alpha)
    beta = np.asfortranarray(np.asarray(beta, dtype=np.float64).reshape(x.shape[1], -1))
    return x, y, alpha, beta


def oracle_categorical_logit_glm(x, y, alpha, beta, propto=False, flags=None):
    """beta: (k, C).  Returns lp, dalpha (C), dbeta (k, C)."""
    x, y, alpha, beta = _prep_categorical(x, y, alpha, beta)
    n, k = x.shape
    ncls = beta.shape[1]
    lp, dalpha, dbeta, aux = np.zeros(1), np.zeros(max(ncls, 1)), np.zeros((max(k, 1), max(ncls, 1)), order="F"), np.zeros(4)
    a = _abi.GlmArgs()
    a.family, a.n, a.k, a.n_classes = _abi.CATEGORICAL_LOGIT, n, k, ncls
    a.flags = (_abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA) if flags is None else flags
    a.x, a.ldx, a.x_row_major, a.y, a.y_is_double = x.ctypes.data, max(n, 1), 0, y.ctypes.data, 0
    a.alpha, a.alpha_len, a.beta = alpha.ctypes.data, alpha.size, beta.ctypes.data
    a.lp, a.dalpha, a.dbeta, a.aux = lp.ctypes.data, dalpha.ctypes.data, dbeta.ctypes.data, aux.ctypes.data
    rc = oracle_lib().oracle_glm_eval(C.byref(a))
    return {"rc": rc, "lp": float(lp[0]), "dalpha": dalpha[:ncls].copy(),
            "dbeta": np.asarray(dbeta).reshape(-1, order="F")[:k * ncls].reshape((k, ncls), order="F").copy()}


class RefError(Exception):
    def __init__(self, rc, msg):
        super().__init__(msg)
        self.rc = rc  # 1 std::domain_error, 2 std::invalid_argument, 3 other


def ref_glm(family, x, y, alpha, beta, sigma=None, propto=False, need_dy=False):
    """Value + reverse-mode gradient through the compiled reference (all parameters var)."""
    fam, x, y, alpha, beta, sigma = prep_glm(family, x, y, alpha, beta, sigma)
    n, k = x.shape
    lp, dbeta, dalpha = np.zeros(1), np.zeros(max(k, 1)), np.zeros(max(alpha.size, 1))
    dsigma = np.zeros(max(sigma.size, 1)) if sigma is not None else np.zeros(1)
    dy = np.zeros(max(n, 1))
    secs = np.zeros(1)
    err = C.create_string_buffer(512)
    sg = sigma if sigma is not None else np.ones(1)
    rc = ref_lib().stanref_glm(fam, int(propto), n, k, x.ctypes.data, y.ctypes.data,
                               alpha.ctypes.data, alpha.size, beta.ctypes.data, sg.ctypes.data,
                               sg.size, int(need_dy), lp.ctypes.data, dbeta.ctypes.data,
                               dalpha.ctypes.data, dsigma.ctypes.data, dy.ctypes.data,
                               secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    res = {"rc": 0, "lp": float(lp[0]), "dbeta": dbeta[:k].copy(), "dalpha": dalpha[:alpha.size].copy(),
           "seconds": float(secs[0])}
    if fam in _abi.HAS_SCALE:
        res["dsigma"] = dsigma[:sg.size].copy()
    if need_dy:
        res["dy"] = dy[:n].copy()
    return res


def ref_ordered_logistic_glm(x, y, beta, cuts, propto=False):
    fam, x, y, alpha, beta, _ = prep_glm(_abi.ORDERED_LOGISTIC, x, y, 0.0, beta, None)
    cuts = _f64(cuts) if np.size(cuts) else np.zeros(0)
    n, k = x.shape
    lp, dbeta, dcuts, secs = np.zeros(1), np.zeros(max(k, 1)), np.zeros(max(cuts.size, 1)), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_ordered_logistic_glm(int(propto), n, k, x.ctypes.data, y.ctypes.data, beta.ctypes.data,
                                                cuts.ctypes.data, cuts.size, lp.ctypes.data, dbeta.ctypes.data,
                                                dcuts.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"rc": 0, "lp": float(lp[0]), "dbeta": dbeta[:k].copy(), "dcuts": dcuts[:cuts.size].copy(),
            "seconds": float(secs[0])}


def ref_categorical_logit_glm(x, y, alpha, beta, propto=False):
    x, y, alpha, beta = _prep_categorical(x, y, alpha, beta)
    n, k = x.shape
    ncls = beta.shape[1]
    lp, dalpha, dbeta, secs = np.zeros(1), np.zeros(max(ncls, 1)), np.zeros(max(k * ncls, 1)), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_categorical_logit_glm(int(propto), n, k, ncls, x.ctypes.data, y.ctypes.data,
                                                 alpha.ctypes.data, beta.ctypes.data, lp.ctypes.data,
                                                 dalpha.ctypes.data, dbeta.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"rc": 0, "lp": float(lp[0]), "dalpha": dalpha[:ncls].copy(),
            "dbeta": dbeta[:k * ncls].reshape((k, ncls), order="F").copy(), "seconds": float(secs[0])}


def ref_glm_value(family, x, y, alpha, beta, sigma=None, propto=False):
    fam, x, y, alpha, beta, sigma = prep_glm(family, x, y, alpha, beta, sigma)
    n, k = x.shape
    lp = np.zeros(1)
    err = C.create_string_buffer(512)
    sg = sigma if sigma is not None else np.ones(1)
    rc = ref_lib().stanref_glm_value(fam, int(propto), n, k, x.ctypes.data, y.ctypes.data,
                                     alpha.ctypes.data, alpha.size, beta.ctypes.data,
                                     sg.ctypes.data, sg.size, lp.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return float(lp[0])


# ------------------------------------------------------------------------------------------
def normal_flags(propto=False, need_dy=False):
    f = _abi.NEED_DALPHA | _abi.NEED_DSIGMA | (_abi.NEED_DY if need_dy else 0)
    return f | (_abi.LP_SCALE | _abi.LP_MAIN if propto else _abi.LP_ALL)


def normal_args(y, mu, sigma, flags, y_is_data=True):
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    n = max(y.size, mu.size, sigma.size) if min(y.size, mu.size, sigma.size) > 0 else 0
    out = {"lp": np.zeros(1), "dmu": np.zeros(max(mu.size, 1)), "dsigma": np.zeros(max(sigma.size, 1)),
           "dy": np.zeros(max(y.size, 1)), "_keep": (y, mu, sigma)}
    a = _abi.NormalArgs()
    a.flags, a.y_is_data, a.n = flags, int(y_is_data), n
    a.y, a.y_len, a.mu, a.mu_len = y.ctypes.data, y.size, mu.ctypes.data, mu.size
    a.sigma, a.sigma_len = sigma.ctypes.data, sigma.size
    a.lp, a.dmu, a.dsigma, a.dy = (out["lp"].ctypes.data, out["dmu"].ctypes.data,
                                   out["dsigma"].ctypes.data, out["dy"].ctypes.data)
    return a, out, (y, mu, sigma)


def _finish_normal(rc, out, y, mu, sigma, need_dy):
    res = {"rc": rc, "lp": float(out["lp"][0]), "dmu": out["dmu"][:mu.size].copy(),
           "dsigma": out["dsigma"][:sigma.size].copy()}
    if need_dy:
        res["dy"] = out["dy"][:y.size].copy()
    return res


def oracle_normal_lpdf(y, mu, sigma, propto=False, need_dy=False, flags=None):
    if flags is None:
        flags = normal_flags(propto, need_dy)
    a, out, (y, mu, sigma) = normal_args(y, mu, sigma, flags)
    rc = oracle_lib().oracle_normal_lpdf_eval(C.byref(a))
    return _finish_normal(rc, out, y, mu, sigma, need_dy)


def ref_normal_lpdf(y, mu, sigma, propto=False, need_dy=False):
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    n = max(y.size, mu.size, sigma.size)
    if y.size == 1 and n > 1:
        y = np.full(n, y[0])
    lp, dmu, dsigma, dy, secs = np.zeros(1), np.zeros(mu.size), np.zeros(sigma.size), np.zeros(n), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_normal_lpdf(int(propto), n, y.ctypes.data, mu.ctypes.data, mu.size,
                                       sigma.ctypes.data, sigma.size, int(need_dy), lp.ctypes.data,
                                       dmu.ctypes.data, dsigma.ctypes.data, dy.ctypes.data,
                                       secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    res = {"rc": 0, "lp": float(lp[0]), "dmu": dmu, "dsigma": dsigma, "seconds": float(secs[0])}
    if need_dy:
        res["dy"] = dy
    return res


# ------------------------------------------------------------------------------------------
# normal_lpdf / cauchy_lpdf / lognormal_lpdf (y | mu, sigma).  var_mask bit 0: y, 1: mu, 2: sigma are autodiff variables.
def locscale_flags(density, propto=False, var_mask=0b110):
    """The summands the reference keeps under include_summand<propto, ...> (normal_lpdf.hpp:89-97, cauchy_lpdf.hpp:87-95,
    lognormal_lpdf.hpp:92-118) and the partials of the operands that are variables."""
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


def oracle_locscale_lpdf(density, y, mu, sigma, propto=False, var_mask=0b110, flags=None):
    d = _abi.LOCSCALES[density] if isinstance(density, str) else density
    if flags is None:
        flags = locscale_flags(d, propto, var_mask)
    a, out, (y, mu, sigma) = normal_args(y, mu, sigma, flags)
    rc = oracle_lib().oracle_locscale_lpdf_eval(d, C.byref(a))
    res = {"rc": rc, "lp": float(out["lp"][0])}
    for key, bit, op in (("dy", _abi.NEED_DY, y), ("dmu", _abi.NEED_DALPHA, mu), ("dsigma", _abi.NEED_DSIGMA, sigma)):
        if flags & bit:
            res[key] = out[key][:op.size].copy()
    return res


def ref_locscale_lpdf(density, y, mu, sigma, propto=False, var_mask=0b110):
    """The compiled reference itself (oracle/ref_shim_locscale.cpp); raises RefError where it throws."""
    d = _abi.LOCSCALES[density] if isinstance(density, str) else density
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    lp, secs = np.zeros(1), np.zeros(1)
    dy, dmu, dsigma = np.zeros(max(y.size, 1)), np.zeros(max(mu.size, 1)), np.zeros(max(sigma.size, 1))
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_locscale_lpdf(d, int(propto), int(var_mask), y.ctypes.data, y.size, mu.ctypes.data, mu.size,
                                         sigma.ctypes.data, sigma.size, lp.ctypes.data, dy.ctypes.data, dmu.ctypes.data,
                                         dsigma.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    res = {"rc": 0, "lp": float(lp[0]), "seconds": float(secs[0])}
    for key, bit, g, op in (("dy", 1, dy, y), ("dmu", 2, dmu, mu), ("dsigma", 4, dsigma, sigma)):
        if var_mask & bit:
            res[key] = g[:op.size].copy()
    return res


# ------------------------------------------------------------------------------------------
# student_t_lpdf(y | nu, mu, sigma).  var_mask bit 0: y, 1: nu, 2: mu, 3: sigma are autodiff variables.
def student_t_flags(propto=False, var_mask=0b1110):
    """The summands the reference keeps (student_t_lpdf.hpp:158-169 under include_summand<propto, ...>) and the
    partials it builds (:171-198 under !is_constant_all<...>) for this mix of data / variables."""
    need = ((_abi.NEED_DY if var_mask & 1 else 0) | (_abi.NEED_DNU if var_mask & 2 else 0)
            | (_abi.NEED_DALPHA if var_mask & 4 else 0) | (_abi.NEED_DSIGMA if var_mask & 8 else 0))
    if not propto:
        return need | _abi.LP_CONST | _abi.LP_DOF | _abi.LP_SCALE | _abi.LP_MAIN
    return need | ((_abi.LP_DOF if var_mask & 2 else 0) | (_abi.LP_SCALE if var_mask & 8 else 0)
                   | (_abi.LP_MAIN if var_mask else 0))


def student_t_args(y, nu, mu, sigma, flags, y_is_data=True):
    ops = [_f64(v) for v in (y, nu, mu, sigma)]
    n = max(o.size for o in ops) if min(o.size for o in ops) > 0 else 0
    out = {"lp": np.zeros(1), "_keep": ops}
    for name, o in zip(("dy", "dnu", "dmu", "dsigma"), ops):
        out[name] = np.zeros(max(o.size, 1))
    a = _abi.StudentTArgs()
    a.flags, a.y_is_data, a.n = flags, int(y_is_data), n
    a.y, a.y_len, a.nu, a.nu_len = ops[0].ctypes.data, ops[0].size, ops[1].ctypes.data, ops[1].size
    a.mu, a.mu_len, a.sigma, a.sigma_len = ops[2].ctypes.data, ops[2].size, ops[3].ctypes.data, ops[3].size
    a.lp, a.dy, a.dnu, a.dmu, a.dsigma = (out[k].ctypes.data for k in ("lp", "dy", "dnu", "dmu", "dsigma"))
    return a, out, ops


def finish_student_t(rc, out, ops, flags):
    res = {"rc": rc, "lp": float(out["lp"][0])}
    for name, bit, o in zip(("dy", "dnu", "dmu", "dsigma"),
                            (_abi.NEED_DY, _abi.NEED_DNU, _abi.NEED_DALPHA, _abi.NEED_DSIGMA), ops):
        if flags & bit:
            res[name] = out[name][:o.size].copy()
    return res


def oracle_student_t_lpdf(y, nu, mu, sigma, propto=False, var_mask=0b1110, flags=None):
    if flags is None:
        flags = student_t_flags(propto, var_mask)
    a, out, ops = student_t_args(y, nu, mu, sigma, flags)
    rc = oracle_lib().oracle_student_t_lpdf_eval(C.byref(a))
    return finish_student_t(rc, out, ops, flags)


def ref_student_t_lpdf(y, nu, mu, sigma, propto=False, var_mask=0b1110):
    """The compiled reference itself (oracle/ref_shim_student_t.cpp); raises RefError where it throws."""
    ops = [_f64(v) for v in (y, nu, mu, sigma)]
    lp, secs = np.zeros(1), np.zeros(1)
    grads = [np.zeros(max(o.size, 1)) for o in ops]
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_student_t_lpdf(int(propto), int(var_mask), ops[0].ctypes.data, ops[0].size,
                                          ops[1].ctypes.data, ops[1].size, ops[2].ctypes.data, ops[2].size,
                                          ops[3].ctypes.data, ops[3].size, lp.ctypes.data, grads[0].ctypes.data,
                                          grads[1].ctypes.data, grads[2].ctypes.data, grads[3].ctypes.data,
                                          secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    res = {"rc": 0, "lp": float(lp[0]), "seconds": float(secs[0])}
    for bit, name, g, o in zip((1, 2, 4, 8), ("dy", "dnu", "dmu", "dsigma"), grads, ops):
        if var_mask & bit:
            res[name] = g[:o.size]
    return res


def ref_normal_lpdf_value(y, mu, sigma, propto=False):
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    lp = np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_normal_lpdf_value(int(propto), y.size, y.ctypes.data, mu.ctypes.data,
                                             mu.size, sigma.ctypes.data, sigma.size,
                                             lp.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return float(lp[0])


# ------------------------------------------------------------------------------------------
class RefShards:
    """Row-sharded GLM (or J-shard hierarchical GLM) evaluated through the reference's map_rect
    with TBB threads: the north star's CPU baseline (Eigen + TBB map_rect_concurrent)."""

    def __init__(self, family, x, y, n_shards, n_threads=-1):
        fam, x, y, _, _, _ = prep_glm(family, x, y, [0.0], np.zeros(np.asarray(x).shape[1]))
        n, k = x.shape
        if n % n_shards:
            raise ValueError("map_rect needs equal-sized jobs: n %% n_shards != 0")
        self.fam, self.n, self.k, self.n_shards = fam, n, k, n_shards
        self.threads = ref_lib().stanref_init_threads(n_threads)
        self.h = ref_lib().stanref_shards_create(fam, n, k, x.ctypes.data, y.ctypes.data, n_shards)

    def eval(self, beta, alpha=None, job_alpha=None, sigma=None):
        beta = _f64(beta)
        alpha = _f64(alpha) if alpha is not None else None
        job_alpha = _f64(job_alpha) if job_alpha is not None else None
        sigma = _f64(sigma) if sigma is not None else np.ones(1)
        lp, dbeta = np.zeros(1), np.zeros(self.k)
        dalpha = np.zeros(self.n_shards if job_alpha is not None else 1)
        dsigma, secs = np.zeros(1), np.zeros(1)
        err = C.create_string_buffer(512)
        rc = ref_lib().stanref_shards_eval(self.h, beta.ctypes.data, _ptr(alpha), _ptr(job_alpha),
                                           sigma.ctypes.data, lp.ctypes.data, dbeta.ctypes.data,
                                           dalpha.ctypes.data, dsigma.ctypes.data, secs.ctypes.data,
                                           err, 512)
        if rc != 0:
            raise RefError(rc, err.value.decode())
        return {"lp": float(lp[0]), "dbeta": dbeta, "dalpha": dalpha, "dsigma": dsigma,
                "seconds": float(secs[0])}

    def close(self):
        if self.h:
            ref_lib().stanref_shards_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------
# vectorised elementwise densities (bernoulli_logit_lpmf, poisson_log_lpmf)
def density_flags(propto=False, theta_var=True):
    """include_summand<propto> keeps the data-only term (poisson: -sum lgamma(n + 1)); the parameter term
    stays whenever theta is a var (include_summand<propto, T>)."""
    f = (_abi.LP_MAIN | _abi.NEED_DALPHA) if theta_var else (0 if propto else _abi.LP_MAIN)
    return f | (0 if propto else _abi.LP_CONST)


def density_args(density, y, theta, flags):
    d = _abi.DENSITIES[density] if isinstance(density, str) else density
    y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    theta = _f64(theta)
    n = max(y.size, theta.size) if min(y.size, theta.size) > 0 else 0
    out = {"lp": np.zeros(1), "dtheta": np.zeros(max(theta.size, 1)), "_keep": (y, theta)}
    a = _abi.DensityArgs()
    a.density, a.flags, a.n = d, flags, n
    a.y, a.y_len, a.theta, a.theta_len = y.ctypes.data, y.size, theta.ctypes.data, theta.size
    a.lp, a.dtheta = out["lp"].ctypes.data, out["dtheta"].ctypes.data
    return a, out, (y, theta)


def oracle_density(density, y, theta, propto=False, flags=None):
    if flags is None:
        flags = density_flags(propto)
    a, out, (y, theta) = density_args(density, y, theta, flags)
    rc = oracle_lib().oracle_density_eval(C.byref(a))
    return {"rc": rc, "lp": float(out["lp"][0]), "dtheta": out["dtheta"][:theta.size].copy()}


def ref_density(density, y, theta, propto=False, theta_var=True):
    """The compiled reference; raises RefError where it throws."""
    d = _abi.DENSITIES[density] if isinstance(density, str) else density
    y = np.ascontiguousarray(np.atleast_1d(np.asarray(y, dtype=np.int32)))
    theta = _f64(theta)
    n = max(y.size, theta.size)
    lp, dth, secs = np.zeros(1), np.zeros(theta.size), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_density(d, int(propto), n, y.ctypes.data, y.size, theta.ctypes.data, theta.size,
                                   int(theta_var), lp.ctypes.data, dth.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"lp": float(lp[0]), "dtheta": dth, "seconds": float(secs[0])}


# ------------------------------------------------------------------------------------------
# dense product
def _fmat(a):
    a = np.asarray(a, dtype=np.float64)
    return np.asfortranarray(a.reshape(a.shape[0], -1) if a.ndim == 2 else a.reshape(-1, 1))


def oracle_multiply(a, b, trans_a=False, trans_b=False):
    """op(a) @ op(b) by the plain-C triple loop (oracle/glm_oracle.c: oracle_multiply)."""
    a, b = _fmat(a), _fmat(b)
    m, ka = (a.shape[1], a.shape[0]) if trans_a else a.shape
    kb, n = (b.shape[1], b.shape[0]) if trans_b else b.shape
    assert ka == kb
    c = np.zeros((m, n), order="F")
    g = _abi.MultiplyArgs()
    g.m, g.n, g.k = m, n, ka
    g.a, g.lda, g.trans_a = a.ctypes.data, max(a.shape[0], 1), int(trans_a)
    g.b, g.ldb, g.trans_b = b.ctypes.data, max(b.shape[0], 1), int(trans_b)
    g.c, g.ldc = c.ctypes.data, max(m, 1)
    rc = oracle_lib().oracle_multiply(C.byref(g))
    assert rc == 0
    return c


def ref_multiply(a, b):
    """stan::math::multiply(A, B) of the compiled reference (double matrices: Eigen's product)."""
    a, b = _fmat(a), _fmat(b)
    (m, k), n = a.shape, b.shape[1]
    c, secs = np.zeros((m, n), order="F"), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_multiply(m, n, k, a.ctypes.data, b.ctypes.data, c.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return c, float(secs[0])


def ref_multiply_rev(a, b, w, var_mask=3):
    """Reverse mode through the compiled reference's multiply_mat_vari: C, dA = W B', dB = A' W for
    f = sum(W * C); var_mask bit 0: A is a matrix of var, bit 1: B."""
    a, b, w = _fmat(a), _fmat(b), _fmat(w)
    (m, k), n = a.shape, b.shape[1]
    c, da, db, secs = np.zeros((m, n), order="F"), np.zeros((m, k), order="F"), np.zeros((k, n), order="F"), np.zeros(2)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_multiply_rev(m, n, k, a.ctypes.data, b.ctypes.data, w.ctypes.data, int(var_mask), c.ctypes.data,
                                        da.ctypes.data, db.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"c": c, "da": da if var_mask & 1 else None, "db": db if var_mask & 2 else None,
            "forward_s": float(secs[0]), "grad_s": float(secs[1])}


# ------------------------------------------------------------------------------------------
# Cholesky factor and its reverse-mode adjoint
def oracle_cholesky(a):
    """(rc, L): lower factor by the plain-C column algorithm; rc = NOT_POS_DEF where the reference throws."""
    a = _fmat(a)
    n = a.shape[0]
    l = np.zeros((n, n), order="F")
    rc = oracle_lib().oracle_cholesky_decompose(n, a.ctypes.data, max(n, 1), l.ctypes.data, max(n, 1))
    return rc, l


def oracle_cholesky_adjoint(l, l_adj):
    """adjoint of the lower triangle of A given L and the adjoint of its lower triangle (unblocked loops)."""
    l, l_adj = _fmat(l), _fmat(l_adj)
    n = l.shape[0]
    out = np.zeros((n, n), order="F")
    rc = oracle_lib().oracle_cholesky_adjoint(n, l.ctypes.data, max(n, 1), l_adj.ctypes.data, max(n, 1), out.ctypes.data, max(n, 1))
    assert rc == 0
    return out


def ref_cholesky(a):
    a = _fmat(a)
    n = a.shape[0]
    l, secs = np.zeros((n, n), order="F"), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_cholesky(n, a.ctypes.data, l.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return l, float(secs[0])


def ref_cholesky_rev(a, w):
    """L = cholesky_decompose(A_v), f = sum(tril(W) * L): returns L, dA (lower triangle) and the two timings."""
    a, w = _fmat(a), _fmat(w)
    n = a.shape[0]
    l, a_adj, secs = np.zeros((n, n), order="F"), np.zeros((n, n), order="F"), np.zeros(2)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_cholesky_rev(n, a.ctypes.data, w.ctypes.data, l.ctypes.data, a_adj.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"l": l, "a_adj": a_adj, "forward_s": float(secs[0]), "grad_s": float(secs[1])}


# ------------------------------------------------------------------------------------------
# triangular solve (mdivide_left_tri)
def oracle_tri_solve(a, b, lower=True, trans=False):
    """op(A)^-1 b by plain-C substitution; b a matrix or a vector."""
    a, b = _fmat(a), _fmat(b)
    n, m = a.shape[0], b.shape[1]
    x = np.zeros((n, m), order="F")
    rc = oracle_lib().oracle_tri_solve(int(lower), int(trans), n, m, a.ctypes.data, max(n, 1), b.ctypes.data, max(n, 1),
                                       x.ctypes.data, max(n, 1))
    assert rc == 0
    return x


def ref_mdivide_left_tri(a, b, lower=True):
    a, b = _fmat(a), _fmat(b)
    n, m = a.shape[0], b.shape[1]
    x, secs = np.zeros((n, m), order="F"), np.zeros(1)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_mdivide_left_tri(int(lower), n, m, a.ctypes.data, b.ctypes.data, x.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return x, float(secs[0])


def ref_mdivide_left_tri_rev(a, b, w, lower=True, var_mask=3):
    a, b, w = _fmat(a), _fmat(b), _fmat(w)
    n, m = a.shape[0], b.shape[1]
    x, da, db, secs = np.zeros((n, m), order="F"), np.zeros((n, n), order="F"), np.zeros((n, m), order="F"), np.zeros(2)
    err = C.create_string_buffer(512)
    rc = ref_lib().stanref_mdivide_left_tri_rev(int(lower), int(var_mask), n, m, a.ctypes.data, b.ctypes.data, w.ctypes.data,
                                                x.ctypes.data, da.ctypes.data, db.ctypes.data, secs.ctypes.data, err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return {"x": x, "da": da if var_mask & 1 else None, "db": db if var_mask & 2 else None,
            "forward_s": float(secs[0]), "grad_s": float(secs[1])}


# ------------------------------------------------------------------------------------------
# gp_exp_quad_cov
def oracle_gp_exp_quad_cov(x1, sigma, length_scale, x2=None):
    """internal::gp_exp_quad_cov restated (oracle/glm_oracle.c); argument meaning as math_b200.dense.gp_exp_quad_cov"""
    def pack(x):
        x = np.asarray(x, dtype=np.float64)
        pts = x.reshape(-1, 1) if x.ndim == 1 else x
        if np.ndim(length_scale) > 0:
            pts = pts / np.asarray(length_scale, dtype=np.float64)[None, :]
        return np.ascontiguousarray(pts.T)
    p1 = pack(x1)
    p2 = None if x2 is None else pack(x2)
    dim, n1 = p1.shape
    n2 = n1 if p2 is None else p2.shape[1]
    factor = -0.5 if np.ndim(length_scale) > 0 else -0.5 / (float(length_scale) * float(length_scale))
    cov = np.zeros((n1, n2), order="F")
    rc = oracle_lib().oracle_gp_exp_quad_cov(n1, n2, dim, p1.ctypes.data, None if p2 is None else p2.ctypes.data,
                                             float(sigma) * float(sigma), factor, cov.ctypes.data, max(n1, 1))
    assert rc == 0
    return cov
