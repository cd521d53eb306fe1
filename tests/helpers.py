"""Shared helpers of the test-suite: golden-case access, the parity metric, synthetic inputs."""
import numpy as np

from math_b200 import _abi

RTOL = 1e-10  # north star: 1e-10 relative on the log-density and every gradient component


def rel_err(got, ref):
    """SURVEY 8(d) parity gate: |d| / max(|ref|, 1e-6 * ||ref||_inf), worst component."""
    got, ref = np.atleast_1d(np.asarray(got, float)), np.atleast_1d(np.asarray(ref, float))
    assert got.shape == ref.shape, (got.shape, ref.shape)
    if ref.size == 0:
        return 0.0
    scale = np.maximum(np.abs(ref), 1e-6 * np.max(np.abs(ref)))
    scale = np.where(scale == 0, 1.0, scale)
    return float(np.max(np.abs(got - ref) / scale))


def assert_close(got, ref, what, rtol=RTOL):
    e = rel_err(got, ref)
    assert e <= rtol, "%s: rel err %.3e > %.1e" % (what, e, rtol)


def glm_case(golden, name):
    p = name + "/"
    fam = int(golden[p + "family"])
    c = {"family": fam, "propto": bool(golden[p + "propto"]), "need_dy": bool(golden[p + "need_dy"]),
         "x": golden["x/" + str(golden[p + "xkey"])], "y": golden[p + "y"], "alpha": golden[p + "alpha"],
         "beta": golden[p + "beta"], "sigma": golden[p + "sigma"] if fam in _abi.HAS_SCALE else None,
         "lp": float(golden[p + "lp"]), "dbeta": golden[p + "dbeta"], "dalpha": golden[p + "dalpha"]}
    if fam in _abi.HAS_SCALE:
        c["dsigma"] = golden[p + "dsigma"]
    if c["need_dy"]:
        c["dy"] = golden[p + "dy"]
    return c


def normal_case(golden, name):
    p = name + "/"
    c = {k: golden[p + k] for k in ("y", "mu", "sigma", "dmu", "dsigma")}
    c["propto"], c["need_dy"], c["lp"] = bool(golden[p + "propto"]), bool(golden[p + "need_dy"]), float(golden[p + "lp"])
    if c["need_dy"]:
        c["dy"] = golden[p + "dy"]
    if p + "known_answer" in golden.files:
        c["known_answer"] = float(golden[p + "known_answer"])
    return c


def student_t_case(golden_st, name):
    p = "st/%s/" % name
    c = {k: golden_st[p + k] for k in ("y", "nu", "mu", "sigma")}
    c["propto"], c["var_mask"], c["lp"] = bool(golden_st[p + "propto"]), int(golden_st[p + "var_mask"]), float(golden_st[p + "lp"])
    for k in ("dy", "dnu", "dmu", "dsigma"):
        if p + k in golden_st.files:
            c[k] = golden_st[p + k]
    if p + "known_answer" in golden_st.files:
        c["known_answer"] = float(golden_st[p + "known_answer"])
    return c


def check_student_t_result(r, c, what, rtol=RTOL):
    """d/dy, d/dmu, d/dsigma: the north-star gate; lp: the same plus the lgamma floor below.  d/dnu (student_t_lpdf.hpp:177-182) is the sum of
    0.5 * digamma(nu/2 + 1/2) - 0.5 * digamma(nu/2) - 0.5 / nu (terms of size |digamma(nu/2)| that cancel to O(1/nu^2))
    and the data terms: every implementation -- the reference's boost::math::digamma included -- carries an ABSOLUTE
    error of a few ulp of digamma(nu/2) per element there, which no relative gate on the cancelled result can hold.
    The gate for d/dnu is therefore |got - ref| <= rtol * scale + 32 ulp(max(1, |digamma(nu/2)|)) per element summed."""
    n = max(np.size(c[k]) for k in ("y", "nu", "mu", "sigma"))
    nu_all = np.broadcast_to(np.asarray(c["nu"], float), (n,)) if np.size(c["nu"]) in (1, n) else np.asarray(c["nu"], float)
    # lp holds lgamma(nu/2 + 1/2) - lgamma(nu/2) per element (:161-163): two numbers of size |lgamma(nu/2)| whose
    # difference is O(log nu) -- any two lgamma implementations (glibc's on the host, CUDA's on the device) differ by
    # a few ulp OF lgamma(nu/2) there.  Invisible for nu < 100, 1e-7 per element at nu = 1e8.
    lgamma_size = np.maximum(1.0, nu_all / 2 * np.maximum(np.log(nu_all / 2), 1.0))
    lp_floor = float(np.sum(8 * np.finfo(float).eps * lgamma_size))
    assert abs(r["lp"] - c["lp"]) <= rtol * abs(c["lp"]) + lp_floor, "%s lp: got %r ref %r" % (what, r["lp"], c["lp"])
    for k in ("dy", "dmu", "dsigma"):
        if k in c:
            assert_close(r[k], c[k], what + " " + k, rtol)
    if "dnu" in c:
        got, ref = np.asarray(r["dnu"], float), np.asarray(c["dnu"], float)
        assert got.shape == ref.shape
        nu = np.asarray(c["nu"], float)
        digamma_size = np.maximum(1.0, np.maximum(np.abs(np.log(nu / 2)), 2.0 / nu))   # |digamma(x)| ~ max(|log x|, 1/x)
        floor = 32 * np.finfo(float).eps * digamma_size * (n if nu.size == 1 else 1)
        scale = np.maximum(np.abs(ref), 1e-6 * np.max(np.abs(ref)))
        excess = np.abs(got - ref) - (rtol * scale + floor)
        assert np.all(excess <= 0), "%s dnu: |got - ref| exceeds rtol * scale + digamma ulp floor by %.3e" % (what, float(np.max(excess)))


def check_glm_result(r, c, what, rtol=RTOL):
    assert_close(r["lp"], c["lp"], what + " lp", rtol)
    assert_close(r["dbeta"], c["dbeta"], what + " dbeta", rtol)
    assert_close(r["dalpha"], c["dalpha"], what + " dalpha", rtol)
    if "dsigma" in c:
        assert_close(r["dsigma"], c["dsigma"], what + " dsigma", rtol)
    if "dy" in c:
        assert_close(r["dy"], c["dy"], what + " dy", rtol)


def synth_glm(family, n, k, seed, dtype_y=None):
    """Synthetic inputs of SURVEY 8(d): X ~ N(0,1); y drawn from the model at (alpha_true,
    beta_true); evaluation point perturbed away from the truth so no gradient cancels."""
    fam = _abi.FAMILIES[family] if isinstance(family, str) else family
    rng = np.random.default_rng(seed)
    x = np.asfortranarray(rng.standard_normal((n, k)))
    sd = 0.1 if k <= 32 else 0.05
    beta_true = rng.standard_normal(k) * sd
    eta = 0.2 + x @ beta_true
    if fam == _abi.BERNOULLI_LOGIT:
        y = (rng.random(n) < 1.0 / (1.0 + np.exp(-eta))).astype(np.int32)
    elif fam == _abi.POISSON_LOG:
        y = rng.poisson(np.exp(eta)).astype(np.int32)
    elif fam == _abi.NEG_BINOMIAL_2_LOG:
        y = rng.negative_binomial(3.0, 3.0 / (3.0 + np.exp(eta))).astype(np.int32)
    else:
        y = eta + 0.7 * rng.standard_normal(n)
    beta = beta_true + 0.02 * rng.standard_normal(k)
    return {"family": fam, "x": x, "y": y, "alpha": np.array([0.3]), "beta": beta,
            "sigma": np.array([1.0]) if fam == _abi.NORMAL_ID else (np.array([2.5]) if fam == _abi.NEG_BINOMIAL_2_LOG else None)}


def ord_case(golden, name):
    p = "ord/%s/" % name
    c = {k: golden[p + k] for k in ("x", "y", "beta", "cuts", "dbeta", "dcuts")}
    c["lp"], c["propto"] = float(golden[p + "lp"]), bool(golden[p + "propto"])
    return c


def check_ord_result(r, c, what, rtol=RTOL):
    assert_close(r["lp"], c["lp"], what + " lp", rtol)
    assert_close(r["dbeta"], c["dbeta"], what + " dbeta", rtol)
    assert_close(r["dcuts"], c["dcuts"], what + " dcuts", rtol)


def synth_ordered(n, k, n_classes, seed):
    """Ordinal outcomes drawn from the model at (beta_true, cuts); evaluated at a perturbed beta."""
    rng = np.random.default_rng(seed)
    x = np.asfortranarray(rng.standard_normal((n, k)))
    beta_true = rng.standard_normal(k) * (1.0 / np.sqrt(k))
    cuts = np.sort(rng.standard_normal(n_classes - 1) * 1.2)
    if cuts.size > 1:
        cuts += np.arange(cuts.size) * 0.05   # strictly ascending
    lat = x @ beta_true + rng.logistic(size=n)
    y = (1 + np.searchsorted(cuts, lat)).astype(np.int32)
    return {"x": x, "y": y, "beta": beta_true + 0.05 * rng.standard_normal(k), "cuts": cuts}


def cat_case(golden, name):
    p = "cat/%s/" % name
    c = {k: golden[p + k] for k in ("x", "y", "alpha", "beta", "dalpha", "dbeta")}
    c["lp"], c["propto"] = float(golden[p + "lp"]), bool(golden[p + "propto"])
    return c


def check_cat_result(r, c, what, rtol=RTOL):
    assert_close(r["lp"], c["lp"], what + " lp", rtol)
    assert_close(r["dalpha"], c["dalpha"], what + " dalpha", rtol)
    assert_close(np.asarray(r["dbeta"]).ravel(order="F"), np.asarray(c["dbeta"]).ravel(order="F"), what + " dbeta", rtol)


def synth_categorical(n, k, n_classes, seed):
    rng = np.random.default_rng(seed)
    x = np.asfortranarray(rng.standard_normal((n, k)))
    beta_true = rng.standard_normal((k, n_classes)) * (1.0 / np.sqrt(k))
    alpha = rng.standard_normal(n_classes) * 0.5
    z = x @ beta_true + alpha + rng.gumbel(size=(n, n_classes))
    y = (1 + np.argmax(z, axis=1)).astype(np.int32)
    return {"x": x, "y": y, "alpha": alpha + 0.1, "beta": beta_true + 0.05 * rng.standard_normal((k, n_classes))}


def dens_case(golden, name):
    p = "dens/%s/" % name
    c = {k: golden[p + k] for k in ("y", "theta", "dtheta")}
    c["density"], c["propto"], c["lp"] = int(golden[p + "density"]), bool(golden[p + "propto"]), float(golden[p + "lp"])
    if p + "known_answer" in golden.files:
        c["known_answer"] = float(golden[p + "known_answer"])
    return c


def check_dens_result(r, c, what, rtol=RTOL):
    assert_close(r["lp"], c["lp"], what + " lp", rtol)
    assert_close(r["dtheta"], c["dtheta"], what + " dtheta", rtol)   # the parity metric of SURVEY 8(d)
    # and element by element (the partials of a vector operand do not share a scale): relative to the element
    # itself, floored at 1e-6 of the terms it is a difference of (poisson: n - exp(alpha))
    got, ref = np.asarray(r["dtheta"], float), np.asarray(c["dtheta"], float)
    if ref.size > 1:
        theta, y = np.asarray(c["theta"], float), np.asarray(c["y"], float)
        with np.errstate(over="ignore"):
            scale = np.abs(y) + np.exp(theta) if c["density"] == _abi.DENSITY_POISSON_LOG else np.ones(ref.size)
        err = np.abs(got - ref) / np.maximum(np.maximum(np.abs(ref), 1e-6 * np.minimum(scale, 1e300)), 1e-300)
        assert not (err > rtol).any(), "%s dtheta: %d components off, worst %.3e" % (what, int((err > rtol).sum()), float(err.max()))
