/* oracle/glm_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C CPU restatement of the reference's algorithm for the hot path, one function per
 * reference function.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
 * may load it (as the checker); the product path (math_b200/, stan/math/cuda/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   (a) the reference's own known-answer vectors for normal_lpdf
 *       (test/prob/normal/normal_test.hpp:11-40),
 *   (b) the fixed fixtures of the reference's GLM unit tests
 *       (test/unit/math/rev/mat/prob/{bernoulli_logit,poisson_log,normal_id}_glm_*_test.cpp),
 *       whose expected values are stored in tests/golden/ as produced by the reference itself
 *       (oracle/_ref/libstanref.so, generator tests/golden/make_golden.py),
 *   (c) the compiled reference on random inputs whenever oracle/_ref/libstanref.so is present.
 *   (d) round 2: the reference-generated vectors of the widened rows -- tests/golden/student_t_golden.npz,
 *       density2_golden.npz (poisson_lpmf, bernoulli_lpmf), locscale_golden.npz (cauchy, lognormal, gamma, exponential,
 *       beta, weibull, inv_gamma), dense_golden.npz (multiply, cholesky_decompose and its reverse sweep, mdivide_left_tri,
 *       gp_exp_quad_cov) -- each with the literal known answers of the reference's own test fixtures (test/prob/...)
 *       stored beside the reference's outputs, generators committed next to them (tests/golden/make_*_golden.py);
 *       tests/test_oracle.py, test_student_t.py, test_locscale.py and test_dense.py run the comparisons.
 *
 * It takes the same argument structs as the product's C ABI (include/stan_cuda.h) so a test can
 * hand identical arguments to both.  Sums run sequentially in double, in row order; the
 * reference sums in Eigen's packet order -- both are within a few ulp * sqrt(N) of the exact
 * sum, far inside the 1e-10 parity bar.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

#include "../include/stan_cuda.h"

/* stan/math/prim/scal/fun/constants.hpp: NEG_LOG_SQRT_TWO_PI = -log(sqrt(2*pi)) */
static const double ORACLE_NEG_LOG_SQRT_TWO_PI = -0.91893853320467274178032973640561763986;

static double x_at(const stan_cuda_glm_args* a, int64_t i, int64_t j) {
  return a->x_row_major ? a->x[i * a->ldx + j] : a->x[j * a->ldx + i];
}

static double y_at(const stan_cuda_glm_args* a, int64_t i) {
  return a->y_is_double ? ((const double*)a->y)[i] : (double)((const int32_t*)a->y)[i];
}

/* eta_i = sum_j x_ij beta_j, accumulated column by column exactly like Eigen's col-major GEMV
 * (bernoulli_logit_glm_lpmf.hpp:92, poisson_log_glm_lpmf.hpp:91, normal_id_glm_lpdf.hpp:108) */
static double* linear_predictor(const stan_cuda_glm_args* a) {
  double* eta = (double*)calloc((size_t)(a->n > 0 ? a->n : 1), sizeof(double));
  if (!eta) return NULL;
  for (int64_t j = 0; j < a->k; ++j) {
    const double b = a->beta[j];
    for (int64_t i = 0; i < a->n; ++i) eta[i] += x_at(a, i, j) * b;
  }
  return eta;
}

/* dbeta = X^T d  (bernoulli :123, poisson :117, normal_id :125-127) */
static void xt_times(const stan_cuda_glm_args* a, const double* d, double* out) {
  for (int64_t j = 0; j < a->k; ++j) {
    double s = 0.0;
    for (int64_t i = 0; i < a->n; ++i) s += x_at(a, i, j) * d[i];
    out[j] = s;
  }
}

static void write_dalpha(const stan_cuda_glm_args* a, const double* d, double sum_d) {
  if (!(a->flags & STAN_CUDA_NEED_DALPHA) || !a->dalpha) return;
  if (a->alpha_len == 1) {
    a->dalpha[0] = sum_d;
  } else {
    for (int64_t i = 0; i < a->n; ++i) a->dalpha[i] = d[i];
  }
}

/* stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:80-135 */
static int bernoulli_logit_glm(const stan_cuda_glm_args* a, double* d) {
  static const double cutoff = 20.0; /* :98 */
  double logp = 0.0, sum_d = 0.0;
  double* eta = linear_predictor(a);
  if (!eta) return STAN_CUDA_ERR_ARG;
  for (int64_t i = 0; i < a->n; ++i) {
    const double sign = 2.0 * y_at(a, i) - 1.0;                              /* :90 */
    const double al = a->alpha_len == 1 ? a->alpha[0] : a->alpha[i];
    const double ytheta = sign * (eta[i] + al);                              /* :93-94 */
    const double e = exp(-ytheta);                                           /* :100 */
    logp += ytheta > cutoff ? -e : (ytheta < -cutoff ? ytheta : -log1p(e)); /* :101-104 */
    d[i] = ytheta > cutoff ? -e : (ytheta < -cutoff ? sign : sign * e / (e + 1.0)); /* :115-121 */
    sum_d += d[i];
  }
  free(eta);
  if (a->lp) *a->lp = (a->flags & STAN_CUDA_LP_MAIN) ? logp : 0.0;
  if (a->aux) { a->aux[0] = sum_d; a->aux[1] = 0; a->aux[2] = 0; a->aux[3] = 0; }
  write_dalpha(a, d, sum_d);
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;               /* :106 */
}

/* stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp:82-129 */
static int poisson_log_glm(const stan_cuda_glm_args* a, double* d) {
  double sum_d = 0.0, lg = 0.0, main_term = 0.0;
  double* theta = linear_predictor(a);
  if (!theta) return STAN_CUDA_ERR_ARG;
  for (int64_t i = 0; i < a->n; ++i) {
    const double y = y_at(a, i);
    const double th = theta[i] + (a->alpha_len == 1 ? a->alpha[0] : a->alpha[i]); /* :92 */
    const double e = exp(th);
    d[i] = y - e;                                                            /* :94-95 */
    sum_d += d[i];                                                           /* :96 */
    lg += lgamma(y + 1.0);                                                   /* :104 */
    main_term += y * th - e;                                                 /* :110-111 */
  }
  free(theta);
  double logp = 0.0;
  if (a->flags & STAN_CUDA_LP_CONST) logp -= lg;                             /* :102-108 */
  if (a->flags & STAN_CUDA_LP_MAIN) logp += main_term;                       /* :109-112 */
  if (a->lp) *a->lp = logp;
  if (a->aux) { a->aux[0] = sum_d; a->aux[1] = 0; a->aux[2] = 0; a->aux[3] = lg; }
  write_dalpha(a, d, sum_d);
  return isfinite(sum_d) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;               /* :97 */
}

/* stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp:89-172 */
static int normal_id_glm(const stan_cuda_glm_args* a, double* m) {
  double sum_m = 0.0, sum_s2 = 0.0, sum_log_sigma = 0.0;
  double* mu = linear_predictor(a);
  if (!mu) return STAN_CUDA_ERR_ARG;
  for (int64_t i = 0; i < a->n; ++i) {
    const double sg = a->sigma_len == 1 ? a->sigma[0] : a->sigma[i];
    const double inv_sigma = 1.0 / sg;                                       /* :100 */
    const double al = a->alpha_len == 1 ? a->alpha[0] : a->alpha[i];
    const double s = (y_at(a, i) - mu[i] - al) * inv_sigma;                  /* :108-111 */
    m[i] = inv_sigma * s;                                                    /* :117 */
    sum_m += m[i];
    sum_s2 += s * s;                                                         /* :138,:141 */
    if (a->sigma_len != 1) sum_log_sigma += log(sg);                         /* :164 */
    if ((a->flags & STAN_CUDA_NEED_DY) && a->dy) a->dy[i] = -m[i];            /* :118-120 */
    if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma && a->sigma_len != 1)
      a->dsigma[i] = (s * s - 1.0) * inv_sigma;                              /* :139 */
  }
  free(mu);
  if (a->sigma_len == 1) {
    sum_log_sigma = (double)a->n * log(a->sigma[0]);                         /* :166 */
    if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma)
      a->dsigma[0] = (sum_s2 - (double)a->n) * (1.0 / a->sigma[0]);          /* :142-143 */
  }
  double logp = 0.0;
  if (a->flags & STAN_CUDA_LP_CONST) logp += ORACLE_NEG_LOG_SQRT_TWO_PI * (double)a->n; /* :161 */
  if (a->flags & STAN_CUDA_LP_SCALE) logp -= sum_log_sigma;                  /* :163-168 */
  if (a->flags & STAN_CUDA_LP_MAIN) logp -= 0.5 * sum_s2;                    /* :170 */
  if (a->lp) *a->lp = logp;
  if (a->aux) { a->aux[0] = sum_m; a->aux[1] = sum_s2; a->aux[2] = sum_log_sigma; a->aux[3] = 0; }
  write_dalpha(a, m, sum_m);
  return isfinite(sum_s2) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;              /* :150 */
}

/* digamma: the reference calls boost::math::digamma (stan/math/prim/scal/fun/digamma.hpp:50-52,
 * Boost 1.69, absent as a compiled library here but header-only in /root/reference/lib).  Restated
 * with the textbook algorithm: recurrence psi(x) = psi(x + 1) - 1/x up to x >= 10, then the
 * asymptotic series ln x - 1/(2x) - sum B_2k / (2k x^2k) (7 terms: error < 1e-17 at x = 10).
 * Pinned against the compiled reference in tests/test_oracle.py. */
static double oracle_digamma(double x) {
  double r = 0.0;
  while (x < 10.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  const double f = 1.0 / (x * x);
  const double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0
                 + f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + log(x) - 0.5 / x + t;
}

/* stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp:108-197; phi arrives in the sigma slot */
static int neg_binomial_2_log_glm(const stan_cuda_glm_args* a, double* d) {
  double sum_d = 0.0, lg_y1 = 0.0, scale_term = 0.0, main_term = 0.0, lin_term = 0.0, lg_ypp = 0.0;
  double dphi_sum = 0.0, witness = 0.0;
  double* theta = linear_predictor(a);
  if (!theta) return STAN_CUDA_ERR_ARG;
  const int phi_vec = a->sigma_len != 1;
  for (int64_t i = 0; i < a->n; ++i) {
    const double y = y_at(a, i);
    const double phi = phi_vec ? a->sigma[i] : a->sigma[0];
    const double th = theta[i] + (a->alpha_len == 1 ? a->alpha[0] : a->alpha[i]);    /* :118-119 */
    const double log_phi = log(phi);                                                   /* :121 */
    const double lse = th > log_phi ? th + log1p(exp(log_phi - th))                    /* :122-125 */
                                    : log_phi + log1p(exp(th - log_phi));
    const double ypp = y + phi;                                                        /* :127 */
    lg_y1 += lgamma(y + 1.0);                                                          /* :131 */
    if (phi_vec) scale_term += (phi == 0 ? 0.0 : phi * log_phi) - lgamma(phi);         /* :137 */
    main_term += ypp * lse;                                                            /* :146 */
    lin_term += y * th;                                                                /* :149 */
    lg_ypp += lgamma(ypp);                                                             /* :152 */
    const double te = exp(th);                                                         /* :159 */
    d[i] = y - te * ypp / (te + phi);                                                  /* :162 */
    sum_d += d[i];
    const double dp = -ypp / (te + phi) + log_phi - lse + oracle_digamma(ypp) - oracle_digamma(phi);
    if (phi_vec) {
      if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma) a->dsigma[i] = 1.0 + dp;    /* :180-183 */
    } else {
      dphi_sum += dp;                                                                  /* :185-190 */
    }
    witness += th;
  }
  free(theta);
  if (!phi_vec) {
    const double phi = a->sigma[0];
    scale_term = (double)a->n * ((phi == 0 ? 0.0 : phi * log(phi)) - lgamma(phi));     /* :140-142 */
    if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma) a->dsigma[0] = (double)a->n + dphi_sum; /* :185 */
  }
  double logp = 0.0;
  if (a->flags & STAN_CUDA_LP_CONST) logp -= lg_y1;                                    /* :130-132 */
  if (a->flags & STAN_CUDA_LP_SCALE) logp += scale_term;                               /* :133-144 */
  if (a->flags & STAN_CUDA_LP_MAIN) logp -= main_term;                                 /* :145-147 */
  if (a->flags & STAN_CUDA_LP_LINEAR) logp += lin_term;                                /* :148-150 */
  if (a->flags & STAN_CUDA_LP_SCALE) logp += lg_ypp;                                   /* :151-153 */
  if (a->lp) *a->lp = logp;
  if (a->aux) { a->aux[0] = sum_d; a->aux[1] = witness; a->aux[2] = 0; a->aux[3] = lg_y1; }
  write_dalpha(a, d, sum_d);
  return isfinite(witness) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;                       /* :120 check_finite(theta) */
}

/* stan/math/prim/scal/fun/log1m_exp.hpp:44-54 */
static double oracle_log1m_exp(double a) {
  if (a >= 0) return NAN;
  if (a > -0.693147) return log(-expm1(a));
  return log1p(-exp(a));
}

/* ordered_logistic_glm_lpmf.hpp:93-199: location = x*beta (no intercept); d = d1 - d2 */
static int ordered_logistic_glm(const stan_cuda_glm_args* a, double* d) {
  const int64_t ncls = a->cuts_len + 1;
  double logp = 0.0, sum_d = 0.0, witness = 0.0;
  double* loc = linear_predictor(a);
  if (!loc) return STAN_CUDA_ERR_ARG;
  const int want_dcuts = (a->flags & STAN_CUDA_NEED_DCUTS) && a->dcuts;
  if (want_dcuts) for (int64_t j = 0; j < a->cuts_len; ++j) a->dcuts[j] = 0.0;
  for (int64_t i = 0; i < a->n; ++i) {
    const int64_t c = (int64_t)y_at(a, i);
    const double cy1 = c != ncls ? a->cuts[c - 1] : INFINITY;                       /* :104-115 */
    const double cy2 = c != 1 ? a->cuts[c - 2] : -INFINITY;
    const double cut2 = loc[i] - cy2, cut1 = loc[i] - cy1;                          /* :124-125 */
    const double m1 = (cut1 > 0.0 ? -cut1 : 0.0) - log1p(exp(-fabs(cut1)));        /* :128-129 */
    const double m2 = (cut2 <= 0.0 ? cut2 : 0.0) - log1p(exp(-fabs(cut2)));        /* :130-131 */
    if (c == 1) logp += m1;                                                         /* :137-144 */
    else if (c == ncls) logp += m2;
    else logp += m2 + oracle_log1m_exp(cut1 - cut2) + m1;
    const double e1 = exp(-cut1), e2 = exp(-cut2), ecd = exp(cy2 - cy1);            /* :162-164 */
    const double d1 = (cut2 > 0 ? e2 / (1 + e2) : 1 / (1 + exp(cut2))) - ecd / (ecd - 1);  /* :165-167 */
    const double d2 = 1 / (1 - ecd) - (cut1 > 0 ? e1 / (1 + e1) : 1 / (1 + exp(cut1)));    /* :168-171 */
    d[i] = d1 - d2;                                                                 /* :173 */
    sum_d += d[i];
    if (want_dcuts) {                                                               /* :192-200 */
      if (c != ncls) a->dcuts[c - 1] += d2;
      if (c != 1) a->dcuts[c - 2] -= d1;
    }
    witness += loc[i];                                                              /* :119 */
  }
  free(loc);
  if (a->lp) *a->lp = (a->flags & STAN_CUDA_LP_MAIN) ? logp : 0.0;
  if (a->aux) { a->aux[0] = sum_d; a->aux[1] = witness; a->aux[2] = 0; a->aux[3] = 0; }
  return isfinite(witness) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* categorical_logit_glm_lpmf.hpp:74-190: lin = x * beta + alpha' (n x C), row-wise softmax.
 * beta is k x C column-major, dbeta likewise; writes lp, dalpha[C], dbeta[k*C] itself. */
static int categorical_logit_glm(const stan_cuda_glm_args* a) {
  const int64_t n = a->n, k = a->k, C = a->n_classes;
  double* lin = (double*)calloc((size_t)(C > 0 ? C : 1), sizeof(double));
  double* e = (double*)calloc((size_t)(C > 0 ? C : 1), sizeof(double));
  if (!lin || !e) { free(lin); free(e); return STAN_CUDA_ERR_ARG; }
  const int want_da = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha;
  const int want_db = (a->flags & STAN_CUDA_NEED_DBETA) && a->dbeta;
  if (want_da) for (int64_t c = 0; c < C; ++c) a->dalpha[c] = 0.0;
  if (want_db) for (int64_t q = 0; q < k * C; ++q) a->dbeta[q] = 0.0;
  double logp = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    double mx = -INFINITY;
    for (int64_t c = 0; c < C; ++c) {                                               /* :82-83 */
      double s = 0.0;
      for (int64_t j = 0; j < k; ++j) s += x_at(a, i, j) * a->beta[c * k + j];
      lin[c] = s + a->alpha[c];
      if (lin[c] > mx) mx = lin[c];                                                 /* :84-85 */
    }
    double sum = 0.0;
    for (int64_t c = 0; c < C; ++c) { e[c] = exp(lin[c] - mx); sum += e[c]; }       /* :88-91 */
    const double inv = 1.0 / sum;
    const int64_t yc = (int64_t)y_at(a, i) - 1;
    logp += log(inv) - mx + lin[yc];                                                /* :93, :98-104 */
    for (int64_t c = 0; c < C; ++c) {
      const double d = (c == yc ? 1.0 : 0.0) - e[c] * inv;                          /* :147-148, :157-159, :172-178 */
      if (want_da) a->dalpha[c] += d;
      if (want_db) for (int64_t j = 0; j < k; ++j) a->dbeta[c * k + j] += x_at(a, i, j) * d;
    }
  }
  free(lin);
  free(e);
  if (a->lp) *a->lp = (a->flags & STAN_CUDA_LP_MAIN) ? logp : 0.0;
  if (a->aux) { a->aux[0] = 0; a->aux[1] = logp; a->aux[2] = 0; a->aux[3] = 0; }
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;                         /* :109 */
}

int oracle_glm_eval(const stan_cuda_glm_args* a) {
  if (!a || a->n < 0 || a->k < 0) return STAN_CUDA_ERR_ARG;
  if (a->family == STAN_CUDA_CATEGORICAL_LOGIT) {
    if (a->n_classes < 1 || a->alpha_len != a->n_classes) return STAN_CUDA_ERR_ARG;
    return categorical_logit_glm(a);
  }
  if (a->family != STAN_CUDA_ORDERED_LOGISTIC && a->alpha_len != 1 && a->alpha_len != a->n) return STAN_CUDA_ERR_ARG;
  double* d = (double*)calloc((size_t)(a->n > 0 ? a->n : 1), sizeof(double));
  if (!d) return STAN_CUDA_ERR_ARG;
  int rc;
  switch (a->family) {
    case STAN_CUDA_BERNOULLI_LOGIT: rc = bernoulli_logit_glm(a, d); break;
    case STAN_CUDA_POISSON_LOG: rc = poisson_log_glm(a, d); break;
    case STAN_CUDA_NORMAL_ID:
      if (a->sigma_len != 1 && a->sigma_len != a->n) { free(d); return STAN_CUDA_ERR_ARG; }
      rc = normal_id_glm(a, d);
      break;
    case STAN_CUDA_NEG_BINOMIAL_2_LOG:
      if (a->sigma_len != 1 && a->sigma_len != a->n) { free(d); return STAN_CUDA_ERR_ARG; }
      rc = neg_binomial_2_log_glm(a, d);
      break;
    case STAN_CUDA_ORDERED_LOGISTIC:
      if (a->cuts_len < 0 || (a->cuts_len > 0 && !a->cuts)) { free(d); return STAN_CUDA_ERR_ARG; }
      rc = ordered_logistic_glm(a, d);
      break;
    default: free(d); return STAN_CUDA_ERR_ARG;
  }
  if (rc >= 0 && (a->flags & STAN_CUDA_NEED_DBETA) && a->dbeta) xt_times(a, d, a->dbeta);
  /* var x: (beta * d^T)^T, i.e. dx_ij = d_i * beta_j (bernoulli_logit_glm_lpmf.hpp:125-128,
   * poisson_log_glm_lpmf.hpp:119-122, normal_id_glm_lpdf.hpp:121-124, neg_binomial_2_log_glm_lpmf.hpp:163-167,
   * ordered_logistic_glm_lpmf.hpp:186-190) */
  if (rc >= 0 && (a->flags & STAN_CUDA_NEED_DX) && a->dx) {
    const int64_t ld = a->lddx > 0 ? a->lddx : a->n;
    for (int64_t j = 0; j < a->k; ++j)
      for (int64_t i = 0; i < a->n; ++i) a->dx[j * ld + i] = a->beta[j] * d[i];
  }
  free(d);
  return rc;
}

/* stan/math/prim/scal/prob/normal_lpdf.hpp:67-110.  Scalar operands accumulate their partial
 * over n through broadcast_array (prim/scal/meta/broadcast_array.hpp:9-33). */
int oracle_normal_lpdf_eval(const stan_cuda_normal_args* a) {
  if (!a || a->n < 0) return STAN_CUDA_ERR_ARG;
  double logp = 0.0;
  const int ym = a->y_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  for (int64_t n = 0; n < a->n; ++n) {
    const double sg = a->sigma[sm ? n : 0];
    const double inv_sigma = 1.0 / sg;                                         /* :72 */
    const double z = (a->y[ym ? n : 0] - a->mu[mm ? n : 0]) * inv_sigma;        /* :82-83 */
    const double z2 = z * z;
    if (a->flags & STAN_CUDA_LP_CONST) logp += ORACLE_NEG_LOG_SQRT_TWO_PI;     /* :89-91 */
    if (a->flags & STAN_CUDA_LP_SCALE) logp -= log(sg);                        /* :92-94 */
    if (a->flags & STAN_CUDA_LP_MAIN) logp += -0.5 * z2;                       /* :95-97 */
    const double scaled_diff = inv_sigma * z;                                  /* :99 */
    if (need_y) a->dy[ym ? n : 0] -= scaled_diff;                              /* :100-102 */
    if (need_mu) a->dmu[mm ? n : 0] += scaled_diff;                            /* :103-105 */
    if (need_sigma) a->dsigma[sm ? n : 0] += -inv_sigma + inv_sigma * z2;      /* :106-109 */
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* cauchy_lpdf (stan/math/prim/scal/prob/cauchy_lpdf.hpp:35-113) and lognormal_lpdf
 * (stan/math/prim/scal/prob/lognormal_lpdf.hpp:19-133) on the operands of normal_lpdf; density 0 is normal_lpdf
 * itself with its eager checks (normal_lpdf.hpp:51-53) turned into BAD_OPERAND.  Scalar operands accumulate their
 * partial over n (broadcast_array). */
#define ORACLE_NEG_LOG_PI (-1.14472988584940017414342735135)
/* gamma_lpdf(y | alpha, beta) (stan/math/prim/scal/prob/gamma_lpdf.hpp:41-145; alpha in the mu slot, beta in the sigma
 * slot) and exponential_lpdf(y | beta) (exponential_lpdf.hpp:42-95; beta in the sigma slot, mu ignored) */
static int oracle_gamma_exponential(int density, const stan_cuda_normal_args* a) {
  const int gam = density == STAN_CUDA_LOCSCALE_GAMMA;
  for (int64_t i = 0; i < a->y_len; ++i)
    if (isnan(a->y[i]) || (!gam && a->y[i] < 0)) return STAN_CUDA_BAD_OPERAND;              /* gamma :55, exponential :55 */
  if (gam)
    for (int64_t i = 0; i < a->mu_len; ++i)
      if (!(a->mu[i] > 0.0) || !isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;           /* :56 */
  for (int64_t i = 0; i < a->sigma_len; ++i)
    if (!(a->sigma[i] > 0.0) || !isfinite(a->sigma[i])) return STAN_CUDA_BAD_OPERAND;       /* :57 / :56 */
  if (gam)
    for (int64_t i = 0; i < a->y_len; ++i)
      if (a->y[i] < 0) {                                                                    /* :69-74 */
        if (a->lp) *a->lp = -INFINITY;
        return STAN_CUDA_LOG_ZERO;
      }
  const int ym = a->y_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_mu = gam && (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  double logp = 0.0;
  for (int64_t n = 0; n < a->n; ++n) {
    const double y = a->y[ym ? n : 0], beta = a->sigma[sm ? n : 0];
    if (gam) {
      const double alpha = a->mu[mm ? n : 0];
      const double log_y = y > 0 ? log(y) : 0.0;                                            /* :86-90 */
      const double log_beta = log(beta);
      if (a->flags & STAN_CUDA_LP_DOF) logp -= lgamma(alpha);                               /* :118-120 */
      if (a->flags & STAN_CUDA_LP_SCALE) logp += alpha * log_beta;                          /* :121-123 */
      if (a->flags & STAN_CUDA_LP_Y) logp += (alpha - 1.0) * log_y;                         /* :124-126 */
      if (a->flags & STAN_CUDA_LP_MAIN) logp -= beta * y;                                   /* :127-129 */
      if (need_y) a->dy[ym ? n : 0] += (alpha - 1) / y - beta;                              /* :131-133 */
      if (need_mu) a->dmu[mm ? n : 0] += -oracle_digamma(alpha) + log_beta + log_y;         /* :134-137 */
      if (need_sigma) a->dsigma[sm ? n : 0] += alpha / beta - y;                            /* :138-140 */
    } else {
      if (a->flags & STAN_CUDA_LP_SCALE) logp += log(beta);                                 /* :76-78 */
      if (a->flags & STAN_CUDA_LP_MAIN) logp -= beta * y;                                   /* :79-81 */
      if (need_y) a->dy[ym ? n : 0] -= beta;                                                /* :83-85 */
      if (need_sigma) a->dsigma[sm ? n : 0] += 1 / beta - y;                                /* :86-88 */
    }
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* beta_lpdf(y | alpha, beta) (stan/math/prim/scal/prob/beta_lpdf.hpp:40-181) and weibull_lpdf(y | alpha, sigma)
 * (weibull_lpdf.hpp:29-143); alpha in the mu slot, beta / sigma in the sigma slot */
static int oracle_beta_weibull(int density, const stan_cuda_normal_args* a) {
  const int bet = density == STAN_CUDA_LOCSCALE_BETA;
  for (int64_t i = 0; i < a->mu_len; ++i)
    if (!(a->mu[i] > 0.0) || !isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;             /* beta :48, weibull :39 */
  for (int64_t i = 0; i < a->sigma_len; ++i)
    if (!(a->sigma[i] > 0.0) || !isfinite(a->sigma[i])) return STAN_CUDA_BAD_OPERAND;       /* :49 / :40 */
  for (int64_t i = 0; i < a->y_len; ++i) {
    if (bet && (isnan(a->y[i]) || a->y[i] < 0 || a->y[i] > 1)) return STAN_CUDA_BAD_OPERAND; /* :50, :53-54 */
    if (!bet && !isfinite(a->y[i])) return STAN_CUDA_BAD_OPERAND;                           /* :38 */
  }
  if (!bet)
    for (int64_t i = 0; i < a->y_len; ++i)
      if (a->y[i] < 0) {                                                                    /* :56-61 */
        if (a->lp) *a->lp = -INFINITY;
        return STAN_CUDA_LOG_ZERO;
      }
  const int ym = a->y_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  double logp = 0.0;
  for (int64_t n = 0; n < a->n; ++n) {
    const double y = a->y[ym ? n : 0], alpha = a->mu[mm ? n : 0], sg = a->sigma[sm ? n : 0];
    if (bet) {
      const double log_y = log(y), log1m_y = log1p(-y);                                     /* :87-94 */
      if (a->flags & STAN_CUDA_LP_SCALE) logp += lgamma(alpha + sg);                        /* :150-152 */
      if (a->flags & STAN_CUDA_LP_DOF) logp -= lgamma(alpha);                               /* :153-155 */
      if (a->flags & STAN_CUDA_LP_LINEAR) logp -= lgamma(sg);                               /* :156-158 */
      if (a->flags & STAN_CUDA_LP_Y) logp += (alpha - 1.0) * log_y;                         /* :159-161 */
      if (a->flags & STAN_CUDA_LP_MAIN) logp += (sg - 1.0) * log1m_y;                       /* :162-164 */
      if (need_y) a->dy[ym ? n : 0] += (alpha - 1) / y + (sg - 1) / (y - 1);                /* :166-169 */
      if (need_mu) a->dmu[mm ? n : 0] += log_y + oracle_digamma(alpha + sg) - oracle_digamma(alpha);       /* :170-173 */
      if (need_sigma) a->dsigma[sm ? n : 0] += log1m_y + oracle_digamma(alpha + sg) - oracle_digamma(sg);  /* :174-177 */
    } else {
      const double inv_sigma = 1.0 / sg, log_y = log(y), log_sigma = log(sg);
      const double pw = pow(y * inv_sigma, alpha);                                          /* :95-103 */
      if (a->flags & STAN_CUDA_LP_DOF) logp += log(alpha);                                  /* :111-113 */
      if (a->flags & STAN_CUDA_LP_Y) logp += (alpha - 1.0) * log_y;                         /* :114-116 */
      if (a->flags & STAN_CUDA_LP_SCALE) logp -= alpha * log_sigma;                         /* :117-119 */
      if (a->flags & STAN_CUDA_LP_MAIN) logp -= pw;                                         /* :120-122 */
      if (need_y) {
        const double inv_y = 1.0 / y;
        a->dy[ym ? n : 0] += (alpha - 1.0) * inv_y - alpha * pw * inv_y;                    /* :125-130 */
      }
      if (need_mu) a->dmu[mm ? n : 0] += 1.0 / alpha + (1.0 - pw) * (log_y - log_sigma);    /* :131-135 */
      if (need_sigma) a->dsigma[sm ? n : 0] += alpha * inv_sigma * (pw - 1.0);              /* :136-139 */
    }
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* inv_gamma_lpdf(y | alpha, beta), stan/math/prim/scal/prob/inv_gamma_lpdf.hpp:33-146; alpha in the mu slot, beta in the
 * sigma slot */
static int oracle_inv_gamma(const stan_cuda_normal_args* a) {
  for (int64_t i = 0; i < a->y_len; ++i) if (isnan(a->y[i])) return STAN_CUDA_BAD_OPERAND;                 /* :41 */
  for (int64_t i = 0; i < a->mu_len; ++i)
    if (!(a->mu[i] > 0.0) || !isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;                           /* :42 */
  for (int64_t i = 0; i < a->sigma_len; ++i)
    if (!(a->sigma[i] > 0.0) || !isfinite(a->sigma[i])) return STAN_CUDA_BAD_OPERAND;                     /* :43 */
  for (int64_t i = 0; i < a->y_len; ++i)
    if (a->y[i] <= 0) {                                                                                   /* :58-63 */
      if (a->lp) *a->lp = -INFINITY;
      return STAN_CUDA_LOG_ZERO;
    }
  const int ym = a->y_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  double logp = 0.0;
  for (int64_t n = 0; n < a->n; ++n) {
    const double y = a->y[ym ? n : 0], alpha = a->mu[mm ? n : 0], beta = a->sigma[sm ? n : 0];
    const double log_y = log(y), inv_y = 1.0 / y, log_beta = log(beta);                      /* :76-110 */
    if (a->flags & STAN_CUDA_LP_DOF) logp -= lgamma(alpha);                                 /* :115-117 */
    if (a->flags & STAN_CUDA_LP_SCALE) logp += alpha * log_beta;                            /* :118-120 */
    if (a->flags & STAN_CUDA_LP_Y) logp -= (alpha + 1.0) * log_y;                           /* :121-123 */
    if (a->flags & STAN_CUDA_LP_MAIN) logp -= beta * inv_y;                                 /* :124-126 */
    if (need_y) a->dy[ym ? n : 0] += -(alpha + 1) * inv_y + beta * inv_y * inv_y;           /* :128-131 */
    if (need_mu) a->dmu[mm ? n : 0] += -oracle_digamma(alpha) + log_beta - log_y;           /* :132-135 */
    if (need_sigma) a->dsigma[sm ? n : 0] += alpha / beta - inv_y;                          /* :136-138 */
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

int oracle_locscale_lpdf_eval(int density, const stan_cuda_normal_args* a) {
  if (!a || a->n < 0 || density < STAN_CUDA_LOCSCALE_NORMAL || density > STAN_CUDA_LOCSCALE_INV_GAMMA) return STAN_CUDA_ERR_ARG;
  if (a->n == 0) {
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  if (density == STAN_CUDA_LOCSCALE_INV_GAMMA) return oracle_inv_gamma(a);
  if (density >= STAN_CUDA_LOCSCALE_BETA) return oracle_beta_weibull(density, a);
  if (density >= STAN_CUDA_LOCSCALE_GAMMA) return oracle_gamma_exponential(density, a);
  const int lognormal = density == STAN_CUDA_LOCSCALE_LOGNORMAL, normal = density == STAN_CUDA_LOCSCALE_NORMAL;
  for (int64_t i = 0; i < a->y_len; ++i)
    if (isnan(a->y[i]) || (lognormal && a->y[i] < 0)) return STAN_CUDA_BAD_OPERAND;          /* cauchy :47, lognormal :24-25 */
  for (int64_t i = 0; i < a->mu_len; ++i) if (!isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;            /* :48 / :26 */
  for (int64_t i = 0; i < a->sigma_len; ++i)
    if (!(a->sigma[i] > 0.0) || (!normal && !isfinite(a->sigma[i]))) return STAN_CUDA_BAD_OPERAND;          /* :49 / :27 */
  if (normal) return oracle_normal_lpdf_eval(a);
  if (lognormal)
    for (int64_t i = 0; i < a->y_len; ++i)
      if (a->y[i] <= 0) {                                                                                   /* :41-45 */
        if (a->lp) *a->lp = -INFINITY;
        return STAN_CUDA_LOG_ZERO;
      }
  const int ym = a->y_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  double logp = 0.0;
  if (lognormal && (a->flags & STAN_CUDA_LP_CONST)) logp += a->n * ORACLE_NEG_LOG_SQRT_TWO_PI;              /* :92-94 */
  for (int64_t n = 0; n < a->n; ++n) {
    const double y = a->y[ym ? n : 0], mu = a->mu[mm ? n : 0], sg = a->sigma[sm ? n : 0];
    const double inv_sigma = 1.0 / sg;
    if (!lognormal) {                                                                /* cauchy_lpdf.hpp:77-110 */
      const double sigma_squared = sg * sg;
      const double y_minus_mu = y - mu;
      const double y_minus_mu_squared = y_minus_mu * y_minus_mu;
      const double y_minus_mu_over_sigma = y_minus_mu * inv_sigma;
      const double y_minus_mu_over_sigma_squared = y_minus_mu_over_sigma * y_minus_mu_over_sigma;
      if (a->flags & STAN_CUDA_LP_CONST) logp += ORACLE_NEG_LOG_PI;
      if (a->flags & STAN_CUDA_LP_SCALE) logp -= log(sg);
      if (a->flags & STAN_CUDA_LP_MAIN) logp -= log1p(y_minus_mu_over_sigma_squared);
      if (need_y) a->dy[ym ? n : 0] -= 2 * y_minus_mu / (sigma_squared + y_minus_mu_squared);
      if (need_mu) a->dmu[mm ? n : 0] += 2 * y_minus_mu / (sigma_squared + y_minus_mu_squared);
      if (need_sigma)
        a->dsigma[sm ? n : 0] += (y_minus_mu_squared - sigma_squared) * inv_sigma / (sigma_squared + y_minus_mu_squared);
    } else {                                                                         /* lognormal_lpdf.hpp:96-131 */
      const double inv_sigma_sq = inv_sigma * inv_sigma;
      const double log_y = log(y), inv_y = 1 / y;
      const double logy_m_mu = log_y - mu;
      const double logy_m_mu_sq = logy_m_mu * logy_m_mu;
      const double logy_m_mu_div_sigma = logy_m_mu * inv_sigma_sq;
      if (a->flags & STAN_CUDA_LP_SCALE) logp -= log(sg);
      if (a->flags & STAN_CUDA_LP_Y) logp -= log_y;
      if (a->flags & STAN_CUDA_LP_MAIN) logp -= 0.5 * logy_m_mu_sq * inv_sigma_sq;
      if (need_y) a->dy[ym ? n : 0] -= (1 + logy_m_mu_div_sigma) * inv_y;
      if (need_mu) a->dmu[mm ? n : 0] += logy_m_mu_div_sigma;
      if (need_sigma) a->dsigma[sm ? n : 0] += (logy_m_mu_div_sigma * logy_m_mu - 1) * inv_sigma;
    }
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* stan/math/prim/scal/prob/student_t_lpdf.hpp:55-201.  The eager checks (:67-70) become BAD_OPERAND; scalar
 * operands accumulate their partial over n (prim/scal/meta/broadcast_array.hpp:9-33). */
#define ORACLE_NEG_LOG_SQRT_PI (-0.5723649429247000870717135)
int oracle_student_t_lpdf_eval(const stan_cuda_student_t_args* a) {
  if (!a || a->n < 0) return STAN_CUDA_ERR_ARG;
  if (a->n == 0) {                                                              /* :57-59 */
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  for (int64_t i = 0; i < a->y_len; ++i) if (isnan(a->y[i])) return STAN_CUDA_BAD_OPERAND;                       /* :67 */
  for (int64_t i = 0; i < a->nu_len; ++i) if (!(a->nu[i] > 0.0) || !isfinite(a->nu[i])) return STAN_CUDA_BAD_OPERAND;   /* :68 */
  for (int64_t i = 0; i < a->mu_len; ++i) if (!isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;                 /* :69 */
  for (int64_t i = 0; i < a->sigma_len; ++i)
    if (!(a->sigma[i] > 0.0) || !isfinite(a->sigma[i])) return STAN_CUDA_BAD_OPERAND;                            /* :70 */
  const int ym = a->y_len != 1, nm = a->nu_len != 1, mm = a->mu_len != 1, sm = a->sigma_len != 1;
  const int need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  const int need_nu = (a->flags & STAN_CUDA_NEED_DNU) && a->dnu;
  const int need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const int need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  if (need_y) for (int64_t i = 0; i < a->y_len; ++i) a->dy[i] = 0.0;
  if (need_nu) for (int64_t i = 0; i < a->nu_len; ++i) a->dnu[i] = 0.0;
  if (need_mu) for (int64_t i = 0; i < a->mu_len; ++i) a->dmu[i] = 0.0;
  if (need_sigma) for (int64_t i = 0; i < a->sigma_len; ++i) a->dsigma[i] = 0.0;
  double logp = 0.0;
  for (int64_t n = 0; n < a->n; ++n) {
    const double y = a->y[ym ? n : 0], nu = a->nu[nm ? n : 0], mu = a->mu[mm ? n : 0], sg = a->sigma[sm ? n : 0];
    const double half_nu = 0.5 * nu;                                                                 /* :90 */
    const double zs = (y - mu) / sg;
    const double sq = zs * zs / nu;                                                                  /* :145-146 */
    const double l1p = log1p(sq);                                                                    /* :147 */
    if (a->flags & STAN_CUDA_LP_CONST) logp += ORACLE_NEG_LOG_SQRT_PI;                               /* :158-160 */
    if (a->flags & STAN_CUDA_LP_DOF) logp += lgamma(half_nu + 0.5) - lgamma(half_nu) - 0.5 * log(nu);  /* :161-163 */
    if (a->flags & STAN_CUDA_LP_SCALE) logp -= log(sg);                                              /* :164-166 */
    if (a->flags & STAN_CUDA_LP_MAIN) logp -= (half_nu + 0.5) * l1p;                                 /* :167-169 */
    if (need_y)                                                                                      /* :171-176 */
      a->dy[ym ? n : 0] += -(half_nu + 0.5) * 1.0 / (1.0 + sq) * (2.0 * (y - mu) / (sg * sg) / nu);
    if (need_nu) {                                                                                   /* :177-185 */
      const double inv_nu = 1.0 / nu;
      a->dnu[nm ? n : 0] += 0.5 * oracle_digamma(half_nu + 0.5) - 0.5 * oracle_digamma(half_nu) - 0.5 * inv_nu - 0.5 * l1p
                            + (half_nu + 0.5) * (1.0 / (1.0 + sq) * sq * inv_nu);
    }
    if (need_mu)                                                                                     /* :186-191 */
      a->dmu[mm ? n : 0] -= (half_nu + 0.5) / (1.0 + sq) * (2.0 * (mu - y) / (sg * sg * nu));
    if (need_sigma) {                                                                                /* :192-198 */
      const double inv_sigma = 1.0 / sg;
      a->dsigma[sm ? n : 0] += -inv_sigma + (nu + 1.0) / (1.0 + sq) * (sq * inv_sigma);
    }
  }
  if (a->lp) *a->lp = logp;
  return isfinite(logp) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

/* Dense product C = op(A) * op(B): what Eigen computes for the reference's `multiply`
 * (stan/math/prim/mat/fun/multiply.hpp:66-83) and for the three products of multiply_mat_vari
 * (stan/math/rev/mat/fun/multiply.hpp:99, 128-131), as the textbook triple loop with k ascending. */
int oracle_multiply(const stan_cuda_multiply_args* g) {
  if (!g || g->m < 0 || g->n < 0 || g->k < 0) return STAN_CUDA_ERR_ARG;
  for (int64_t j = 0; j < g->n; ++j)
    for (int64_t i = 0; i < g->m; ++i) {
      double s = 0.0;
      for (int64_t l = 0; l < g->k; ++l) {
        const double av = g->trans_a ? g->a[i * g->lda + l] : g->a[l * g->lda + i];
        const double bv = g->trans_b ? g->b[l * g->ldb + j] : g->b[j * g->ldb + l];
        s += av * bv;
      }
      g->c[j * g->ldc + i] = s;
    }
  return STAN_CUDA_OK;
}

/* Cholesky factor, lower: what Eigen::LLT computes for the reference's cholesky_decompose
 * (stan/math/prim/mat/fun/cholesky_decompose.hpp:34-62; rev: stan/math/rev/mat/fun/cholesky_decompose.hpp:396-405),
 * as the textbook column algorithm.  Reads the lower triangle of a; l gets zeros above the diagonal.  NOT_POS_DEF
 * where check_pos_definite throws (a pivot that is not positive, NaN included). */
int oracle_cholesky_decompose(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl) {
  if (n < 0) return STAN_CUDA_ERR_ARG;
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = 0; i < n; ++i) l[j * ldl + i] = 0.0;
  for (int64_t j = 0; j < n; ++j) {
    double s = a[j * lda + j];
    for (int64_t k = 0; k < j; ++k) s -= l[k * ldl + j] * l[k * ldl + j];
    if (!(s > 0.0)) return STAN_CUDA_NOT_POS_DEF;
    const double ljj = sqrt(s);
    l[j * ldl + j] = ljj;
    for (int64_t i = j + 1; i < n; ++i) {
      double t = a[j * lda + i];
      for (int64_t k = 0; k < j; ++k) t -= l[k * ldl + i] * l[k * ldl + j];
      l[j * ldl + i] = t / ljj;
    }
  }
  return STAN_CUDA_OK;
}

/* Reverse sweep of the Cholesky factor, unblocked: the loops of cholesky_scalar::chain()
 * (stan/math/rev/mat/fun/cholesky_decompose.hpp:236-252), on full n x n arrays instead of the packed ones.
 * Reads the lower triangles of l and l_adj, writes the lower triangle of a_adj (zeros above). */
int oracle_cholesky_adjoint(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj, double* a_adj,
                            int64_t lda) {
  if (n < 0) return STAN_CUDA_ERR_ARG;
  double* adjL = (double*)calloc((size_t)(n * n > 0 ? n * n : 1), sizeof(double));
  if (!adjL) return STAN_CUDA_ERR_ARG;
#define ADJL(i, j) adjL[(j) * n + (i)]
#define LA(i, j) l[(j) * ldl + (i)]
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = 0; i < n; ++i) {
      ADJL(i, j) = i >= j ? l_adj[j * ldadj + i] : 0.0;
      a_adj[j * lda + i] = 0.0;
    }
  for (int64_t i = n - 1; i >= 0; --i) {
    for (int64_t j = i; j >= 0; --j) {
      double adjA;
      if (i == j) {
        adjA = 0.5 * ADJL(i, j) / LA(i, j);                                  /* :239-240 */
      } else {
        adjA = ADJL(i, j) / LA(j, j);                                        /* :242 */
        ADJL(j, j) -= ADJL(i, j) * LA(i, j) / LA(j, j);                      /* :243-244 */
      }
      for (int64_t k = j - 1; k >= 0; --k) {                                 /* :246-249 */
        ADJL(i, k) -= adjA * LA(j, k);
        ADJL(j, k) -= adjA * LA(i, k);
      }
      a_adj[j * lda + i] = adjA;                                             /* :250 */
    }
  }
#undef ADJL
#undef LA
  free(adjL);
  return STAN_CUDA_OK;
}

/* Triangular solve x = op(A)^-1 b by substitution: what Eigen's triangularView<>.solve computes for
 * mdivide_left_tri (stan/math/prim/mat/fun/mdivide_left_tri.hpp:34-36, 75; rev: stan/math/rev/mat/fun/mdivide_left_tri.hpp
 * :78, :98-101).  lower: the lower triangle of a is read, else the upper; trans: op(A) = A'. */
int oracle_tri_solve(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b, int64_t ldb,
                     double* x, int64_t ldx) {
  if (n < 0 || m < 0) return STAN_CUDA_ERR_ARG;
  const int forward = (lower != 0) != (trans != 0);  /* op(A) is lower triangular */
  for (int64_t c = 0; c < m; ++c) {
    for (int64_t i = 0; i < n; ++i) x[c * ldx + i] = b[c * ldb + i];
    for (int64_t s = 0; s < n; ++s) {
      const int64_t i = forward ? s : n - 1 - s;
      double v = x[c * ldx + i];
      if (forward) {
        for (int64_t k = 0; k < i; ++k) v -= (trans ? a[i * lda + k] : a[k * lda + i]) * x[c * ldx + k];
      } else {
        for (int64_t k = i + 1; k < n; ++k) v -= (trans ? a[i * lda + k] : a[k * lda + i]) * x[c * ldx + k];
      }
      x[c * ldx + i] = v / a[i * lda + i];
    }
  }
  return STAN_CUDA_OK;
}

/* Squared-exponential covariance: internal::gp_exp_quad_cov of
 * stan/math/prim/mat/fun/gp_exp_quad_cov.hpp:40-170 -- cov(i, j) = sigma_sq * exp(squared_distance(x1_i, x2_j) * factor),
 * the diagonal of the one-sample form (x2 == NULL) set to sigma_sq (:53, :88).  Points dimension-major. */
int oracle_gp_exp_quad_cov(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2, double sigma_sq, double factor,
                           double* cov, int64_t ldc) {
  if (n1 < 0 || n2 < 0 || dim < 1) return STAN_CUDA_ERR_ARG;
  const int symmetric = x2 == NULL;
  if (symmetric) {
    x2 = x1;
    n2 = n1;
  }
  for (int64_t j = 0; j < n2; ++j)
    for (int64_t i = 0; i < n1; ++i) {
      if (symmetric && i == j) {
        cov[j * ldc + i] = sigma_sq;
        continue;
      }
      double d2 = 0.0;
      for (int64_t d = 0; d < dim; ++d) {
        const double diff = x1[d * n1 + i] - x2[d * n2 + j];
        d2 += diff * diff;
      }
      cov[j * ldc + i] = sigma_sq * exp(d2 * factor);
    }
  return STAN_CUDA_OK;
}

/* Vectorised elementwise densities with an integer outcome.
 *   bernoulli_logit_lpmf  stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:51-80
 *   poisson_log_lpmf      stan/math/prim/scal/prob/poisson_log_lpmf.hpp:51-93
 *   poisson_lpmf          stan/math/prim/scal/prob/poisson_lpmf.hpp:20-77   (oracle_density_poisson below)
 *   bernoulli_lpmf        stan/math/prim/scal/prob/bernoulli_lpmf.hpp:28-112 (oracle_density_bernoulli below)
 * Return codes as the product's: BAD_OPERAND where the reference's eager checks throw (:42-43 / :38-39),
 * LOG_ZERO where poisson_log_lpmf returns LOG_ZERO before building partials (:51-62). */
/* poisson_lpmf(n | lambda), stan/math/prim/scal/prob/poisson_lpmf.hpp:20-77 */
static int oracle_density_poisson(const stan_cuda_density_args* a) {
  const int ym = a->y_len != 1, tm = a->theta_len != 1;
  const int need_d = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dtheta;
  for (int64_t i = 0; i < a->y_len; ++i)
    if (a->y[i] < 0) return STAN_CUDA_BAD_OPERAND;                                  /* :31 */
  for (int64_t i = 0; i < a->theta_len; ++i)
    if (isnan(a->theta[i]) || a->theta[i] < 0) return STAN_CUDA_BAD_OPERAND;        /* :32-33 */
  for (int64_t i = 0; i < a->n; ++i)
    if (isinf(a->theta[tm ? i : 0])) {                                              /* :45-49 */
      if (a->lp) *a->lp = -INFINITY;
      return STAN_CUDA_LOG_ZERO;
    }
  for (int64_t i = 0; i < a->n; ++i)
    if (a->theta[tm ? i : 0] == 0 && a->y[ym ? i : 0] != 0) {                       /* :50-54 */
      if (a->lp) *a->lp = -INFINITY;
      return STAN_CUDA_LOG_ZERO;
    }
  double logp = 0.0;
  if (need_d) for (int64_t i = 0; i < a->theta_len; ++i) a->dtheta[i] = 0.0;
  for (int64_t i = 0; i < a->n; ++i) {
    const int y = a->y[ym ? i : 0];
    const double lam = a->theta[tm ? i : 0];
    if (!(lam == 0 && y == 0)) {                                                    /* :59 */
      if (a->flags & STAN_CUDA_LP_CONST) logp -= lgamma(y + 1.0);                   /* :60-62 */
      if (a->flags & STAN_CUDA_LP_MAIN)                                             /* :63-66; multiply_log(0, 0) = 0 */
        logp += ((y == 0 && lam == 0) ? 0.0 : y * log(lam)) - lam;
    }
    if (need_d) a->dtheta[tm ? i : 0] += y / lam - 1.0;                             /* :69-72 */
  }
  if (a->lp) *a->lp = logp;
  return STAN_CUDA_OK;
}

/* bernoulli_lpmf(n | theta), stan/math/prim/scal/prob/bernoulli_lpmf.hpp:28-112 */
static int oracle_density_bernoulli(const stan_cuda_density_args* a) {
  const int ym = a->y_len != 1, tm = a->theta_len != 1;
  const int need_d = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dtheta;
  const int main_term = (a->flags & STAN_CUDA_LP_MAIN) != 0;
  for (int64_t i = 0; i < a->y_len; ++i)
    if (a->y[i] < 0 || a->y[i] > 1) return STAN_CUDA_BAD_OPERAND;                   /* :41 */
  for (int64_t i = 0; i < a->theta_len; ++i)
    if (!isfinite(a->theta[i]) || a->theta[i] < 0.0 || a->theta[i] > 1.0) return STAN_CUDA_BAD_OPERAND;  /* :42-43 */
  double logp = 0.0;
  if (need_d) for (int64_t i = 0; i < a->theta_len; ++i) a->dtheta[i] = 0.0;
  if (!tm) {                                                                        /* :58-91 */
    int64_t sum = 0;
    for (int64_t i = 0; i < a->n; ++i) sum += a->y[ym ? i : 0];
    const double th = a->theta[0], N = (double)a->n;
    double d = 0.0;
    if (sum == a->n) {
      logp += N * log(th);
      d += N / th;
    } else if (sum == 0) {
      logp += N * log1p(-th);
      d += N / (th - 1);
    } else {
      logp += sum * log(th);
      logp += (a->n - sum) * log1p(-th);
      d += sum / th;
      d += (a->n - sum) / (th - 1);
    }
    if (need_d) a->dtheta[0] = d;
  } else {                                                                          /* :92-107 */
    for (int64_t i = 0; i < a->n; ++i) {
      const int y = a->y[ym ? i : 0];
      const double th = a->theta[i];
      logp += y == 1 ? log(th) : log1p(-th);
      if (need_d) a->dtheta[i] += y == 1 ? 1.0 / th : 1.0 / (th - 1);
    }
  }
  if (a->lp) *a->lp = main_term ? logp : 0.0;
  return STAN_CUDA_OK;
}

int oracle_density_eval(const stan_cuda_density_args* a) {
  if (!a || a->n < 0) return STAN_CUDA_ERR_ARG;
  if (a->density == STAN_CUDA_DENSITY_POISSON || a->density == STAN_CUDA_DENSITY_BERNOULLI) {
    if (a->n == 0) {
      if (a->lp) *a->lp = 0.0;
      return STAN_CUDA_OK;
    }
    return a->density == STAN_CUDA_DENSITY_POISSON ? oracle_density_poisson(a) : oracle_density_bernoulli(a);
  }
  const int ym = a->y_len != 1, tm = a->theta_len != 1;
  const int need_d = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dtheta;
  const int bern = a->density == STAN_CUDA_DENSITY_BERNOULLI_LOGIT;
  if (!bern && a->density != STAN_CUDA_DENSITY_POISSON_LOG) return STAN_CUDA_ERR_ARG;
  if (a->n == 0) {
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  for (int64_t i = 0; i < a->y_len; ++i)
    if (a->y[i] < 0 || (bern && a->y[i] > 1)) return STAN_CUDA_BAD_OPERAND;
  for (int64_t i = 0; i < a->theta_len; ++i)
    if (isnan(a->theta[i])) return STAN_CUDA_BAD_OPERAND;
  if (!bern) {
    for (int64_t i = 0; i < a->n; ++i) {
      const double al = a->theta[tm ? i : 0];
      if (al == INFINITY || (al == -INFINITY && a->y[ym ? i : 0] != 0)) {            /* :51-62 */
        if (a->lp) *a->lp = -INFINITY;
        return STAN_CUDA_LOG_ZERO;
      }
    }
  }
  double logp = 0.0;
  if (need_d) for (int64_t i = 0; i < a->theta_len; ++i) a->dtheta[i] = 0.0;
  for (int64_t i = 0; i < a->n; ++i) {
    const int y = a->y[ym ? i : 0];
    const double th = a->theta[tm ? i : 0];
    double d;
    if (bern) {
      const int sign = 2 * y - 1;                                                  /* :58 */
      const double ntheta = sign * th;
      const double e = exp(-ntheta);                                               /* :60 */
      if (a->flags & STAN_CUDA_LP_MAIN) {
        if (ntheta > 20.0) logp -= e;                                              /* :63-69 */
        else if (ntheta < -20.0) logp += ntheta;
        else logp -= log1p(e);
      }
      if (ntheta > 20.0) d = -e;                                                   /* :72-79 */
      else if (ntheta < -20.0) d = sign;
      else d = sign * e / (e + 1);
    } else {
      const double e = exp(th);                                                    /* :70-74 */
      if (!(th == -INFINITY && y == 0)) {                                          /* :77-86 */
        if (a->flags & STAN_CUDA_LP_CONST) logp -= lgamma(y + 1.0);
        if (a->flags & STAN_CUDA_LP_MAIN) logp += y * th - e;
      }
      d = y - e;                                                                   /* :88-90 */
    }
    if (need_d) a->dtheta[tm ? i : 0] += d;
  }
  if (a->lp) *a->lp = logp;
  return STAN_CUDA_OK;
}
