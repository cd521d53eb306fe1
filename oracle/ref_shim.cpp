// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A C shim around the UNMODIFIED reference (Stan Math 3.0.0 under /root/reference).
// It is compiled from the reference headers where they lie (recipe: oracle/Makefile,
// output oracle/_ref/libstanref.so) and is used only by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference leg: it validates the plain-C
// restatement (oracle/glm_oracle.c), generates the golden fixtures under tests/golden/
// and is timed as the CPU baseline.  Nothing under math_b200/ or stan/math/cuda/
// may link or load it.
//
// Every entry point builds stan::math::var operands, calls the reference function
// through its public signature, runs grad() and copies value + adjoints out:
//   bernoulli_logit_glm_lpmf  stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:46-138
//   poisson_log_glm_lpmf      stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp:44-131
//   normal_id_glm_lpdf        stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp:49-174
//   normal_lpdf               stan/math/prim/scal/prob/normal_lpdf.hpp:36-112
//   neg_binomial_2_log_glm_lpmf stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp:60-197
//   ordered_logistic_glm_lpmf stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:46-222
//   categorical_logit_glm_lpmf stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:38-193
//   map_rect                  stan/math/prim/mat/functor/map_rect.hpp:120-177
#include <stan/math.hpp>

#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

using stan::math::var;
typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_v;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_d;
typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> matrix_d;

enum { FAM_BERNOULLI_LOGIT = 0, FAM_POISSON_LOG = 1, FAM_NORMAL_ID = 2, FAM_NEG_BINOMIAL_2_LOG = 3 };

void set_err(char* err, int errlen, const char* what) {
  if (err && errlen > 0) {
    std::strncpy(err, what, errlen - 1);
    err[errlen - 1] = 0;
  }
}

// Runs f() under the reference's exception contract and maps it onto the shim's
// return codes: 0 ok, 1 std::domain_error, 2 std::invalid_argument, 3 anything else.
template <typename F>
int guarded(F&& f, char* err, int errlen) {
  int rc = 0;
  try {
    f();
  } catch (const std::domain_error& e) {
    set_err(err, errlen, e.what());
    rc = 1;
  } catch (const std::invalid_argument& e) {
    set_err(err, errlen, e.what());
    rc = 2;
  } catch (const std::exception& e) {
    set_err(err, errlen, e.what());
    rc = 3;
  }
  stan::math::recover_memory();
  return rc;
}

vector_v make_vars(const double* p, long n) {
  vector_v v(n);
  for (long i = 0; i < n; ++i) v[i] = p[i];
  return v;
}

void read_adj(const vector_v& v, double* out) {
  if (!out) return;
  for (long i = 0; i < v.size(); ++i) out[i] = v[i].adj();
}

template <bool propto, typename T_y, typename T_alpha, typename T_sigma>
var call_family(int family, const T_y& y, const std::vector<int>& yi,
                const matrix_d& x, const T_alpha& alpha, const vector_v& beta,
                const T_sigma& sigma) {
  switch (family) {
    case FAM_BERNOULLI_LOGIT:
      return stan::math::bernoulli_logit_glm_lpmf<propto>(yi, x, alpha, beta);
    case FAM_POISSON_LOG:
      return stan::math::poisson_log_glm_lpmf<propto>(yi, x, alpha, beta);
    case FAM_NEG_BINOMIAL_2_LOG:  // the precision phi travels in the sigma slot
      return stan::math::neg_binomial_2_log_glm_lpmf<propto>(yi, x, alpha, beta, sigma);
    default:
      return stan::math::normal_id_glm_lpdf<propto>(y, x, alpha, beta, sigma);
  }
}

}  // namespace

extern "C" {

// GLM value + reverse-mode gradient through the reference.
//   x            n*k doubles, column-major (Eigen::MatrixXd layout)
//   y            int32[n] for bernoulli/poisson, double[n] for normal_id
//   alpha_len    1 (var scalar) or n (var vector);  sigma_len 1 or n (normal_id only)
//   y_is_var     normal_id only: make y a var vector and return dy
// Outputs may be NULL.  seconds (optional) receives the wall time of call + grad().
int stanref_glm(int family, int propto, long n, long k, const double* x,
                const void* y, const double* alpha, long alpha_len,
                const double* beta, const double* sigma, long sigma_len,
                int y_is_var, double* lp, double* dbeta, double* dalpha,
                double* dsigma, double* dy, double* seconds, char* err,
                int errlen) {
  return guarded(
      [&]() {
        Eigen::Map<const matrix_d> xmap(x, n, k);
        matrix_d xm = xmap;
        std::vector<int> yi;
        vector_d yd(0);
        if (family == FAM_NORMAL_ID) {
          yd = Eigen::Map<const vector_d>(static_cast<const double*>(y), n);
        } else {
          const int* p = static_cast<const int*>(y);
          yi.assign(p, p + n);
        }
        vector_v beta_v = make_vars(beta, k);
        vector_v alpha_vec = make_vars(alpha, alpha_len);
        const bool has_scale = family == FAM_NORMAL_ID || family == FAM_NEG_BINOMIAL_2_LOG;
        vector_v sigma_vec = make_vars(sigma, has_scale ? sigma_len : 0);
        vector_v y_v = make_vars(yd.data(), y_is_var ? n : 0);
        auto t0 = std::chrono::steady_clock::now();
        var res;
        const bool av = alpha_len != 1, sv = (has_scale && sigma_len != 1);
        var a0 = alpha_len >= 1 ? alpha_vec[0] : var(0.0);
        var s0 = sigma_vec.size() >= 1 ? sigma_vec[0] : var(1.0);
#define STANREF_CALL(P, YARG)                                                        \
  (av ? (sv ? call_family<P>(family, YARG, yi, xm, alpha_vec, beta_v, sigma_vec)     \
            : call_family<P>(family, YARG, yi, xm, alpha_vec, beta_v, s0))           \
      : (sv ? call_family<P>(family, YARG, yi, xm, a0, beta_v, sigma_vec)            \
            : call_family<P>(family, YARG, yi, xm, a0, beta_v, s0)))
        if (propto) {
          res = y_is_var ? STANREF_CALL(true, y_v) : STANREF_CALL(true, yd);
        } else {
          res = y_is_var ? STANREF_CALL(false, y_v) : STANREF_CALL(false, yd);
        }
#undef STANREF_CALL
        res.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = res.val();
        read_adj(beta_v, dbeta);
        if (dalpha) {
          if (av) read_adj(alpha_vec, dalpha); else dalpha[0] = a0.adj();
        }
        if (dsigma && has_scale) {
          if (sv) read_adj(sigma_vec, dsigma); else dsigma[0] = s0.adj();
        }
        if (dy && y_is_var) read_adj(y_v, dy);
      },
      err, errlen);
}

// ordered_logistic_glm_lpmf(y, x, beta, cuts) with var beta and cuts: value + gradient.
int stanref_ordered_logistic_glm(int propto, long n, long k, const double* x, const int* y,
                                 const double* beta, const double* cuts, long ncuts, double* lp,
                                 double* dbeta, double* dcuts, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        matrix_d xm = Eigen::Map<const matrix_d>(x, n, k);
        std::vector<int> yi(y, y + n);
        vector_v beta_v = make_vars(beta, k);
        vector_v cuts_v = make_vars(cuts, ncuts);
        auto t0 = std::chrono::steady_clock::now();
        var res = propto ? stan::math::ordered_logistic_glm_lpmf<true>(yi, xm, beta_v, cuts_v)
                         : stan::math::ordered_logistic_glm_lpmf<false>(yi, xm, beta_v, cuts_v);
        res.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = res.val();
        read_adj(beta_v, dbeta);
        read_adj(cuts_v, dcuts);
      },
      err, errlen);
}

// categorical_logit_glm_lpmf(y, x, alpha, beta) with var alpha (C) and beta (k x C, column-major).
int stanref_categorical_logit_glm(int propto, long n, long k, long n_classes, const double* x, const int* y,
                                  const double* alpha, const double* beta, double* lp, double* dalpha,
                                  double* dbeta, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        matrix_d xm = Eigen::Map<const matrix_d>(x, n, k);
        std::vector<int> yi(y, y + n);
        vector_v alpha_v = make_vars(alpha, n_classes);
        Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic> beta_v(k, n_classes);
        for (long q = 0; q < k * n_classes; ++q) beta_v(q) = beta[q];
        auto t0 = std::chrono::steady_clock::now();
        var res = propto ? stan::math::categorical_logit_glm_lpmf<true>(yi, xm, alpha_v, beta_v)
                         : stan::math::categorical_logit_glm_lpmf<false>(yi, xm, alpha_v, beta_v);
        res.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = res.val();
        read_adj(alpha_v, dalpha);
        if (dbeta) for (long q = 0; q < k * n_classes; ++q) dbeta[q] = beta_v(q).adj();
      },
      err, errlen);
}

// Same call with every argument double (no autodiff): the value only.
int stanref_glm_value(int family, int propto, long n, long k, const double* x,
                      const void* y, const double* alpha, long alpha_len,
                      const double* beta, const double* sigma, long sigma_len,
                      double* lp, char* err, int errlen) {
  return guarded(
      [&]() {
        Eigen::Map<const matrix_d> xmap(x, n, k);
        matrix_d xm = xmap;
        vector_d b = Eigen::Map<const vector_d>(beta, k);
        vector_d a = Eigen::Map<const vector_d>(alpha, alpha_len);
        double r = 0;
        if (family == FAM_NORMAL_ID) {
          vector_d yd = Eigen::Map<const vector_d>(static_cast<const double*>(y), n);
          vector_d s = Eigen::Map<const vector_d>(sigma, sigma_len);
          if (alpha_len == 1 && sigma_len == 1)
            r = propto ? stan::math::normal_id_glm_lpdf<true>(yd, xm, a[0], b, s[0])
                       : stan::math::normal_id_glm_lpdf<false>(yd, xm, a[0], b, s[0]);
          else if (alpha_len == 1)
            r = propto ? stan::math::normal_id_glm_lpdf<true>(yd, xm, a[0], b, s)
                       : stan::math::normal_id_glm_lpdf<false>(yd, xm, a[0], b, s);
          else if (sigma_len == 1)
            r = propto ? stan::math::normal_id_glm_lpdf<true>(yd, xm, a, b, s[0])
                       : stan::math::normal_id_glm_lpdf<false>(yd, xm, a, b, s[0]);
          else
            r = propto ? stan::math::normal_id_glm_lpdf<true>(yd, xm, a, b, s)
                       : stan::math::normal_id_glm_lpdf<false>(yd, xm, a, b, s);
        } else {
          const int* p = static_cast<const int*>(y);
          std::vector<int> yi(p, p + n);
          if (family == FAM_BERNOULLI_LOGIT) {
            if (alpha_len == 1)
              r = propto ? stan::math::bernoulli_logit_glm_lpmf<true>(yi, xm, a[0], b)
                         : stan::math::bernoulli_logit_glm_lpmf<false>(yi, xm, a[0], b);
            else
              r = propto ? stan::math::bernoulli_logit_glm_lpmf<true>(yi, xm, a, b)
                         : stan::math::bernoulli_logit_glm_lpmf<false>(yi, xm, a, b);
          } else {
            if (alpha_len == 1)
              r = propto ? stan::math::poisson_log_glm_lpmf<true>(yi, xm, a[0], b)
                         : stan::math::poisson_log_glm_lpmf<false>(yi, xm, a[0], b);
            else
              r = propto ? stan::math::poisson_log_glm_lpmf<true>(yi, xm, a, b)
                         : stan::math::poisson_log_glm_lpmf<false>(yi, xm, a, b);
          }
        }
        if (lp) *lp = r;
      },
      err, errlen);
}

// Vectorised normal_lpdf(y | mu, sigma): y double data (or var when y_is_var),
// mu var vector (mu_len n) or var scalar (mu_len 1), sigma var scalar or var vector.
int stanref_normal_lpdf(int propto, long n, const double* y, const double* mu,
                        long mu_len, const double* sigma, long sigma_len,
                        int y_is_var, double* lp, double* dmu, double* dsigma,
                        double* dy, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        vector_d yd = Eigen::Map<const vector_d>(y, n);
        vector_v y_v = make_vars(y, y_is_var ? n : 0);
        vector_v mu_v = make_vars(mu, mu_len);
        vector_v sigma_v = make_vars(sigma, sigma_len);
        const bool mv = mu_len != 1, sv = sigma_len != 1;
        var m0 = mu_len >= 1 ? mu_v[0] : var(0.0);
        var s0 = sigma_len >= 1 ? sigma_v[0] : var(1.0);
        auto t0 = std::chrono::steady_clock::now();
        var res;
#define STANREF_NCALL(P, YARG)                                                  \
  (mv ? (sv ? stan::math::normal_lpdf<P>(YARG, mu_v, sigma_v)                   \
            : stan::math::normal_lpdf<P>(YARG, mu_v, s0))                       \
      : (sv ? stan::math::normal_lpdf<P>(YARG, m0, sigma_v)                     \
            : stan::math::normal_lpdf<P>(YARG, m0, s0)))
        if (propto) {
          res = y_is_var ? STANREF_NCALL(true, y_v) : STANREF_NCALL(true, yd);
        } else {
          res = y_is_var ? STANREF_NCALL(false, y_v) : STANREF_NCALL(false, yd);
        }
#undef STANREF_NCALL
        res.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = res.val();
        if (dmu) {
          if (mv) read_adj(mu_v, dmu); else dmu[0] = m0.adj();
        }
        if (dsigma) {
          if (sv) read_adj(sigma_v, dsigma); else dsigma[0] = s0.adj();
        }
        if (dy && y_is_var) read_adj(y_v, dy);
      },
      err, errlen);
}

// Vectorised bernoulli_logit_lpmf (density 1) / poisson_log_lpmf (2) / poisson_lpmf (3) / bernoulli_lpmf (4): n int data (n_len 1 or n),
// theta a var vector (theta_len n) or var scalar; theta_is_var == 0 evaluates the all-double form (value only).
int stanref_density(int density, int propto, long n, const int* y, long y_len, const double* theta,
                    long theta_len, int theta_is_var, double* lp, double* dtheta, double* seconds,
                    char* err, int errlen) {
  return guarded(
      [&]() {
        std::vector<int> yv(y, y + y_len);
        const bool ym = y_len != 1, tm = theta_len != 1;
        auto t0 = std::chrono::steady_clock::now();
        if (!theta_is_var) {
          vector_d th = Eigen::Map<const vector_d>(theta, theta_len);
          double r;
#define STANREF_DCALL(F, P) (ym ? (tm ? F<P>(yv, th) : F<P>(yv, th[0])) : (tm ? F<P>(yv[0], th) : F<P>(yv[0], th[0])))
          if (density == 1)
            r = propto ? STANREF_DCALL(stan::math::bernoulli_logit_lpmf, true) : STANREF_DCALL(stan::math::bernoulli_logit_lpmf, false);
          else if (density == 3)
            r = propto ? STANREF_DCALL(stan::math::poisson_lpmf, true) : STANREF_DCALL(stan::math::poisson_lpmf, false);
          else if (density == 4)
            r = propto ? STANREF_DCALL(stan::math::bernoulli_lpmf, true) : STANREF_DCALL(stan::math::bernoulli_lpmf, false);
          else
            r = propto ? STANREF_DCALL(stan::math::poisson_log_lpmf, true) : STANREF_DCALL(stan::math::poisson_log_lpmf, false);
          if (lp) *lp = r;
          return;
        }
        vector_v th = make_vars(theta, theta_len);
        var t0v = th[0];
        var res;
#define STANREF_DVCALL(F, P) (ym ? (tm ? F<P>(yv, th) : F<P>(yv, t0v)) : (tm ? F<P>(yv[0], th) : F<P>(yv[0], t0v)))
        if (density == 1)
          res = propto ? STANREF_DVCALL(stan::math::bernoulli_logit_lpmf, true) : STANREF_DVCALL(stan::math::bernoulli_logit_lpmf, false);
        else if (density == 3)
          res = propto ? STANREF_DVCALL(stan::math::poisson_lpmf, true) : STANREF_DVCALL(stan::math::poisson_lpmf, false);
        else if (density == 4)
          res = propto ? STANREF_DVCALL(stan::math::bernoulli_lpmf, true) : STANREF_DVCALL(stan::math::bernoulli_lpmf, false);
        else
          res = propto ? STANREF_DVCALL(stan::math::poisson_log_lpmf, true) : STANREF_DVCALL(stan::math::poisson_log_lpmf, false);
#undef STANREF_DCALL
#undef STANREF_DVCALL
        res.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = res.val();
        if (dtheta) {
          if (tm) read_adj(th, dtheta); else dtheta[0] = t0v.adj();
        }
      },
      err, errlen);
}

// All-double normal_lpdf (value only; the reference's known-answer tests use this form,
// test/prob/normal/normal_test.hpp:11-40).
int stanref_normal_lpdf_value(int propto, long n, const double* y, const double* mu,
                              long mu_len, const double* sigma, long sigma_len,
                              double* lp, char* err, int errlen) {
  return guarded(
      [&]() {
        vector_d yd = Eigen::Map<const vector_d>(y, n);
        vector_d m = Eigen::Map<const vector_d>(mu, mu_len);
        vector_d s = Eigen::Map<const vector_d>(sigma, sigma_len);
        double r;
        if (mu_len != 1 && sigma_len != 1)
          r = propto ? stan::math::normal_lpdf<true>(yd, m, s) : stan::math::normal_lpdf<false>(yd, m, s);
        else if (mu_len != 1)
          r = propto ? stan::math::normal_lpdf<true>(yd, m, s[0]) : stan::math::normal_lpdf<false>(yd, m, s[0]);
        else if (sigma_len != 1)
          r = propto ? stan::math::normal_lpdf<true>(yd, m[0], s) : stan::math::normal_lpdf<false>(yd, m[0], s);
        else
          r = propto ? stan::math::normal_lpdf<true>(yd, m[0], s[0]) : stan::math::normal_lpdf<false>(yd, m[0], s[0]);
        if (lp) *lp = r;
      },
      err, errlen);
}

}  // extern "C"

// ---------------------------------------------------------------------------------
// map_rect baselines (stan/math/rev/mat/functor/map_rect_concurrent.hpp:21-71 under
// -DSTAN_THREADS => tbb::parallel_for over jobs).
// ---------------------------------------------------------------------------------
namespace {

// Row-shard functor: one job = one contiguous block of rows of a GLM.
//   shared = [beta(k), alpha, (sigma)]   job params: empty
//   x_r[j] = shard of X, column-major rows_j x k  (+ y doubles for normal_id)
//   x_i[j] = [family, rows_j, k, y ints...]
struct glm_shard_functor {
  template <typename T1, typename T2>
  Eigen::Matrix<typename stan::return_type<T1, T2>::type, Eigen::Dynamic, 1>
  operator()(const Eigen::Matrix<T1, Eigen::Dynamic, 1>& shared,
             const Eigen::Matrix<T2, Eigen::Dynamic, 1>& job,
             const std::vector<double>& x_r, const std::vector<int>& x_i,
             std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int family = x_i[0], rows = x_i[1], k = x_i[2];
    // the functor has to materialise an Eigen matrix: value_of_rec only takes Eigen::Matrix
    matrix_d x = Eigen::Map<const matrix_d>(x_r.data(), rows, k);
    Eigen::Matrix<T1, Eigen::Dynamic, 1> beta = shared.head(k);
    // hierarchical form (config 5): per-job intercept in job(0), else shared(k)
    T alpha = job.size() > 0 ? T(job(0)) : T(shared(k));
    Eigen::Matrix<T, Eigen::Dynamic, 1> res(1);
    if (family == FAM_NORMAL_ID) {
      vector_d y = Eigen::Map<const vector_d>(x_r.data() + static_cast<size_t>(rows) * k, rows);
      T1 sigma = shared(k + 1);
      res(0) = stan::math::normal_id_glm_lpdf(y, x, alpha, beta, sigma);
    } else {
      std::vector<int> y(x_i.begin() + 3, x_i.begin() + 3 + rows);
      if (family == FAM_BERNOULLI_LOGIT)
        res(0) = stan::math::bernoulli_logit_glm_lpmf(y, x, alpha, beta);
      else
        res(0) = stan::math::poisson_log_glm_lpmf(y, x, alpha, beta);
    }
    return res;
  }
};

struct shard_set {
  int family;
  long k;
  std::vector<std::vector<double>> x_r;
  std::vector<std::vector<int>> x_i;
};

bool tbb_ready = false;

}  // namespace

extern "C" {

// STAN_NUM_THREADS is only honoured if init_threadpool_tbb() runs before TBB is first
// used (stan/math/prim/core/init_threadpool_tbb.hpp:83-90).  n_threads: -1 = all cores.
// Returns the number of threads in effect, or <0 on error.
int stanref_init_threads(int n_threads) {
  try {
    if (!tbb_ready) {
      std::string s = std::to_string(n_threads);
      setenv("STAN_NUM_THREADS", s.c_str(), 1);
      stan::math::init_threadpool_tbb();
      tbb_ready = true;
    }
    return stan::math::internal::get_num_threads();
  } catch (const std::exception&) {
    return -1;
  }
}

// Build the job data for a row-sharded GLM once (the reference treats x_r/x_i as
// immutable between evaluations, map_rect.hpp:97-100).  X column-major n x k.
void* stanref_shards_create(int family, long n, long k, const double* x, const void* y,
                            int n_shards) {
  shard_set* s = new shard_set();
  s->family = family;
  s->k = k;
  s->x_r.resize(n_shards);
  s->x_i.resize(n_shards);
  // map_rect insists on equal-sized jobs (map_rect.hpp:139-163): n must divide evenly
  const long rows = n / n_shards;
  for (int j = 0; j < n_shards; ++j) {
    const long r0 = j * rows;
    std::vector<double>& xr = s->x_r[j];
    xr.resize(static_cast<size_t>(rows) * k + (family == FAM_NORMAL_ID ? rows : 0));
    for (long c = 0; c < k; ++c)
      std::memcpy(&xr[static_cast<size_t>(c) * rows], x + static_cast<size_t>(c) * n + r0,
                  rows * sizeof(double));
    std::vector<int>& xi = s->x_i[j];
    xi.assign(3, 0);
    xi[0] = family; xi[1] = static_cast<int>(rows); xi[2] = static_cast<int>(k);
    if (family == FAM_NORMAL_ID) {
      std::memcpy(&xr[static_cast<size_t>(rows) * k], static_cast<const double*>(y) + r0,
                  rows * sizeof(double));
    } else {
      const int* yi = static_cast<const int*>(y);
      xi.insert(xi.end(), yi + r0, yi + r0 + rows);
    }
  }
  return s;
}

void stanref_shards_free(void* h) { delete static_cast<shard_set*>(h); }

// One evaluation: sum over shards of the GLM log-density, value + gradient, through
// map_rect (TBB threads).  job_alpha == NULL: shared scalar alpha (row-sharded GLM);
// otherwise one intercept per shard (hierarchical model, config 5) and dalpha has
// n_shards entries.
int stanref_shards_eval(void* h, const double* beta, const double* alpha,
                        const double* job_alpha, const double* sigma, double* lp,
                        double* dbeta, double* dalpha, double* dsigma, double* seconds,
                        char* err, int errlen) {
  shard_set* s = static_cast<shard_set*>(h);
  return guarded(
      [&]() {
        const long k = s->k;
        const int J = static_cast<int>(s->x_r.size());
        const bool normal = s->family == FAM_NORMAL_ID;
        vector_v shared(k + 1 + (normal ? 1 : 0));
        for (long i = 0; i < k; ++i) shared[i] = beta[i];
        shared[k] = alpha ? alpha[0] : 0.0;
        if (normal) shared[k + 1] = sigma[0];
        std::vector<vector_v> job(J, vector_v(job_alpha ? 1 : 0));
        if (job_alpha)
          for (int j = 0; j < J; ++j) job[j][0] = job_alpha[j];
        auto t0 = std::chrono::steady_clock::now();
        vector_v out = stan::math::map_rect<1, glm_shard_functor>(shared, job, s->x_r, s->x_i);
        var total = stan::math::sum(out);
        total.grad();
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        if (lp) *lp = total.val();
        if (dbeta) for (long i = 0; i < k; ++i) dbeta[i] = shared[i].adj();
        if (dalpha) {
          if (job_alpha) for (int j = 0; j < J; ++j) dalpha[j] = job[j][0].adj();
          else dalpha[0] = shared[k].adj();
        }
        if (dsigma && normal) dsigma[0] = shared[k + 1].adj();
      },
      err, errlen);
}

const char* stanref_version(void) {
  return "Stan Math " STAN_STRING(STAN_MATH_MAJOR) "." STAN_STRING(STAN_MATH_MINOR) "." STAN_STRING(STAN_MATH_PATCH)
#ifdef STAN_THREADS
         " STAN_THREADS"
#endif
      ;
}

}  // extern "C"
