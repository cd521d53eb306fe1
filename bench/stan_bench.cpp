// bench/stan_bench.cpp -- the five BASELINE configs through the reference's OWN C++ signatures.
//
// One binary, built with the overlay include path and -DSTAN_CUDA -DSTAN_THREADS (test/Makefile):
//   * stan::math::<fn>(...)                with var arguments  -> the B200 path (this repo)
//   * stan::math::<fn>_reference_cpu(...)  the unmodified reference body (same header, renamed)
//   * stan::math::map_rect<id, F>          -> shards dispatched to devices;
//     internal::map_rect_concurrent_reference_cpu<id, F> -> the reference's TBB job loop
// One evaluation = forward call returning a var + grad() + reading every adjoint, in a
// nested tape that is recovered afterwards; var construction is outside the timed region on both
// sides.  X and y are cached on the device by the first (untimed) GPU evaluation.
//
//   stan_bench --config c1|c2|c3|c4|c5|d1|d2|d3|d4|l1|l2|l3|l4|l5|l6|s1|m1|m2|k1|k2|v2 [--n ROWS] [--reps R] [--cpu-reps R] [--threads T] [--no-cpu]
// prints one JSON line: GPU evals/s (median), CPU single-thread and TBB map_rect evals/s, parity.
#include <stan/math.hpp>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::MatrixXd;
using Eigen::VectorXd;
using stan::math::var;
typedef Matrix<var, Dynamic, 1> vector_v;

namespace {

double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- fast deterministic generator (splitmix64 + Box-Muller), one stream per column ------------
struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double uni() { return (double(next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
  double normal() {
    const double u = uni(), v = uni();
    return std::sqrt(-2.0 * std::log(u)) * std::cos(6.283185307179586 * v);
  }
  int poisson(double lam) {  // Knuth for small lambda, normal approximation above 30
    if (lam > 30) {
      const double v = lam + std::sqrt(lam) * normal();
      return v < 0 ? 0 : int(v + 0.5);
    }
    const double L = std::exp(-lam);
    int k = 0;
    double p = 1.0;
    do {
      ++k;
      p *= uni();
    } while (p > L);
    return k - 1;
  }
};

template <typename F>
void parallel_chunks(int64_t n, int threads, F&& f) {
  std::vector<std::thread> ts;
  const int64_t per = (n + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    const int64_t a = t * per, b = std::min<int64_t>(n, a + per);
    if (a >= b) break;
    ts.emplace_back([=, &f] { f(a, b, t); });
  }
  for (auto& t : ts) t.join();
}

struct Problem {
  int family;  // 0 bernoulli, 1 poisson, 2 normal
  int64_t n, k;
  MatrixXd x;
  std::vector<int> y_int;
  VectorXd y_real;
  VectorXd beta;
  double alpha, sigma;
};

void make_problem(Problem* p, int family, int64_t n, int64_t k, uint64_t seed, int threads) {
  p->family = family;
  p->n = n;
  p->k = k;
  p->x.resize(n, k);
  parallel_chunks(k, threads, [&](int64_t a, int64_t b, int) {
    for (int64_t j = a; j < b; ++j) {
      Rng r(seed * 1000003ull + j);
      double* col = p->x.data() + j * n;
      for (int64_t i = 0; i < n; ++i) col[i] = r.normal();
    }
  });
  Rng rb(seed + 99);
  const double sd = k <= 32 ? 0.1 : 0.05;
  VectorXd beta_true(k);
  for (int64_t j = 0; j < k; ++j) beta_true[j] = sd * rb.normal();
  p->beta = beta_true;
  for (int64_t j = 0; j < k; ++j) p->beta[j] += 0.02 * rb.normal();
  p->alpha = 0.3;   // truth 0.2
  p->sigma = 1.0;   // truth 0.7
  if (family == 2) p->y_real.resize(n); else p->y_int.resize(n);
  parallel_chunks(n, threads, [&](int64_t a, int64_t b, int t) {
    Rng r(seed * 7919ull + 13 * t + 5);
    for (int64_t i = a; i < b; ++i) {
      double eta = 0.2;
      for (int64_t j = 0; j < k; ++j) eta += p->x.data()[j * n + i] * beta_true[j];
      if (family == 0) p->y_int[i] = r.uni() < 1.0 / (1.0 + std::exp(-eta)) ? 1 : 0;
      else if (family == 1) p->y_int[i] = r.poisson(std::exp(eta));
      else p->y_real[i] = eta + 0.7 * r.normal();
    }
  });
}

struct Result {
  double lp = 0;
  VectorXd grad;  // [dbeta..., dalpha, (dsigma)] or config-specific
};

double rel_err(const Result& a, const Result& b) {
  double e = std::fabs(a.lp - b.lp) / std::max(std::fabs(b.lp), 1e-300);
  const double scale = b.grad.size() ? b.grad.cwiseAbs().maxCoeff() : 0.0;
  for (int i = 0; i < b.grad.size(); ++i) {
    const double den = std::max(std::fabs(b.grad[i]), 1e-6 * scale);
    if (den > 0) e = std::max(e, std::fabs(a.grad[i] - b.grad[i]) / den);
  }
  return e;
}

// One GLM evaluation through the public signature (GPU = true) or the reference body.
template <bool GPU>
double glm_eval_once(const Problem& p, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v beta(p.k);
  for (int64_t j = 0; j < p.k; ++j) beta[j] = p.beta[j];
  var alpha = p.alpha, sigma = p.sigma;
  const double t0 = now();
  var lp;
  if (p.family == 0) {
    lp = GPU ? bernoulli_logit_glm_lpmf(p.y_int, p.x, alpha, beta)
             : bernoulli_logit_glm_lpmf_reference_cpu<false>(p.y_int, p.x, alpha, beta);
  } else if (p.family == 1) {
    lp = GPU ? poisson_log_glm_lpmf(p.y_int, p.x, alpha, beta)
             : poisson_log_glm_lpmf_reference_cpu<false>(p.y_int, p.x, alpha, beta);
  } else {
    lp = GPU ? normal_id_glm_lpdf(p.y_real, p.x, alpha, beta, sigma)
             : normal_id_glm_lpdf_reference_cpu<false>(p.y_real, p.x, alpha, beta, sigma);
  }
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(p.k + 2);
  for (int64_t j = 0; j < p.k; ++j) out->grad[j] = beta[j].adj();
  out->grad[p.k] = alpha.adj();
  out->grad[p.k + 1] = p.family == 2 ? sigma.adj() : 0.0;
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// the per-row variant -- bernoulli_logit_glm_lpmf with the intercept a var VECTOR of length N
// (varying intercepts); one evaluation = forward + grad() + reading the K + N adjoints
template <bool GPU>
double glm_vector_alpha_once(const Problem& p, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v beta(p.k), alpha(p.n);
  for (int64_t j = 0; j < p.k; ++j) beta[j] = p.beta[j];
  for (int64_t i = 0; i < p.n; ++i) alpha[i] = p.alpha + 0.001 * double(i % 97);
  const double t0 = now();
  var lp = GPU ? bernoulli_logit_glm_lpmf(p.y_int, p.x, alpha, beta)
               : bernoulli_logit_glm_lpmf_reference_cpu<false>(p.y_int, p.x, alpha, beta);
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(p.k + p.n);
  for (int64_t j = 0; j < p.k; ++j) out->grad[j] = beta[j].adj();
  for (int64_t i = 0; i < p.n; ++i) out->grad[p.k + i] = alpha[i].adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// ---- map_rect functors ----------------------------------------------------------------------------
// shared = beta (K) [ + alpha, sigma when SHARED_ALPHA ]; job = alpha_j (1) or empty;
// x_i = [rows, k, y...] (int outcomes) ; x_r = X col-major rows*k [ + y (normal) ]
template <bool GPU, int FAMILY, bool SHARED_ALPHA>
struct shard_functor {
  template <typename T1, typename T2>
  Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(
      const Matrix<T1, Dynamic, 1>& shared, const Matrix<T2, Dynamic, 1>& job, const std::vector<double>& x_r,
      const std::vector<int>& x_i, std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    using namespace stan::math;  // NOLINT
    const int rows = x_i[0], k = x_i[1];
    // what a Stan functor does: to_matrix(x_r, rows, k) -- a fresh copy on every call
    MatrixXd x = Eigen::Map<const MatrixXd>(x_r.data(), rows, k);
    Matrix<T1, Dynamic, 1> beta = shared.head(k);
    Matrix<T, Dynamic, 1> res(1);
    if (FAMILY == 2) {
      VectorXd y = Eigen::Map<const VectorXd>(x_r.data() + size_t(rows) * k, rows);
      T1 alpha = shared[k], sigma = shared[k + 1];
      res(0) = GPU ? normal_id_glm_lpdf(y, x, alpha, beta, sigma)
                   : normal_id_glm_lpdf_reference_cpu<false>(y, x, alpha, beta, sigma);
      return res;
    }
    std::vector<int> y(x_i.begin() + 2, x_i.end());
    if (SHARED_ALPHA) {
      T1 alpha = shared[k];
      if (FAMILY == 0)
        res(0) = GPU ? bernoulli_logit_glm_lpmf(y, x, alpha, beta)
                     : bernoulli_logit_glm_lpmf_reference_cpu<false>(y, x, alpha, beta);
      else
        res(0) = GPU ? poisson_log_glm_lpmf(y, x, alpha, beta)
                     : poisson_log_glm_lpmf_reference_cpu<false>(y, x, alpha, beta);
    } else {
      T2 alpha = job[0];
      if (FAMILY == 0)
        res(0) = GPU ? bernoulli_logit_glm_lpmf(y, x, alpha, beta)
                     : bernoulli_logit_glm_lpmf_reference_cpu<false>(y, x, alpha, beta);
      else
        res(0) = GPU ? poisson_log_glm_lpmf(y, x, alpha, beta)
                     : poisson_log_glm_lpmf_reference_cpu<false>(y, x, alpha, beta);
    }
    return res;
  }
};

// The same shard with x passed as a zero-copy view of x_r (cuda::data_matrix_view, an extension of
// the CUDA subsystem): no per-call copy of the design matrix on the host.
struct shard_functor_view {
  template <typename T1, typename T2>
  Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(
      const Matrix<T1, Dynamic, 1>& shared, const Matrix<T2, Dynamic, 1>& job, const std::vector<double>& x_r,
      const std::vector<int>& x_i, std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int rows = x_i[0], k = x_i[1];
    stan::math::cuda::data_matrix_view x(x_r.data(), rows, k);
    Eigen::Map<const Matrix<int, Dynamic, 1>> y_map(x_i.data() + 2, rows);
    Matrix<int, Dynamic, 1> y = y_map;
    Matrix<T, Dynamic, 1> res(1);
    T2 alpha = job[0];
    res(0) = stan::math::bernoulli_logit_glm_lpmf(y, x, alpha, shared);
    return res;
  }
};

// What the unmodified functor does BEFORE it calls the GLM, and nothing else: the copy of x_r into a matrix and of the
// outcomes into a vector.  Timed through the same map_rect to show how much of `gpu_map_rect` is this user code.
struct shard_functor_copy_only {
  template <typename T1, typename T2>
  Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(
      const Matrix<T1, Dynamic, 1>& shared, const Matrix<T2, Dynamic, 1>& job, const std::vector<double>& x_r,
      const std::vector<int>& x_i, std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int rows = x_i[0], k = x_i[1];
    MatrixXd x = Eigen::Map<const MatrixXd>(x_r.data(), rows, k);
    std::vector<int> y(x_i.begin() + 2, x_i.end());
    Matrix<T, Dynamic, 1> res(1);
    res(0) = job[0] * (x(rows - 1, k - 1) + double(y[rows - 1])) + shared[0] * x(0, 0);  // keeps both copies alive
    return res;
  }
};

struct ShardSet {
  std::vector<std::vector<double>> x_r;
  std::vector<std::vector<int>> x_i;
  int J = 0;
};

// rows [0, J*rows) of p split into J equal jobs (map_rect needs equal-sized jobs, map_rect.hpp:139-163)
void make_shards(const Problem& p, int J, ShardSet* s) {
  const int64_t rows = p.n / J;
  s->J = J;
  s->x_r.resize(J);
  s->x_i.resize(J);
  parallel_chunks(J, std::min(J, 16), [&](int64_t a, int64_t b, int) {
    for (int64_t j = a; j < b; ++j) {
      std::vector<double>& xr = s->x_r[j];
      xr.resize(size_t(rows) * p.k + (p.family == 2 ? rows : 0));
      for (int64_t c = 0; c < p.k; ++c)
        std::memcpy(xr.data() + c * rows, p.x.data() + c * p.n + j * rows, size_t(rows) * sizeof(double));
      std::vector<int>& xi = s->x_i[j];
      xi.assign(2, 0);
      xi[0] = int(rows);
      xi[1] = int(p.k);
      if (p.family == 2) {
        std::memcpy(xr.data() + size_t(rows) * p.k, p.y_real.data() + j * rows, size_t(rows) * sizeof(double));
      } else {
        xi.insert(xi.end(), p.y_int.begin() + j * rows, p.y_int.begin() + (j + 1) * rows);
      }
    }
  });
}

// map_rect evaluation.  MODE 0: this repo's path through stan::math::map_rect (device dispatch);
// MODE 1: the reference's TBB loop with the reference bodies.
template <int MODE, int FAMILY, bool SHARED_ALPHA, int CALL_ID>
double map_rect_once(const Problem& p, const ShardSet& s, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  const int64_t k = p.k;
  vector_v shared(SHARED_ALPHA ? k + 2 : k);
  for (int64_t j = 0; j < k; ++j) shared[j] = p.beta[j];
  if (SHARED_ALPHA) {
    shared[k] = p.alpha;
    shared[k + 1] = p.sigma;
  }
  std::vector<vector_v> job(s.J, vector_v(SHARED_ALPHA ? 0 : 1));
  if (!SHARED_ALPHA)
    for (int j = 0; j < s.J; ++j) job[j][0] = 0.01 * j;
  const double t0 = now();
  vector_v res;
  if (MODE == 0) {
    res = map_rect<CALL_ID, shard_functor<true, FAMILY, SHARED_ALPHA>>(shared, job, s.x_r, s.x_i);
  } else if (MODE == 2) {
    res = map_rect<CALL_ID + 200, shard_functor_view>(shared, job, s.x_r, s.x_i);
  } else if (MODE == 3) {
    res = map_rect<CALL_ID + 300, shard_functor_copy_only>(shared, job, s.x_r, s.x_i);
  } else {
    res = internal::map_rect_concurrent_reference_cpu<CALL_ID + 100, shard_functor<false, FAMILY, SHARED_ALPHA>>(
        shared, job, s.x_r, s.x_i, nullptr);
  }
  var total = sum(res);
  grad(total.vi_);
  out->lp = total.val();
  out->grad.resize(shared.size() + (SHARED_ALPHA ? 0 : s.J));
  for (int64_t j = 0; j < shared.size(); ++j) out->grad[j] = shared[j].adj();
  if (!SHARED_ALPHA)
    for (int j = 0; j < s.J; ++j) out->grad[shared.size() + j] = job[j][0].adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

template <bool GPU>
double normal_once(const VectorXd& y, const VectorXd& mu_val, double sigma_val, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v mu(mu_val.size());
  for (int64_t i = 0; i < mu_val.size(); ++i) mu[i] = mu_val[i];
  var sigma = sigma_val;
  const double t0 = now();
  var lp = GPU ? normal_lpdf(y, mu, sigma) : normal_lpdf_reference_cpu<false>(y, mu, sigma);
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(mu.size() + 1);
  for (int64_t i = 0; i < mu.size(); ++i) out->grad[i] = mu[i].adj();
  out->grad[mu.size()] = sigma.adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// cauchy_lpdf (DENSITY 1) / lognormal_lpdf (2) as a model states them -- data y, var vector mu, var sigma
template <bool GPU, int DENSITY>
double locscale_once(const VectorXd& y, const VectorXd& mu_val, double sigma_val, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v mu(mu_val.size());
  for (int64_t i = 0; i < mu_val.size(); ++i) mu[i] = mu_val[i];
  var sigma = sigma_val;
  const double t0 = now();
  // (3: gamma_lpdf(y | alpha = the scalar, beta = the vector), 4: exponential_lpdf(y | beta = the vector))
  var lp = DENSITY == 1   ? (GPU ? cauchy_lpdf(y, mu, sigma) : cauchy_lpdf_reference_cpu<false>(y, mu, sigma))
           : DENSITY == 2 ? (GPU ? lognormal_lpdf(y, mu, sigma) : lognormal_lpdf_reference_cpu<false>(y, mu, sigma))
           : DENSITY == 3 ? (GPU ? gamma_lpdf(y, sigma, mu) : gamma_lpdf_reference_cpu<false>(y, sigma, mu))
           : DENSITY == 5 ? (GPU ? beta_lpdf(y, sigma, mu) : beta_lpdf_reference_cpu<false>(y, sigma, mu))
           : DENSITY == 6 ? (GPU ? weibull_lpdf(y, sigma, mu) : weibull_lpdf_reference_cpu<false>(y, sigma, mu))
                          : (GPU ? exponential_lpdf(y, mu) + 0.0 * sigma : exponential_lpdf_reference_cpu<false>(y, mu) + 0.0 * sigma);
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(mu.size() + 1);
  for (int64_t i = 0; i < mu.size(); ++i) out->grad[i] = mu[i].adj();
  out->grad[mu.size()] = sigma.adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// student_t_lpdf as a model states it -- data y, var nu, var vector mu, var sigma; one evaluation =
// forward + grad() + reading the N + 2 adjoints
template <bool GPU>
double student_t_once(const VectorXd& y, double nu_val, const VectorXd& mu_val, double sigma_val, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v mu(mu_val.size());
  for (int64_t i = 0; i < mu_val.size(); ++i) mu[i] = mu_val[i];
  var nu = nu_val, sigma = sigma_val;
  const double t0 = now();
  var lp = GPU ? student_t_lpdf(y, nu, mu, sigma) : student_t_lpdf_reference_cpu<false>(y, nu, mu, sigma);
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(mu.size() + 2);
  for (int64_t i = 0; i < mu.size(); ++i) out->grad[i] = mu[i].adj();
  out->grad[mu.size()] = sigma.adj();
  out->grad[mu.size() + 1] = nu.adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// dense multiply.  prim: C = multiply(A, B) for double matrices; rev: AB = multiply(A_v, B_v) with
// both operands matrices of var, f = sum(AB), grad() and the adjoints of A and B read back (forward + the two
// products of chain()).
template <bool GPU>
double multiply_prim_once(const MatrixXd& A, const MatrixXd& B, Result* out) {
  const double t0 = now();
  MatrixXd C = GPU ? stan::math::multiply(A, B) : stan::math::multiply_reference_cpu(A, B);
  const double t1 = now();
  out->lp = C.sum();
  out->grad = Eigen::Map<VectorXd>(C.data(), std::min<int64_t>(C.size(), 4096));
  return t1 - t0;
}

template <bool GPU>
double multiply_rev_once(const MatrixXd& A, const MatrixXd& B, Result* out) {
  using namespace stan::math;  // NOLINT
  typedef Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic> matrix_v;
  start_nested();
  matrix_v Av(A.rows(), A.cols()), Bv(B.rows(), B.cols());
  for (int64_t i = 0; i < A.size(); ++i) Av(i) = A(i);
  for (int64_t i = 0; i < B.size(); ++i) Bv(i) = B(i);
  const double t0 = now();
  matrix_v AB = GPU ? multiply(Av, Bv) : multiply_reference_cpu(Av, Bv);
  var f = sum(AB);
  grad(f.vi_);
  out->lp = f.val();
  out->grad.resize(A.size() + B.size());
  for (int64_t i = 0; i < A.size(); ++i) out->grad[i] = Av(i).adj();
  for (int64_t i = 0; i < B.size(); ++i) out->grad[A.size() + i] = Bv(i).adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// cholesky_decompose.  prim: L = cholesky_decompose(A) for a double matrix; rev: A a matrix of var,
// f = sum(L), grad() and the adjoints of the lower triangle of A read back (the factorisation + the reverse sweep).
template <bool GPU>
double cholesky_prim_once(const MatrixXd& A, Result* out) {
  const double t0 = now();
  MatrixXd L = GPU ? stan::math::cholesky_decompose(A) : stan::math::cholesky_decompose_reference_cpu(A);
  const double t1 = now();
  out->lp = L.sum();
  out->grad = Eigen::Map<VectorXd>(L.data(), std::min<int64_t>(L.size(), 4096));
  return t1 - t0;
}

template <bool GPU>
double cholesky_rev_once(const MatrixXd& A, Result* out) {
  using namespace stan::math;  // NOLINT
  typedef Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic> matrix_v;
  start_nested();
  matrix_v Av(A.rows(), A.cols());
  for (int64_t i = 0; i < A.size(); ++i) Av(i) = A(i);
  const double t0 = now();
  matrix_v L = GPU ? cholesky_decompose(Av) : cholesky_decompose_reference_cpu(Av);
  var f = sum(L);
  grad(f.vi_);
  out->lp = f.val();
  const int64_t n = A.rows();
  out->grad.resize(n * (n + 1) / 2);
  int64_t pos = 0;
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = j; i < n; ++i) out->grad[pos++] = Av(i, j).adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

// vectorised lpmf with an integer outcome and a var vector operand (DENSITY 1: bernoulli_logit_lpmf,
// 2: poisson_log_lpmf, 3: poisson_lpmf, 4: bernoulli_lpmf); one evaluation = forward + grad() + reading the N adjoints
template <bool GPU, int DENSITY>
double int_density_once(const std::vector<int>& y, const VectorXd& th_val, Result* out) {
  using namespace stan::math;  // NOLINT
  start_nested();
  vector_v th(th_val.size());
  for (int64_t i = 0; i < th_val.size(); ++i) th[i] = th_val[i];
  const double t0 = now();
  var lp = DENSITY == 1   ? (GPU ? bernoulli_logit_lpmf(y, th) : bernoulli_logit_lpmf_reference_cpu<false>(y, th))
           : DENSITY == 2 ? (GPU ? poisson_log_lpmf(y, th) : poisson_log_lpmf_reference_cpu<false>(y, th))
           : DENSITY == 3 ? (GPU ? poisson_lpmf(y, th) : poisson_lpmf_reference_cpu<false>(y, th))
                          : (GPU ? bernoulli_lpmf(y, th) : bernoulli_lpmf_reference_cpu<false>(y, th));
  grad(lp.vi_);
  out->lp = lp.val();
  out->grad.resize(th.size());
  for (int64_t i = 0; i < th.size(); ++i) out->grad[i] = th[i].adj();
  const double t1 = now();
  recover_memory_nested();
  return t1 - t0;
}

struct Timing {
  double median = 0, best = 0, p10 = 0, p90 = 0;
  int reps = 0;
};

template <typename F>
Timing time_reps(int warm, int reps, double budget_s, F&& f) {
  for (int i = 0; i < warm; ++i) f();
  std::vector<double> t;
  const double start = now();
  for (int i = 0; i < reps; ++i) {
    t.push_back(f());
    if (now() - start > budget_s && t.size() >= 2) break;
  }
  std::sort(t.begin(), t.end());
  Timing r;
  r.reps = int(t.size());
  r.best = t.front();
  r.median = t[t.size() / 2];
  r.p10 = t[size_t(0.1 * double(t.size() - 1) + 0.5)];
  r.p90 = t[size_t(0.9 * double(t.size() - 1) + 0.5)];
  return r;
}

void print_timing(const char* name, const Timing& t) {
  std::printf("\"%s\": {\"evals_per_s\": %.6g, \"median_ms\": %.6g, \"p10_ms\": %.6g, \"p90_ms\": %.6g, \"best_ms\": %.6g, "
              "\"reps\": %d}",
              name, 1.0 / t.median, t.median * 1e3, t.p10 * 1e3, t.p90 * 1e3, t.best * 1e3, t.reps);
}

}  // namespace

int run(int argc, char** argv);

int main(int argc, char** argv) {
  const int rc = run(argc, argv);
  // The reference's static ad_tape_observer is destroyed by exit() while TBB workers may still be
  // calling its on_scheduler_exit (stan/math/rev/core/init_chainablestack.hpp:48-59): leave without
  // running static destructors.
  std::fflush(stdout);
  std::fflush(stderr);
  std::_Exit(rc);
}

int run(int argc, char** argv) {
  std::string config = "c3";
  int64_t n_override = 0;
  int reps = 50, cpu_reps = 3, threads = int(std::thread::hardware_concurrency());
  bool no_cpu = false, cpu_only = false;  // --cpu-only: debugging aid for boxes without a GPU
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a == "--config" && i + 1 < argc) config = argv[++i];
    else if (a == "--n" && i + 1 < argc) n_override = std::atoll(argv[++i]);
    else if (a == "--reps" && i + 1 < argc) reps = std::atoi(argv[++i]);
    else if (a == "--cpu-reps" && i + 1 < argc) cpu_reps = std::atoi(argv[++i]);
    else if (a == "--threads" && i + 1 < argc) threads = std::atoi(argv[++i]);
    else if (a == "--no-cpu") no_cpu = true;
    else if (a == "--cpu-only") cpu_only = true;
  }
  if (threads < 1) threads = 1;
  // the reference only honours STAN_NUM_THREADS if init_threadpool_tbb() runs before TBB starts
  // (stan/math/prim/core/init_threadpool_tbb.hpp:83-90)
  static std::string env = "STAN_NUM_THREADS=" + std::to_string(threads);
  putenv(&env[0]);
  stan::math::init_threadpool_tbb();
  const int n_dev = cpu_only ? 0 : stan::math::cuda::device_count();
  if (n_dev < 1 && !cpu_only) {
    std::fprintf(stderr, "stan_bench: no CUDA device: %s\n", stan_cuda_last_error());
    return 2;
  }

  std::printf("{\"config\": \"%s\", \"host_threads\": %d, \"devices\": %d, ", config.c_str(), threads, n_dev);
  if (config == "c1") {
    const int64_t n = n_override ? n_override : 1000000;
    VectorXd y(n), mu(n);
    Rng r(1001);
    for (int64_t i = 0; i < n; ++i) {
      y[i] = r.normal();
      mu[i] = 0.1 * r.normal();
    }
    Result g, c;
    normal_once<true>(y, mu, 1.3, &g);
    Timing tg = time_reps(3, reps, 30, [&] { return normal_once<true>(y, mu, 1.3, &g); });
    std::printf("\"what\": \"normal_lpdf N=%lld, y data, mu var vector, sigma var\", ", (long long)n);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 5), 30, [&] { return normal_once<false>(y, mu, 1.3, &c); });
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_device_bytes\": %lld}\n", (long long)(3 * n * 8));
    return 0;
  }

  if (config.size() == 2 && config[0] == 'l' && config[1] >= '1' && config[1] <= '6') {  // cauchy, lognormal, gamma, exponential, beta, weibull
    const int which = config[1] - '0';
    const bool cauchy = which == 1;
    const int64_t n = n_override ? n_override : 1000000;
    VectorXd y(n), mu(n);
    Rng r(1040 + which);
    for (int64_t i = 0; i < n; ++i) {
      y[i] = cauchy ? 2.0 * r.normal() : std::exp(1.2 * r.normal());
      if (which == 5) y[i] = std::min(std::max(y[i] / (1.0 + y[i]), 1e-9), 1.0 - 1e-9);   // beta: outcomes in (0, 1)
      mu[i] = which >= 3 ? std::exp(0.3 * r.normal()) : 0.1 * r.normal();   // gamma / exponential / beta / weibull: positive
    }
    Result g, c;
    auto gpu = [&] {
      return which == 1 ? locscale_once<true, 1>(y, mu, 1.3, &g) : which == 2 ? locscale_once<true, 2>(y, mu, 1.3, &g)
             : which == 3 ? locscale_once<true, 3>(y, mu, 1.3, &g) : which == 4 ? locscale_once<true, 4>(y, mu, 1.3, &g)
             : which == 5 ? locscale_once<true, 5>(y, mu, 1.3, &g) : locscale_once<true, 6>(y, mu, 1.3, &g);
    };
    auto cpu = [&] {
      return which == 1 ? locscale_once<false, 1>(y, mu, 1.3, &c) : which == 2 ? locscale_once<false, 2>(y, mu, 1.3, &c)
             : which == 3 ? locscale_once<false, 3>(y, mu, 1.3, &c) : which == 4 ? locscale_once<false, 4>(y, mu, 1.3, &c)
             : which == 5 ? locscale_once<false, 5>(y, mu, 1.3, &c) : locscale_once<false, 6>(y, mu, 1.3, &c);
    };
    gpu();
    Timing tg = time_reps(3, reps, 30, gpu);
    static const char* names[7] = {"", "cauchy_lpdf N=%lld, y data, mu var vector, sigma var", "lognormal_lpdf N=%lld, y data, mu var vector, sigma var",
                                   "gamma_lpdf N=%lld, y data, alpha var, beta var vector", "exponential_lpdf N=%lld, y data, beta var vector",
                                   "beta_lpdf N=%lld, y data, alpha var, beta var vector", "weibull_lpdf N=%lld, y data, alpha var, sigma var vector"};
    std::printf("\"what\": \"");
    std::printf(names[which], (long long)n);
    std::printf("\", ");
    if (false) std::printf("%s N=%lld", cauchy ? "cauchy_lpdf" : "lognormal_lpdf", (long long)n);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 5), 30, cpu);
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_device_bytes\": %lld}\n", (long long)(3 * n * 8));
    return 0;
  }

  if (config == "m1" || config == "m2") {
    const bool rev = config == "m2";
    const int64_t n = n_override ? n_override : (rev ? 1024 : 2048);
    MatrixXd A(n, n), B(n, n);
    Rng r(1031);
    for (int64_t i = 0; i < n * n; ++i) {
      A(i) = r.normal();
      B(i) = r.normal();
    }
    Result g, c;
    auto gpu = [&] { return rev ? multiply_rev_once<true>(A, B, &g) : multiply_prim_once<true>(A, B, &g); };
    auto cpu = [&] { return rev ? multiply_rev_once<false>(A, B, &c) : multiply_prim_once<false>(A, B, &c); };
    gpu();
    Timing tg = time_reps(2, reps, 30, gpu);
    std::printf("\"what\": \"multiply %lld x %lld x %lld, %s\", ", (long long)n, (long long)n, (long long)n,
                rev ? "matrices of var: forward + grad() + reading the adjoints" : "matrices of double");
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(0, std::max(std::min(cpu_reps, 3), 2), 60, cpu);
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"flops\": %.6g}\n", (rev ? 6.0 : 2.0) * double(n) * double(n) * double(n));
    return 0;
  }

  if (config == "k1" || config == "k2") {
    const bool rev = config == "k2";
    const int64_t n = n_override ? n_override : 2048;
    MatrixXd X(n, n);
    Rng r(1041);
    for (int64_t i = 0; i < n * n; ++i) X(i) = r.normal();
    MatrixXd A = X * X.transpose() / double(n) + MatrixXd::Identity(n, n);
    A = (0.5 * (A + A.transpose())).eval();
    Result g, c;
    auto gpu = [&] { return rev ? cholesky_rev_once<true>(A, &g) : cholesky_prim_once<true>(A, &g); };
    auto cpu = [&] { return rev ? cholesky_rev_once<false>(A, &c) : cholesky_prim_once<false>(A, &c); };
    gpu();
    Timing tg = time_reps(2, reps, 30, gpu);
    std::printf("\"what\": \"cholesky_decompose n = %lld, %s\", ", (long long)n,
                rev ? "matrix of var: factorisation + grad() + reading the adjoints" : "matrix of double");
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(0, std::max(std::min(cpu_reps, 3), 2), 60, cpu);
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf("}\n");
    return 0;
  }

  if (config == "s1") {
    const int64_t n = n_override ? n_override : 1000000;
    VectorXd y(n), mu(n);
    Rng r(1021);
    for (int64_t i = 0; i < n; ++i) {
      y[i] = 1.5 * r.normal();
      mu[i] = 0.1 * r.normal();
    }
    Result g, c;
    student_t_once<true>(y, 4.0, mu, 1.3, &g);
    Timing tg = time_reps(3, reps, 30, [&] { return student_t_once<true>(y, 4.0, mu, 1.3, &g); });
    std::printf("\"what\": \"student_t_lpdf N=%lld, y data, nu var, mu var vector, sigma var\", ", (long long)n);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 5), 30, [&] { return student_t_once<false>(y, 4.0, mu, 1.3, &c); });
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_device_bytes\": %lld}\n", (long long)(3 * n * 8));
    return 0;
  }

  if (config == "d1" || config == "d2") {
    const bool bern = config == "d1";
    const int64_t n = n_override ? n_override : 1000000;
    std::vector<int> y(n);
    VectorXd th(n);
    Rng r(bern ? 1011 : 1012);
    for (int64_t i = 0; i < n; ++i) {
      th[i] = (bern ? 2.0 : 1.0) * r.normal() + (bern ? 0.0 : 0.5);
      y[i] = bern ? (r.normal() < th[i] ? 1 : 0) : int(std::max(0.0, std::floor(std::exp(th[i]) + r.normal())));
      th[i] += 0.1 * r.normal();
    }
    Result g, c;
    auto gpu = [&] { return bern ? int_density_once<true, 1>(y, th, &g) : int_density_once<true, 2>(y, th, &g); };
    auto cpu = [&] { return bern ? int_density_once<false, 1>(y, th, &c) : int_density_once<false, 2>(y, th, &c); };
    gpu();
    Timing tg = time_reps(3, reps, 30, gpu);
    std::printf("\"what\": \"%s N=%lld, n data, operand var vector\", ", bern ? "bernoulli_logit_lpmf" : "poisson_log_lpmf",
                (long long)n);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 5), 30, cpu);
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_device_bytes\": %lld}\n", (long long)(n * 20));
    return 0;
  }

  if (config == "d3" || config == "d4") {  // poisson_lpmf(n | lambda), bernoulli_lpmf(n | theta)
    const bool pois = config == "d3";
    const int64_t n = n_override ? n_override : 1000000;
    std::vector<int> y(n);
    VectorXd op(n);
    Rng r(pois ? 1013 : 1014);
    for (int64_t i = 0; i < n; ++i) {
      const double z = r.normal();
      if (pois) {
        op[i] = std::exp(0.5 + z);
        y[i] = int(std::max(0.0, std::floor(op[i] + r.normal())));
        op[i] *= std::exp(0.1 * r.normal());
      } else {
        op[i] = 1.0 / (1.0 + std::exp(-2.0 * z));
        y[i] = r.normal() < 2.0 * z ? 1 : 0;
        op[i] = std::min(std::max(op[i], 1e-6), 1.0 - 1e-6);
      }
    }
    Result g, c;
    auto gpu = [&] { return pois ? int_density_once<true, 3>(y, op, &g) : int_density_once<true, 4>(y, op, &g); };
    auto cpu = [&] { return pois ? int_density_once<false, 3>(y, op, &c) : int_density_once<false, 4>(y, op, &c); };
    gpu();
    Timing tg = time_reps(3, reps, 30, gpu);
    std::printf("\"what\": \"%s N=%lld, n data, operand var vector\", ", pois ? "poisson_lpmf" : "bernoulli_lpmf", (long long)n);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 5), 30, cpu);
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_device_bytes\": %lld}\n", (long long)(n * 20));
    return 0;
  }

  if (config == "c5") {
    const int J = 64;
    const int64_t rows = n_override ? n_override : 100000, k = 32;
    Problem p;
    make_problem(&p, 0, J * rows, k, 1005, threads);
    ShardSet s;
    make_shards(p, J, &s);
    Result g, c;
    std::printf("\"what\": \"map_rect J=%d shards x %lld rows x K=%lld bernoulli_logit_glm, shared beta, job alpha\", ",
                J, (long long)rows, (long long)k);
    if (!cpu_only) {
      map_rect_once<0, 0, false, 5>(p, s, &g);  // cache fill
      Timing tg = time_reps(3, reps, 60, [&] { return map_rect_once<0, 0, false, 5>(p, s, &g); });
      print_timing("gpu_map_rect", tg);
      Result gv;
      map_rect_once<2, 0, false, 5>(p, s, &gv);
      Timing tv = time_reps(3, reps, 60, [&] { return map_rect_once<2, 0, false, 5>(p, s, &gv); });
      std::printf(", ");
      print_timing("gpu_map_rect_zero_copy_view", tv);
      std::printf(", \"view_vs_copy_rel_err\": %.3e, ", rel_err(gv, g));
      Result gc;
      Timing tcp = time_reps(3, reps, 60, [&] { return map_rect_once<3, 0, false, 5>(p, s, &gc); });
      print_timing("functor_copies_only_no_glm", tcp);
      std::printf(", ");
    }
    if (!no_cpu) {
      Timing tc = time_reps(1, cpu_reps, 60, [&] { return map_rect_once<1, 0, false, 5>(p, s, &c); });
      print_timing("cpu_reference_tbb_map_rect", tc);
      if (!cpu_only) std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf("%s\"algorithmic_bytes\": %lld}\n", no_cpu ? "" : ", ", (long long)(J * (rows * k * 8 + rows * 4)));
    return 0;
  }

  if (config == "v2") {  // C2 with a var intercept vector
    const int64_t n = n_override ? n_override : 1000000, k = 32;
    Problem p;
    make_problem(&p, 0, n, k, 1000, threads);
    Result g, c;
    glm_vector_alpha_once<true>(p, &g);  // cache fill
    Timing tg = time_reps(3, reps, 60, [&] { return glm_vector_alpha_once<true>(p, &g); });
    std::printf("\"what\": \"bernoulli_logit_glm_lpmf N=%lld K=%lld, var beta, var intercept VECTOR\", ", (long long)n, (long long)k);
    print_timing("gpu", tg);
    if (!no_cpu) {
      Timing tc = time_reps(1, std::max(cpu_reps, 3), 60, [&] { return glm_vector_alpha_once<false>(p, &c); });
      std::printf(", ");
      print_timing("cpu_reference_1_thread", tc);
      std::printf(", \"parity_rel_err\": %.3e", rel_err(g, c));
    }
    std::printf(", \"algorithmic_bytes\": %lld}\n", (long long)(n * k * 8 + n * 4 + 2 * n * 8));
    return 0;
  }

  int family;
  int64_t n, k;
  if (config == "c2") { family = 0; n = 1000000; k = 32; }
  else if (config == "c3") { family = 1; n = 10000000; k = 128; }
  else if (config == "c4") { family = 2; n = 20000000; k = 256; }
  else { std::fprintf(stderr, "unknown config\n"); return 2; }
  if (n_override) n = n_override;
  Problem p;
  make_problem(&p, family, n, k, 1000 + family, threads);
  Result g, c1, ct;
  const double tf0 = now();
  glm_eval_once<true>(p, &g);  // cache fill + first evaluation
  const double fill_s = now() - tf0;
  Timing tg = time_reps(5, reps, 60, [&] { return glm_eval_once<true>(p, &g); });
  static const char* names[] = {"bernoulli_logit_glm_lpmf", "poisson_log_glm_lpmf", "normal_id_glm_lpdf"};
  std::printf("\"what\": \"%s N=%lld K=%lld, var alpha/beta%s\", \"cache_fill_s\": %.3f, ", names[family], (long long)n,
              (long long)k, family == 2 ? "/sigma" : "", fill_s);
  print_timing("gpu", tg);
  const long long alg = (long long)(n * k * 8 + n * (family == 2 ? 8 : 4));
  std::printf(", \"algorithmic_bytes\": %lld, \"gpu_e2e_GBps\": %.1f", alg, alg / tg.median / 1e9);
  if (!no_cpu) {
    Timing t1 = time_reps(0, cpu_reps, 60, [&] { return glm_eval_once<false>(p, &c1); });
    std::printf(", ");
    print_timing("cpu_reference_1_thread", t1);
    std::printf(", \"parity_rel_err_vs_1_thread\": %.3e", rel_err(g, c1));
    // TBB map_rect over row shards (north star's CPU baseline): 4 shards per thread
    int J = threads * 4;
    Problem* pp = &p;
    ShardSet s;
    make_shards(*pp, J, &s);
    Timing tt;
    if (family == 0) tt = time_reps(1, cpu_reps, 60, [&] { return map_rect_once<1, 0, true, 20>(*pp, s, &ct); });
    else if (family == 1) tt = time_reps(1, cpu_reps, 60, [&] { return map_rect_once<1, 1, true, 21>(*pp, s, &ct); });
    else tt = time_reps(1, cpu_reps, 60, [&] { return map_rect_once<1, 2, true, 22>(*pp, s, &ct); });
    const double covered = double(int64_t(n / J) * J) / double(n);
    tt.median /= covered;  // rows dropped by the equal-size split (< J rows) are scaled back in
    tt.best /= covered;
    std::printf(", ");
    print_timing("cpu_reference_tbb_map_rect", tt);
    std::printf(", \"tbb_shards\": %d", J);
  }
  std::printf("}\n");
  return 0;
}
