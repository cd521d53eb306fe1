// cauchy_lpdf / lognormal_lpdf / gamma_lpdf / exponential_lpdf / beta_lpdf / weibull_lpdf on the device through the reference's own C++ signatures, against the
// unmodified reference bodies (reachable in the same binary as *_reference_cpu, see overlay/), at the north star's
// tolerance: every mix of var / double and vector / scalar operands that the overlay routes, propto on and off, the
// reference's exceptions and its LOG_ZERO.
#ifdef STAN_CUDA
#include <stan/math.hpp>
#include <gtest/gtest.h>
#include <random>
#include <string>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::VectorXd;
using stan::math::var;
typedef Matrix<var, Dynamic, 1> vector_v;

namespace {

void expect_rel(double got, double ref, double scale, const char* what) {
  const double denom = std::max(std::fabs(ref), 1e-8 * scale);
  if (denom == 0) {
    EXPECT_EQ(got, ref) << what;
  } else {
    EXPECT_LE(std::fabs(got - ref) / denom, 1e-10) << what << " got " << got << " ref " << ref;
  }
}
void expect_vec(const VectorXd& got, const VectorXd& ref, const char* what) {
  ASSERT_EQ(got.size(), ref.size()) << what;
  const double scale = ref.size() ? ref.cwiseAbs().maxCoeff() : 0.0;
  for (int i = 0; i < ref.size(); ++i) expect_rel(got[i], ref[i], scale, what);
}
VectorXd adj(const vector_v& v) {
  VectorXd a(v.size());
  for (int i = 0; i < v.size(); ++i) a[i] = v[i].adj();
  return a;
}
VectorXd adj(const var& v) { return VectorXd::Constant(1, v.adj()); }
vector_v to_v(const VectorXd& d) {
  vector_v v(d.size());
  for (int i = 0; i < d.size(); ++i) v[i] = d[i];
  return v;
}

struct Data {
  VectorXd y, y_pos, mu, sigma;
};
Data make(int n, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::normal_distribution<double> nd(0, 1);
  std::uniform_real_distribution<double> ud(0.3, 3.0);
  Data d;
  d.y.resize(n), d.y_pos.resize(n), d.mu.resize(n), d.sigma.resize(n);
  for (int i = 0; i < n; ++i) {
    d.y[i] = 2.0 * nd(rng);
    d.y_pos[i] = std::exp(1.5 * nd(rng));
    d.mu[i] = 0.7 * nd(rng);
    d.sigma[i] = ud(rng);
  }
  return d;
}

}  // namespace

#define STAN_CUDA_PARITY(NAME, GPU_CALL, CPU_CALL, ...)                                     \
  {                                                                                         \
    var lp_gpu = GPU_CALL;                                                                  \
    lp_gpu.grad();                                                                          \
    const double v_gpu = lp_gpu.val();                                                      \
    std::vector<VectorXd> g_gpu = {__VA_ARGS__};                                            \
    stan::math::set_zero_all_adjoints();                                                    \
    var lp_cpu = CPU_CALL;                                                                  \
    lp_cpu.grad();                                                                          \
    std::vector<VectorXd> g_cpu = {__VA_ARGS__};                                            \
    expect_rel(v_gpu, lp_cpu.val(), std::fabs(lp_cpu.val()), NAME " value");                \
    for (size_t q = 0; q < g_cpu.size(); ++q) expect_vec(g_gpu[q], g_cpu[q], NAME " grad"); \
    stan::math::set_zero_all_adjoints();                                                    \
  }

class CudaLocScale : public ::testing::TestWithParam<int> {};

TEST_P(CudaLocScale, cauchy_matches_reference_loop) {
  const int n = GetParam();
  Data d = make(n, 100 + n);
  vector_v y = to_v(d.y), mu = to_v(d.mu), sigma = to_v(d.sigma);
  var mu0 = 0.2, sigma0 = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("cauchy, all var vectors", cauchy_lpdf(y, mu, sigma), cauchy_lpdf_reference_cpu<false>(y, mu, sigma), adj(y),
                   adj(mu), adj(sigma));
  STAN_CUDA_PARITY("cauchy, the model statement", cauchy_lpdf(d.y, mu, sigma0), cauchy_lpdf_reference_cpu<false>(d.y, mu, sigma0),
                   adj(mu), adj(sigma0));
  STAN_CUDA_PARITY("cauchy propto, data y", cauchy_lpdf<true>(d.y, mu, sigma0), cauchy_lpdf_reference_cpu<true>(d.y, mu, sigma0),
                   adj(mu), adj(sigma0));
  STAN_CUDA_PARITY("cauchy propto, data sigma", cauchy_lpdf<true>(d.y, mu, d.sigma), cauchy_lpdf_reference_cpu<true>(d.y, mu, d.sigma),
                   adj(mu));
  STAN_CUDA_PARITY("cauchy, scalar location", cauchy_lpdf(y, mu0, d.sigma), cauchy_lpdf_reference_cpu<false>(y, mu0, d.sigma), adj(y),
                   adj(mu0));
  const double v = cauchy_lpdf(d.y, d.mu, d.sigma);
  expect_rel(v, cauchy_lpdf_reference_cpu<false>(d.y, d.mu, d.sigma), std::fabs(v), "cauchy, all double");
  EXPECT_EQ(0.0, cauchy_lpdf<true>(d.y, d.mu, d.sigma));  // nothing to compute: the reference body
  stan::math::recover_memory();
}

TEST_P(CudaLocScale, lognormal_matches_reference_loop) {
  const int n = GetParam();
  Data d = make(n, 200 + n);
  vector_v y = to_v(d.y_pos), mu = to_v(d.mu), sigma = to_v(d.sigma);
  var mu0 = 0.2, sigma0 = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("lognormal, all var vectors", lognormal_lpdf(y, mu, sigma), lognormal_lpdf_reference_cpu<false>(y, mu, sigma),
                   adj(y), adj(mu), adj(sigma));
  STAN_CUDA_PARITY("lognormal, the model statement", lognormal_lpdf(d.y_pos, mu, sigma0),
                   lognormal_lpdf_reference_cpu<false>(d.y_pos, mu, sigma0), adj(mu), adj(sigma0));
  STAN_CUDA_PARITY("lognormal propto, data y", lognormal_lpdf<true>(d.y_pos, mu, sigma0),
                   lognormal_lpdf_reference_cpu<true>(d.y_pos, mu, sigma0), adj(mu), adj(sigma0));
  STAN_CUDA_PARITY("lognormal propto, var y only", lognormal_lpdf<true>(y, d.mu, d.sigma),
                   lognormal_lpdf_reference_cpu<true>(y, d.mu, d.sigma), adj(y));
  STAN_CUDA_PARITY("lognormal, scalar location", lognormal_lpdf(y, mu0, d.sigma), lognormal_lpdf_reference_cpu<false>(y, mu0, d.sigma),
                   adj(y), adj(mu0));
  const double v = lognormal_lpdf(d.y_pos, d.mu, d.sigma);
  expect_rel(v, lognormal_lpdf_reference_cpu<false>(d.y_pos, d.mu, d.sigma), std::fabs(v), "lognormal, all double");
  EXPECT_EQ(0.0, lognormal_lpdf<true>(d.y_pos, d.mu, d.sigma));
  stan::math::recover_memory();
}

TEST_P(CudaLocScale, gamma_and_exponential_match_reference_loop) {
  const int n = GetParam();
  Data d = make(n, 300 + n);
  VectorXd al_d = d.sigma.array() + 0.2;  // shapes in (0.5, 3.2)
  vector_v y = to_v(d.y_pos), alpha = to_v(al_d), beta = to_v(d.sigma);
  var alpha0 = 2.2, beta0 = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("gamma, all var vectors", gamma_lpdf(y, alpha, beta), gamma_lpdf_reference_cpu<false>(y, alpha, beta), adj(y),
                   adj(beta));
  STAN_CUDA_PARITY("gamma, the model statement", gamma_lpdf(d.y_pos, alpha0, beta), gamma_lpdf_reference_cpu<false>(d.y_pos, alpha0, beta),
                   adj(beta));
  STAN_CUDA_PARITY("gamma propto, data y", gamma_lpdf<true>(d.y_pos, alpha0, beta), gamma_lpdf_reference_cpu<true>(d.y_pos, alpha0, beta),
                   adj(beta));
  STAN_CUDA_PARITY("gamma propto, var y only", gamma_lpdf<true>(y, al_d, d.sigma), gamma_lpdf_reference_cpu<true>(y, al_d, d.sigma),
                   adj(y));
  {  // d/dalpha = -digamma(alpha) + log(beta) + log(y): accurate to the magnitude of its terms (see tests/test_locscale.py)
    var lp = gamma_lpdf(d.y_pos, alpha, beta0);
    lp.grad();
    VectorXd g = adj(alpha);
    const double gb = beta0.adj();
    set_zero_all_adjoints();
    var lp_ref = gamma_lpdf_reference_cpu<false>(d.y_pos, alpha, beta0);
    lp_ref.grad();
    expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "gamma, var shape vector: value");
    expect_rel(gb, beta0.adj(), std::fabs(beta0.adj()), "gamma, var shape vector: d/dbeta");
    for (int i = 0; i < n; ++i) {
      const double terms = std::fabs(digamma(al_d[i])) + std::fabs(std::log(1.3)) + std::fabs(std::log(d.y_pos[i]));
      EXPECT_LE(std::fabs(g[i] - alpha[i].adj()), 1e-10 * std::max(std::fabs(alpha[i].adj()), terms)) << "gamma d/dalpha " << i;
    }
    set_zero_all_adjoints();
  }
  const double v = gamma_lpdf(d.y_pos, al_d, d.sigma);
  expect_rel(v, gamma_lpdf_reference_cpu<false>(d.y_pos, al_d, d.sigma), std::fabs(v), "gamma, all double");
  EXPECT_EQ(0.0, gamma_lpdf<true>(d.y_pos, al_d, d.sigma));
  STAN_CUDA_PARITY("exponential, var vectors", exponential_lpdf(y, beta), exponential_lpdf_reference_cpu<false>(y, beta), adj(y),
                   adj(beta));
  STAN_CUDA_PARITY("exponential, the model statement", exponential_lpdf(d.y_pos, beta0), exponential_lpdf_reference_cpu<false>(d.y_pos, beta0),
                   adj(beta0));
  STAN_CUDA_PARITY("exponential propto, var y only", exponential_lpdf<true>(y, d.sigma), exponential_lpdf_reference_cpu<true>(y, d.sigma),
                   adj(y));
  const double w = exponential_lpdf(d.y_pos, d.sigma);
  expect_rel(w, exponential_lpdf_reference_cpu<false>(d.y_pos, d.sigma), std::fabs(w), "exponential, all double");
  EXPECT_EQ(0.0, exponential_lpdf<true>(d.y_pos, d.sigma));
  stan::math::recover_memory();
}

TEST_P(CudaLocScale, inv_gamma_matches_reference_loop) {
  const int n = GetParam();
  Data d = make(n, 500 + n);
  VectorXd al_d = d.sigma.array() + 0.2;
  vector_v y = to_v(d.y_pos), alpha = to_v(al_d), beta = to_v(d.sigma);
  var alpha0 = 2.2, beta0 = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("inv_gamma, var y and beta vectors", inv_gamma_lpdf(y, al_d, beta), inv_gamma_lpdf_reference_cpu<false>(y, al_d, beta),
                   adj(y), adj(beta));
  STAN_CUDA_PARITY("inv_gamma, the model statement", inv_gamma_lpdf(d.y_pos, alpha0, beta),
                   inv_gamma_lpdf_reference_cpu<false>(d.y_pos, alpha0, beta), adj(beta));
  STAN_CUDA_PARITY("inv_gamma propto, var y only", inv_gamma_lpdf<true>(y, al_d, d.sigma),
                   inv_gamma_lpdf_reference_cpu<true>(y, al_d, d.sigma), adj(y));
  {  // d/dalpha = -digamma(alpha) + log(beta) - log(y): against the magnitude of its terms
    var lp = inv_gamma_lpdf(d.y_pos, alpha, beta0);
    lp.grad();
    VectorXd g = adj(alpha);
    set_zero_all_adjoints();
    var lp_ref = inv_gamma_lpdf_reference_cpu<false>(d.y_pos, alpha, beta0);
    lp_ref.grad();
    expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "inv_gamma, var shape vector: value");
    for (int i = 0; i < n; ++i) {
      const double terms = std::fabs(digamma(al_d[i])) + std::fabs(std::log(1.3)) + std::fabs(std::log(d.y_pos[i]));
      EXPECT_LE(std::fabs(g[i] - alpha[i].adj()), 1e-10 * std::max(std::fabs(alpha[i].adj()), terms)) << "inv_gamma d/dalpha " << i;
    }
    set_zero_all_adjoints();
  }
  const double v = inv_gamma_lpdf(d.y_pos, al_d, d.sigma);
  expect_rel(v, inv_gamma_lpdf_reference_cpu<false>(d.y_pos, al_d, d.sigma), std::fabs(v), "inv_gamma, all double");
  EXPECT_EQ(0.0, inv_gamma_lpdf<true>(d.y_pos, al_d, d.sigma));
  VectorXd y_zero = d.y_pos;
  y_zero[0] = 0.0;
  EXPECT_EQ(LOG_ZERO, inv_gamma_lpdf(y_zero, alpha, beta).val());   // :58-63
  vector_v alpha_bad = to_v(al_d);
  alpha_bad[0] = -1.0;
  EXPECT_THROW(inv_gamma_lpdf(d.y_pos, alpha_bad, beta), std::domain_error);
  stan::math::recover_memory();
}

// partials made of terms that may cancel (digamma differences, (1 - p)(log y - log sigma), ...) are compared against the
// magnitude `floor` of those terms, like student_t's d/dnu (tests/test_locscale.py has the same gates)
void expect_floor(double got, double ref, double floor, const char* what, int i) {
  EXPECT_LE(std::fabs(got - ref), 1e-10 * std::max(std::fabs(ref), floor)) << what << " " << i << " got " << got << " ref " << ref;
}

TEST_P(CudaLocScale, beta_and_weibull_match_reference_loop) {
  const int n = GetParam();
  Data d = make(n, 400 + n);
  using namespace stan::math;  // NOLINT
  VectorXd yb(n), a_d = d.sigma.array() + 0.2, b_d = d.sigma.reverse().array() + 0.1;
  for (int i = 0; i < n; ++i) yb[i] = std::min(std::max(1.0 / (1.0 + std::exp(-d.y[i])), 1e-9), 1.0 - 1e-9);
  {
    vector_v y = to_v(yb), alpha = to_v(a_d), beta = to_v(b_d);
    var lp = beta_lpdf(y, alpha, beta);
    lp.grad();
    VectorXd gy = adj(y), ga = adj(alpha), gb = adj(beta);
    set_zero_all_adjoints();
    var lp_ref = beta_lpdf_reference_cpu<false>(y, alpha, beta);
    lp_ref.grad();
    expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "beta value");
    expect_vec(gy, adj(y), "beta d/dy");
    for (int i = 0; i < n; ++i) {
      const double dab = std::fabs(digamma(a_d[i] + b_d[i]));
      expect_floor(ga[i], alpha[i].adj(), std::fabs(std::log(yb[i])) + dab + std::fabs(digamma(a_d[i])), "beta d/dalpha", i);
      expect_floor(gb[i], beta[i].adj(), std::fabs(log1m(yb[i])) + dab + std::fabs(digamma(b_d[i])), "beta d/dbeta", i);
    }
    set_zero_all_adjoints();
    STAN_CUDA_PARITY("beta propto, var y only", beta_lpdf<true>(y, a_d, b_d), beta_lpdf_reference_cpu<true>(y, a_d, b_d), adj(y));
    const double v = beta_lpdf(yb, a_d, b_d);
    expect_rel(v, beta_lpdf_reference_cpu<false>(yb, a_d, b_d), std::fabs(v), "beta, all double");
    EXPECT_EQ(0.0, beta_lpdf<true>(yb, a_d, b_d));
  }
  {
    vector_v y = to_v(d.y_pos), alpha = to_v(a_d), sigma = to_v(d.sigma);
    var lp = weibull_lpdf(y, alpha, sigma);
    lp.grad();
    VectorXd gy = adj(y), ga = adj(alpha), gs = adj(sigma);
    set_zero_all_adjoints();
    var lp_ref = weibull_lpdf_reference_cpu<false>(y, alpha, sigma);
    lp_ref.grad();
    expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "weibull value");
    for (int i = 0; i < n; ++i) {
      const double pw = std::pow(d.y_pos[i] / d.sigma[i], a_d[i]);
      expect_floor(gy[i], y[i].adj(), (std::fabs(a_d[i] - 1.0) + a_d[i] * pw) / d.y_pos[i], "weibull d/dy", i);
      expect_floor(ga[i], alpha[i].adj(),
                   1.0 / a_d[i] + (1.0 + pw) * (std::fabs(std::log(d.y_pos[i])) + std::fabs(std::log(d.sigma[i]))), "weibull d/dalpha", i);
      expect_floor(gs[i], sigma[i].adj(), a_d[i] / d.sigma[i] * (pw + 1.0), "weibull d/dsigma", i);
    }
    set_zero_all_adjoints();
    var s0 = 1.3;
    {
      var lq = weibull_lpdf<true>(d.y_pos, a_d, s0);
      lq.grad();
      const double g0 = s0.adj(), v0 = lq.val();
      set_zero_all_adjoints();
      var lq_ref = weibull_lpdf_reference_cpu<true>(d.y_pos, a_d, s0);
      lq_ref.grad();
      expect_rel(v0, lq_ref.val(), std::fabs(lq_ref.val()), "weibull propto, var scale: value");
      double floor = 0;
      for (int i = 0; i < n; ++i) floor += a_d[i] / 1.3 * (std::pow(d.y_pos[i] / 1.3, a_d[i]) + 1.0);
      expect_floor(g0, s0.adj(), floor, "weibull propto, var scale: d/dsigma", 0);
      set_zero_all_adjoints();
    }
    const double v = weibull_lpdf(d.y_pos, a_d, d.sigma);
    expect_rel(v, weibull_lpdf_reference_cpu<false>(d.y_pos, a_d, d.sigma), std::fabs(v), "weibull, all double");
    EXPECT_EQ(0.0, weibull_lpdf<true>(d.y_pos, a_d, d.sigma));
  }
  stan::math::recover_memory();
}

INSTANTIATE_TEST_CASE_P(Sizes, CudaLocScale, ::testing::Values(1, 3, 1000, 100003));

TEST(CudaLocScaleErrors, beta_and_weibull_same_exceptions_and_log_zero_as_the_reference) {
  using namespace stan::math;  // NOLINT
  const int n = 2000;
  Data d = make(n, 9);
  VectorXd yb(n), a_d = d.sigma.array() + 0.2;
  for (int i = 0; i < n; ++i) yb[i] = std::min(std::max(1.0 / (1.0 + std::exp(-d.y[i])), 1e-9), 1.0 - 1e-9);
  vector_v alpha = to_v(a_d), second = to_v(d.sigma);
  VectorXd yb_hi = yb, yb_nan = yb, yw_neg = d.y_pos, yw_inf = d.y_pos;
  yb_hi[5] = 1.5;
  yb_nan[5] = NAN;
  yw_neg[5] = -0.5;
  yw_inf[5] = INFINITY;
  vector_v alpha_zero = to_v(a_d), second_inf = to_v(d.sigma);
  alpha_zero[17] = 0.0;
  second_inf[9] = INFINITY;
  EXPECT_THROW(beta_lpdf(yb_hi, alpha, second), std::domain_error);
  EXPECT_THROW(beta_lpdf(yb_nan, alpha, second), std::domain_error);
  EXPECT_THROW(beta_lpdf(yb, alpha_zero, second), std::domain_error);
  EXPECT_THROW(beta_lpdf(yb, alpha, second_inf), std::domain_error);
  EXPECT_THROW(weibull_lpdf(yw_inf, alpha, second), std::domain_error);
  EXPECT_THROW(weibull_lpdf(d.y_pos, alpha_zero, second), std::domain_error);
  EXPECT_THROW(weibull_lpdf(d.y_pos, alpha, second_inf), std::domain_error);
  vector_v shorter = to_v(d.sigma.head(n - 1));
  EXPECT_THROW(beta_lpdf(yb, alpha, shorter), std::invalid_argument);
  EXPECT_THROW(weibull_lpdf(d.y_pos, alpha, shorter), std::invalid_argument);
  std::string m_gpu, m_cpu;
  try {
    beta_lpdf(yb_hi, alpha, second);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    beta_lpdf_reference_cpu<false>(yb_hi, alpha, second);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  m_gpu.clear();
  m_cpu.clear();
  try {
    weibull_lpdf(d.y_pos, alpha, second_inf);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    weibull_lpdf_reference_cpu<false>(d.y_pos, alpha, second_inf);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  EXPECT_EQ(LOG_ZERO, weibull_lpdf(yw_neg, alpha, second).val());                       // weibull_lpdf.hpp:56-61
  EXPECT_EQ(LOG_ZERO, weibull_lpdf_reference_cpu<false>(yw_neg, alpha, second).val());
  VectorXd none(0);
  vector_v empty(0);
  EXPECT_EQ(0.0, beta_lpdf(none, empty, empty).val());
  EXPECT_EQ(0.0, weibull_lpdf(none, empty, empty).val());
  stan::math::recover_memory();
}

TEST(CudaLocScaleErrors, gamma_and_exponential_same_exceptions_and_log_zero_as_the_reference) {
  using namespace stan::math;  // NOLINT
  const int n = 2000;
  Data d = make(n, 8);
  VectorXd al_d = d.sigma.array() + 0.2;
  vector_v alpha = to_v(al_d), beta = to_v(d.sigma);
  VectorXd y_nan = d.y_pos, y_neg = d.y_pos;
  y_nan[1999] = NAN;
  y_neg[5] = -0.5;
  vector_v alpha_zero = to_v(al_d), beta_inf = to_v(d.sigma);
  alpha_zero[17] = 0.0;
  beta_inf[9] = INFINITY;
  EXPECT_THROW(gamma_lpdf(y_nan, alpha, beta), std::domain_error);
  EXPECT_THROW(gamma_lpdf(d.y_pos, alpha_zero, beta), std::domain_error);
  EXPECT_THROW(gamma_lpdf(d.y_pos, alpha, beta_inf), std::domain_error);
  EXPECT_THROW(exponential_lpdf(y_nan, beta), std::domain_error);
  EXPECT_THROW(exponential_lpdf(y_neg, beta), std::domain_error);
  EXPECT_THROW(exponential_lpdf(d.y_pos, beta_inf), std::domain_error);
  vector_v shorter = to_v(d.sigma.head(n - 1));
  EXPECT_THROW(gamma_lpdf(d.y_pos, alpha, shorter), std::invalid_argument);
  EXPECT_THROW(exponential_lpdf(d.y_pos, shorter), std::invalid_argument);
  std::string m_gpu, m_cpu;
  try {
    gamma_lpdf(d.y_pos, alpha_zero, beta);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    gamma_lpdf_reference_cpu<false>(d.y_pos, alpha_zero, beta);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  m_gpu.clear();
  m_cpu.clear();
  try {
    exponential_lpdf(y_neg, beta);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    exponential_lpdf_reference_cpu<false>(y_neg, beta);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // a negative outcome: LOG_ZERO (gamma_lpdf.hpp:69-74)
  EXPECT_EQ(LOG_ZERO, gamma_lpdf(y_neg, alpha, beta).val());
  EXPECT_EQ(LOG_ZERO, gamma_lpdf_reference_cpu<false>(y_neg, alpha, beta).val());
  VectorXd none(0);
  vector_v empty(0);
  EXPECT_EQ(0.0, gamma_lpdf(none, empty, empty).val());
  EXPECT_EQ(0.0, exponential_lpdf(none, empty).val());
  stan::math::recover_memory();
}

TEST(CudaLocScaleErrors, same_exceptions_and_log_zero_as_the_reference) {
  using namespace stan::math;  // NOLINT
  const int n = 2000;
  Data d = make(n, 7);
  vector_v mu = to_v(d.mu), sigma = to_v(d.sigma);
  VectorXd y_nan = d.y, y_neg = d.y_pos, y_zero = d.y_pos;
  y_nan[1999] = NAN;
  y_neg[5] = -0.5;
  y_zero[5] = 0.0;
  vector_v mu_inf = to_v(d.mu), sigma_zero = to_v(d.sigma), sigma_inf = to_v(d.sigma);
  mu_inf[17] = INFINITY;
  sigma_zero[9] = 0.0;
  sigma_inf[9] = INFINITY;
  EXPECT_THROW(cauchy_lpdf(y_nan, mu, sigma), std::domain_error);
  EXPECT_THROW(cauchy_lpdf(d.y, mu_inf, sigma), std::domain_error);
  EXPECT_THROW(cauchy_lpdf(d.y, mu, sigma_zero), std::domain_error);
  EXPECT_THROW(cauchy_lpdf(d.y, mu, sigma_inf), std::domain_error);
  EXPECT_THROW(lognormal_lpdf(y_nan, mu, sigma), std::domain_error);
  EXPECT_THROW(lognormal_lpdf(y_neg, mu, sigma), std::domain_error);
  EXPECT_THROW(lognormal_lpdf(d.y_pos, mu_inf, sigma), std::domain_error);
  EXPECT_THROW(lognormal_lpdf(d.y_pos, mu, sigma_inf), std::domain_error);
  vector_v shorter = to_v(d.mu.head(n - 1));
  EXPECT_THROW(cauchy_lpdf(d.y, shorter, sigma), std::invalid_argument);
  EXPECT_THROW(lognormal_lpdf(d.y_pos, shorter, sigma), std::invalid_argument);
  std::string m_gpu, m_cpu;
  try {
    cauchy_lpdf(d.y, mu, sigma_inf);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    cauchy_lpdf_reference_cpu<false>(d.y, mu, sigma_inf);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  m_gpu.clear();
  m_cpu.clear();
  try {
    lognormal_lpdf(y_neg, mu, sigma);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    lognormal_lpdf_reference_cpu<false>(y_neg, mu, sigma);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // an outcome equal to zero: LOG_ZERO (lognormal_lpdf.hpp:41-45)
  EXPECT_EQ(LOG_ZERO, lognormal_lpdf(y_zero, mu, sigma).val());
  EXPECT_EQ(LOG_ZERO, lognormal_lpdf_reference_cpu<false>(y_zero, mu, sigma).val());
  VectorXd none(0);
  vector_v empty(0);
  EXPECT_EQ(0.0, cauchy_lpdf(none, empty, empty).val());
  EXPECT_EQ(0.0, lognormal_lpdf(none, empty, empty).val());
  stan::math::recover_memory();
}

TEST(CudaLocScaleRouting, the_device_is_used) {
  stan_cuda_stats_t s0, s1;
  Data d = make(5000, 3);
  vector_v mu = to_v(d.mu);
  stan_cuda_stats(&s0);
  var lp = stan::math::cauchy_lpdf(d.y, mu, 1.5) + stan::math::lognormal_lpdf(d.y_pos, mu, 1.5);
  stan_cuda_stats(&s1);
  EXPECT_GE(s1.evals + s1.kernel_launches, s0.evals + s0.kernel_launches + 2);
  EXPECT_TRUE(std::isfinite(lp.val()));
  stan::math::recover_memory();
}
#endif
