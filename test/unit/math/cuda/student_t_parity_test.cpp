// student_t_lpdf through the reference's own C++ signature: the STAN_CUDA path against the unmodified reference
// loops (reachable in the same binary as student_t_lpdf_reference_cpu, see overlay/), value and every adjoint,
// for the mixes of data / variable and scalar / container operands a Stan model produces, plus the exception
// contract (stan/math/prim/scal/prob/student_t_lpdf.hpp:57-77).
#ifdef STAN_CUDA
#include <stan/math.hpp>
#include <gtest/gtest.h>
#include <cmath>
#include <random>
#include <string>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::VectorXd;
using stan::math::var;
typedef Matrix<var, Dynamic, 1> vector_v;

namespace {

// north-star gate, 1e-10 relative; `abs_floor` admits the absolute rounding of terms that cancel inside the
// reference's own expression (d/dnu: 0.5 * (digamma(nu/2 + 1/2) - digamma(nu/2)) - 0.5 / nu, see tests/helpers.py)
void expect_close(const VectorXd& got, const VectorXd& ref, const char* what, double abs_floor = 0.0) {
  ASSERT_EQ(got.size(), ref.size()) << what;
  const double scale = ref.size() ? ref.cwiseAbs().maxCoeff() : 0.0;
  for (int i = 0; i < ref.size(); ++i) {
    const double denom = std::max(std::fabs(ref[i]), 1e-6 * scale);
    EXPECT_LE(std::fabs(got[i] - ref[i]), 1e-10 * denom + abs_floor) << what << "[" << i << "] got " << got[i] << " ref " << ref[i];
  }
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

struct Inputs {
  VectorXd y, nu, mu, sigma;
};

Inputs make_inputs(int n, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::normal_distribution<double> nd(0, 1);
  std::uniform_real_distribution<double> ud(0, 1);
  Inputs in;
  in.y.resize(n); in.nu.resize(n); in.mu.resize(n); in.sigma.resize(n);
  for (int i = 0; i < n; ++i) {
    in.y[i] = 3 * nd(rng);
    in.nu[i] = 0.5 + 30 * ud(rng);
    in.mu[i] = nd(rng);
    in.sigma[i] = 0.3 + 3 * ud(rng);
  }
  return in;
}

const double kDigammaFloor = 32 * 2.3e-16 * 4.0;  // 32 ulp of |digamma(nu/2)| <= 4 for nu in [0.5, 60]

#define STAN_CUDA_ST_PARITY(NAME, N, ARGS, GRADS_GPU, GRADS_CPU)                                       \
  {                                                                                                    \
    var lp_gpu = stan::math::student_t_lpdf ARGS;                                                      \
    lp_gpu.grad();                                                                                     \
    const double v_gpu = lp_gpu.val();                                                                 \
    GRADS_GPU;                                                                                         \
    stan::math::set_zero_all_adjoints();                                                               \
    var lp_cpu = stan::math::student_t_lpdf_reference_cpu ARGS;                                        \
    lp_cpu.grad();                                                                                     \
    GRADS_CPU;                                                                                         \
    EXPECT_LE(std::fabs(v_gpu - lp_cpu.val()), 1e-10 * std::fabs(lp_cpu.val())) << NAME " value";      \
    stan::math::set_zero_all_adjoints();                                                               \
  }

}  // namespace

class CudaStudentT : public ::testing::TestWithParam<int> {};

TEST_P(CudaStudentT, vectorised_matches_reference_loops) {
  const int n = GetParam();
  Inputs in = make_inputs(n, 40 + n);
  vector_v y_v = to_v(in.y), nu_v = to_v(in.nu), mu_v = to_v(in.mu), sigma_v = to_v(in.sigma);
  var nu = 4.5, mu = 0.2, sigma = 1.7;
  VectorXd g0, g1, g2, g3, r0, r1, r2, r3;
  // the common model statement: data y, scalar parameters nu and sigma, a vector of locations
  STAN_CUDA_ST_PARITY("y data, mu vector", n, (in.y, nu, mu_v, sigma),
                      (g1 = adj(nu), g2 = adj(mu_v), g3 = adj(sigma)), (r1 = adj(nu), r2 = adj(mu_v), r3 = adj(sigma)));
  expect_close(g1, r1, "dnu (scalar)", n * kDigammaFloor);
  expect_close(g2, r2, "dmu");
  expect_close(g3, r3, "dsigma (scalar)");
  STAN_CUDA_ST_PARITY("propto", n, <true>(in.y, nu, mu_v, sigma), (g1 = adj(nu), g2 = adj(mu_v), g3 = adj(sigma)),
                      (r1 = adj(nu), r2 = adj(mu_v), r3 = adj(sigma)));
  expect_close(g1, r1, "propto dnu", n * kDigammaFloor);
  expect_close(g2, r2, "propto dmu");
  expect_close(g3, r3, "propto dsigma");
  // everything a vector of variables
  STAN_CUDA_ST_PARITY("all var vectors", n, (y_v, nu_v, mu_v, sigma_v),
                      (g0 = adj(y_v), g1 = adj(nu_v), g2 = adj(mu_v), g3 = adj(sigma_v)),
                      (r0 = adj(y_v), r1 = adj(nu_v), r2 = adj(mu_v), r3 = adj(sigma_v)));
  expect_close(g0, r0, "dy");
  expect_close(g1, r1, "dnu", kDigammaFloor);
  expect_close(g2, r2, "dmu");
  expect_close(g3, r3, "dsigma");
  // propto with data nu and sigma drops their summands (:158-163)
  STAN_CUDA_ST_PARITY("propto, only mu variable", n, <true>(in.y, in.nu, mu_v, in.sigma), (g2 = adj(mu_v)), (r2 = adj(mu_v)));
  expect_close(g2, r2, "dmu, data nu and sigma");
  // scalar outcome against vector parameters, std::vector containers
  std::vector<var> nu_s(nu_v.data(), nu_v.data() + n), sg_s(sigma_v.data(), sigma_v.data() + n);
  STAN_CUDA_ST_PARITY("scalar y, std::vector parameters", n, (0.75, nu_s, mu, sg_s),
                      (g1 = adj(nu_v), g2 = adj(mu), g3 = adj(sigma_v)), (r1 = adj(nu_v), r2 = adj(mu), r3 = adj(sigma_v)));
  expect_close(g1, r1, "dnu (std::vector)", kDigammaFloor);
  expect_close(g2, r2, "dmu (scalar)");
  expect_close(g3, r3, "dsigma (std::vector)");
  // all-double call: a double through the same device path
  const double d_gpu = stan::math::student_t_lpdf(in.y, in.nu, in.mu, in.sigma);
  const double d_cpu = stan::math::student_t_lpdf_reference_cpu(in.y, in.nu, in.mu, in.sigma);
  EXPECT_LE(std::fabs(d_gpu - d_cpu), 1e-10 * std::fabs(d_cpu));
  EXPECT_EQ(0.0, stan::math::student_t_lpdf<true>(in.y, in.nu, in.mu, in.sigma));
  stan::math::recover_memory();
}

INSTANTIATE_TEST_CASE_P(Sizes, CudaStudentT, ::testing::Values(1, 3, 257, 20000, 300001));

TEST(CudaStudentTLpdf, known_answers_of_the_reference_fixture) {
  // test/prob/student_t/student_t_test.hpp:13-30 through the vectorised entry
  std::vector<double> y{1.0}, nu{1.0}, mu{0.0}, sigma{1.0};
  EXPECT_NEAR(-1.837877066409345339082, stan::math::student_t_lpdf(y, nu, mu, sigma), 1e-14);
  std::vector<double> y2{-3.0}, nu2{2.0};
  EXPECT_NEAR(-3.596842909197555560041, stan::math::student_t_lpdf(y2, nu2, mu, sigma), 1e-14);
  std::vector<double> yy{1.0, -3.0}, nn{1.0, 2.0};
  EXPECT_NEAR(-1.837877066409345339082 - 3.596842909197555560041, stan::math::student_t_lpdf(yy, nn, 0.0, 1.0), 1e-14);
}

TEST(CudaStudentTLpdf, same_exceptions_as_the_reference) {
  using stan::math::student_t_lpdf;
  const int n = 1000;
  VectorXd y = VectorXd::LinSpaced(n, -1, 1);
  vector_v mu = to_v(VectorXd::Constant(n, 0.1));
  var nu = 3.0, sigma = 1.5;
  VectorXd y_nan = y;
  y_nan[777] = NAN;
  EXPECT_THROW(student_t_lpdf(y_nan, nu, mu, sigma), std::domain_error);
  vector_v mu_inf = to_v(VectorXd::Constant(n, 0.1));
  mu_inf[3] = INFINITY;
  EXPECT_THROW(student_t_lpdf(y, nu, mu_inf, sigma), std::domain_error);
  for (double bad : {0.0, -1.0, static_cast<double>(INFINITY), static_cast<double>(NAN)}) {
    var b = bad;
    EXPECT_THROW(student_t_lpdf(y, b, mu, sigma), std::domain_error) << "nu = " << bad;
    EXPECT_THROW(student_t_lpdf(y, nu, mu, b), std::domain_error) << "sigma = " << bad;
  }
  vector_v nu_v = to_v(VectorXd::Constant(n, 2.0));
  nu_v[999] = -2.0;
  EXPECT_THROW(student_t_lpdf(y, nu_v, mu, sigma), std::domain_error);
  vector_v mu_short = to_v(VectorXd::Constant(n - 1, 0.1));
  EXPECT_THROW(student_t_lpdf(y, nu, mu_short, sigma), std::invalid_argument);
  // a domain error wins over a size error, as in the reference (:67-73)
  VectorXd y_nan_short = y_nan.head(n - 1);
  EXPECT_THROW(student_t_lpdf(y_nan_short, nu, mu, sigma), std::domain_error);
  // empty operand: 0 before any check (:57-59)
  VectorXd none(0);
  EXPECT_EQ(0.0, student_t_lpdf(none, nu, mu, sigma).val());
  // same message as the reference loops
  std::string m_gpu, m_cpu;
  try {
    student_t_lpdf(y, nu_v, mu, sigma);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    stan::math::student_t_lpdf_reference_cpu<false>(y, nu_v, mu, sigma);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // y = inf is not an error (check_not_nan only): the log-density is -inf
  VectorXd y_inf = y;
  y_inf[5] = INFINITY;
  EXPECT_EQ(-INFINITY, student_t_lpdf(y_inf, 3.0, 0.0, 1.0));
  stan::math::recover_memory();
}
#endif
