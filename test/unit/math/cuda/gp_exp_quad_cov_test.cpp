// gp_exp_quad_cov on the device through the reference's own C++ signatures -- the six double
// argument kinds the reference sends to its OpenCL backend (stan/math/prim/mat/fun/gp_exp_quad_cov.hpp:344-655) --
// against the unmodified generic templates (gp_exp_quad_cov_reference_cpu, see overlay/), and the exception
// contract of those templates (:175-342).
#ifdef STAN_CUDA
#include <stan/math/rev/mat.hpp>
#include <gtest/gtest.h>
#include <cmath>
#include <limits>
#include <random>
#include <stdexcept>
#include <vector>

using Eigen::MatrixXd;
using Eigen::VectorXd;

namespace {

std::vector<double> scalars(int n, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::uniform_real_distribution<double> u(-3, 3);
  std::vector<double> x(n);
  for (double& v : x) v = u(rng);
  return x;
}
std::vector<VectorXd> points(int n, int dim, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::uniform_real_distribution<double> u(-2, 2);
  std::vector<VectorXd> x(n, VectorXd(dim));
  for (auto& p : x)
    for (int d = 0; d < dim; ++d) p[d] = u(rng);
  return x;
}
void expect_cov(const MatrixXd& got, const MatrixXd& ref, const char* what) {
  ASSERT_EQ(got.rows(), ref.rows()) << what;
  ASSERT_EQ(got.cols(), ref.cols()) << what;
  const double scale = ref.cwiseAbs().maxCoeff();
  double worst = 0;
  for (int i = 0; i < ref.size(); ++i) worst = std::max(worst, std::fabs(got(i) - ref(i)) / std::max(std::fabs(ref(i)), 1e-6 * scale));
  EXPECT_LE(worst, 1e-10) << what;
}
uint64_t device_calls() {
  stan_cuda_stats_t s;
  stan_cuda_stats(&s);
  return s.evals + s.kernel_launches;
}

}  // namespace

TEST(CudaGpExpQuadCov, six_argument_kinds_match_reference) {
  using namespace stan::math;  // NOLINT
  for (int n : {256, 300, 777}) {
    std::vector<double> x = scalars(n, n), x2 = scalars(n + 13, n + 1);
    std::vector<VectorXd> xv = points(n, 3, n + 2), xv2 = points(n + 5, 3, n + 3);
    std::vector<double> ls{0.7, 1.3, 2.0};
    const uint64_t before = device_calls();
    MatrixXd c1 = gp_exp_quad_cov(x, 1.5, 0.4);
    EXPECT_GT(device_calls(), before);
    expect_cov(c1, gp_exp_quad_cov_reference_cpu(x, 1.5, 0.4), "scalars");
    EXPECT_EQ(0.0, (c1 - c1.transpose()).cwiseAbs().maxCoeff());
    EXPECT_EQ(1.5 * 1.5, c1(7, 7));  // the diagonal is sigma^2 exactly (:53)
    expect_cov(gp_exp_quad_cov(xv, 1.5, 0.4), gp_exp_quad_cov_reference_cpu(xv, 1.5, 0.4), "vectors");
    expect_cov(gp_exp_quad_cov(xv, 1.5, ls), gp_exp_quad_cov_reference_cpu(xv, 1.5, ls), "vectors, length scale per dimension");
    expect_cov(gp_exp_quad_cov(x, x2, 1.5, 0.4), gp_exp_quad_cov_reference_cpu(x, x2, 1.5, 0.4), "two samples of scalars");
    expect_cov(gp_exp_quad_cov(xv, xv2, 1.5, 0.4), gp_exp_quad_cov_reference_cpu(xv, xv2, 1.5, 0.4), "two samples of vectors");
    expect_cov(gp_exp_quad_cov(xv, xv2, 1.5, ls), gp_exp_quad_cov_reference_cpu(xv, xv2, 1.5, ls), "two samples, length scales");
    expect_cov(cov_exp_quad(x, 1.5, 0.4), c1, "cov_exp_quad forwards to gp_exp_quad_cov");
  }
}

TEST(CudaGpExpQuadCov, routing_and_exceptions) {
  using namespace stan::math;  // NOLINT
  std::vector<double> small = scalars(50, 1);
  uint64_t before = device_calls();
  MatrixXd cs = gp_exp_quad_cov(small, 1.5, 0.4);  // 2500 elements: host
  EXPECT_EQ(before, device_calls());
  EXPECT_EQ(0.0, (cs - gp_exp_quad_cov_reference_cpu(small, 1.5, 0.4)).cwiseAbs().maxCoeff());
  std::vector<double> x = scalars(300, 2);
  EXPECT_THROW(gp_exp_quad_cov(x, -1.0, 0.4), std::domain_error);
  EXPECT_THROW(gp_exp_quad_cov(x, 1.0, 0.0), std::domain_error);
  std::vector<double> x_nan = x;
  x_nan[17] = std::numeric_limits<double>::quiet_NaN();
  EXPECT_THROW(gp_exp_quad_cov(x_nan, 1.0, 0.4), std::domain_error);
  EXPECT_THROW(gp_exp_quad_cov(x, x_nan, 1.0, 0.4), std::domain_error);
  std::vector<VectorXd> xv = points(300, 3, 3);
  std::vector<double> ls_short{0.7, 1.3};
  EXPECT_THROW(gp_exp_quad_cov(xv, 1.0, ls_short), std::invalid_argument);
  std::vector<double> ls_bad{0.7, -1.3, 1.0};
  EXPECT_THROW(gp_exp_quad_cov(xv, 1.0, ls_bad), std::domain_error);
  std::string m_gpu, m_cpu;
  try {
    gp_exp_quad_cov(x_nan, 1.0, 0.4);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    gp_exp_quad_cov_reference_cpu(x_nan, 1.0, 0.4);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  // var hyperparameters keep going to the reverse-mode overloads (rev/mat/fun/gp_exp_quad_cov.hpp)
  var sigma = 1.5, l = 0.4;
  Eigen::Matrix<var, -1, -1> cv = gp_exp_quad_cov(x, sigma, l);
  cv(3, 5).grad();
  EXPECT_NE(0.0, l.adj());
  std::vector<double> empty;
  EXPECT_EQ(0, gp_exp_quad_cov(empty, 1.0, 0.4).size());
  recover_memory();
}
#endif
