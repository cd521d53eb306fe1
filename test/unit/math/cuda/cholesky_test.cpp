// cholesky_decompose on the device through the reference's own C++ signatures, against the
// unmodified reference functions (cholesky_decompose_reference_cpu, see overlay/).  Modelled on the reference's
// OpenCL tests (test/unit/math/opencl/cholesky_decompose_test.cpp:12-98: the 3 x 3 / 4 x 4 fixtures, random
// X X' + n I at 10 ... 2000 with the transfer threshold forced down, sizes around the block edges) at 1e-10
// instead of 1e-8, plus the reverse sweep, which the reference tests only through finite differences at n <= 4
// (test/unit/math/mix/mat/fun/cholesky_decompose_test.cpp, also run unmodified against the overlay).
#ifdef STAN_CUDA
#include <stan/math/rev/mat.hpp>
#include <gtest/gtest.h>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

using Eigen::MatrixXd;
using stan::math::var;
typedef Eigen::Matrix<var, -1, -1> matrix_v;

namespace {

struct threshold_override {
  int64_t saved;
  explicit threshold_override(int64_t v) : saved(stan::math::cuda::cholesky_size_worth_transfer()) {
    stan::math::cuda::cholesky_size_worth_transfer() = v;
  }
  ~threshold_override() { stan::math::cuda::cholesky_size_worth_transfer() = saved; }
};

MatrixXd spd(int n, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::uniform_real_distribution<double> u(-1, 1);
  MatrixXd m(n, n);
  for (int i = 0; i < n * n; ++i) m(i) = u(rng);
  MatrixXd a = m * m.transpose() + n * MatrixXd::Identity(n, n);
  return 0.5 * (a + a.transpose());
}

uint64_t device_calls() {
  stan_cuda_stats_t s;
  stan_cuda_stats(&s);
  return s.evals + s.kernel_launches;
}

// the reference's own metric (opencl/cholesky_decompose_test.cpp:52-60): worst relative error over the lower triangle
double lower_rel_err(const MatrixXd& got, const MatrixXd& ref) {
  double worst = 0;
  const double scale = ref.cwiseAbs().maxCoeff();
  for (int i = 0; i < ref.rows(); ++i)
    for (int j = 0; j <= i; ++j)
      worst = std::max(worst, std::fabs(got(i, j) - ref(i, j)) / std::max(std::fabs(ref(i, j)), 1e-6 * scale));
  return worst;
}

}  // namespace

TEST(CudaCholesky, fixtures_of_the_reference_opencl_test) {
  threshold_override force(0);
  MatrixXd m0(3, 3), m1(4, 4);
  m0 << 25, 15, -5, 15, 18, 0, -5, 0, 11;
  m1 << 18, 22, 54, 42, 22, 70, 86, 62, 54, 86, 174, 134, 42, 62, 134, 106;
  const uint64_t before = device_calls();
  MatrixXd l0 = stan::math::cholesky_decompose(m0), l1 = stan::math::cholesky_decompose(m1);
  EXPECT_GT(device_calls(), before);
  MatrixXd expected0(3, 3);
  expected0 << 5, 0, 0, 3, 3, 0, -1, 1, 3;
  EXPECT_LE((l0 - expected0).cwiseAbs().maxCoeff(), 1e-14);
  EXPECT_LE(lower_rel_err(l1, stan::math::cholesky_decompose_reference_cpu(m1)), 1e-10);
  EXPECT_EQ(0.0, l1(0, 3));  // zeros above the diagonal, as LLT::matrixL() returns them
}

class CudaCholeskySizes : public ::testing::TestWithParam<int> {};

TEST_P(CudaCholeskySizes, prim_matches_reference) {
  const int n = GetParam();
  threshold_override force(0);
  MatrixXd A = spd(n, 100 + n);
  MatrixXd L = stan::math::cholesky_decompose(A), Lr = stan::math::cholesky_decompose_reference_cpu(A);
  EXPECT_LE(lower_rel_err(L, Lr), 1e-10);
  EXPECT_EQ(0.0, MatrixXd(L.triangularView<Eigen::StrictlyUpper>()).cwiseAbs().maxCoeff());
}

TEST_P(CudaCholeskySizes, reverse_sweep_matches_reference) {
  const int n = GetParam();
  MatrixXd A = spd(n, 200 + n);
  std::mt19937_64 rng(300 + n);
  std::uniform_real_distribution<double> u(-1, 1);
  MatrixXd W(n, n);
  for (int i = 0; i < n * n; ++i) W(i) = u(rng);
  MatrixXd adj[2], val[2];
  for (int side = 0; side < 2; ++side) {  // 0: device, 1: reference (cholesky_scalar up to n = 35, cholesky_block above)
    threshold_override force(side == 0 ? 0 : INT64_MAX);
    matrix_v Av(n, n);
    for (int i = 0; i < n * n; ++i) Av(i) = A(i);
    matrix_v Lv = stan::math::cholesky_decompose(Av);
    var f = 0;
    for (int j = 0; j < n; ++j)
      for (int i = j; i < n; ++i) f += W(i, j) * Lv(i, j);
    f.grad();
    adj[side].resize(n, n);
    val[side].resize(n, n);
    for (int i = 0; i < n * n; ++i) {
      adj[side](i) = Av(i).adj();
      val[side](i) = Lv(i).val();
    }
    stan::math::recover_memory();
  }
  EXPECT_LE(lower_rel_err(val[0], val[1]), 1e-10);
  EXPECT_LE(lower_rel_err(adj[0], adj[1]), 1e-10);
  // only the lower triangle of A is an operand (:278-282)
  EXPECT_EQ(0.0, MatrixXd(adj[0].triangularView<Eigen::StrictlyUpper>()).cwiseAbs().maxCoeff());
}

INSTANTIATE_TEST_CASE_P(Sizes, CudaCholeskySizes, ::testing::Values(1, 2, 10, 36, 50, 63, 64, 65, 100, 127, 128, 130, 511, 512, 777));

TEST(CudaCholesky, big_sizes_of_the_reference_opencl_test) {
  for (int n : {1251, 1704, 2000}) {
    MatrixXd A = spd(n, n);
    const uint64_t before = device_calls();
    MatrixXd L = stan::math::cholesky_decompose(A);  // above the default threshold
    EXPECT_GT(device_calls(), before);
    MatrixXd Lr = stan::math::cholesky_decompose_reference_cpu(A);
    EXPECT_LE(lower_rel_err(L, Lr), 1e-10) << n;
  }
}

TEST(CudaCholesky, same_exceptions_as_the_reference) {
  threshold_override force(0);
  MatrixXd not_square(2, 3);
  not_square << 1, 2, 3, 4, 5, 6;
  EXPECT_THROW(stan::math::cholesky_decompose(not_square), std::invalid_argument);
  MatrixXd asym = spd(80, 1);
  asym(3, 7) += 1.0;
  EXPECT_THROW(stan::math::cholesky_decompose(asym), std::domain_error);
  MatrixXd indefinite = spd(150, 2);
  indefinite(100, 100) = -5.0;
  std::string m_gpu, m_cpu;
  try {
    stan::math::cholesky_decompose(indefinite);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    stan::math::cholesky_decompose_reference_cpu(indefinite);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  MatrixXd with_nan = spd(90, 3);
  with_nan(50, 50) = std::numeric_limits<double>::quiet_NaN();
  EXPECT_THROW(stan::math::cholesky_decompose(with_nan), std::domain_error);
  matrix_v iv(150, 150);
  for (int i = 0; i < indefinite.size(); ++i) iv(i) = indefinite(i);
  EXPECT_THROW(stan::math::cholesky_decompose(iv), std::domain_error);
  MatrixXd empty(0, 0);
  EXPECT_EQ(0, stan::math::cholesky_decompose(empty).size());
  stan::math::recover_memory();
}

TEST(CudaCholesky, small_matrices_stay_on_the_host) {
  MatrixXd A = spd(100, 9);
  const uint64_t before = device_calls();
  MatrixXd L = stan::math::cholesky_decompose(A);
  EXPECT_EQ(before, device_calls());
  EXPECT_EQ(0.0, (L - stan::math::cholesky_decompose_reference_cpu(A)).cwiseAbs().maxCoeff());
}
#endif
