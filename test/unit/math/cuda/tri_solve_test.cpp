// mdivide_left_tri on the device through the reference's own C++ signatures, against the unmodified
// reference overloads (mdivide_left_tri_reference_cpu, see overlay/).  Modelled on the reference's OpenCL-vs-CPU tests
// (test/unit/math/rev/mat/fun/mdivide_left_tri_test.cpp:20-360: 256 x 256, entries uniform(-5, 5), diagonal 20, the
// transfer threshold forced to 0 / INT_MAX, value and adjoints for vv / dv / vd, lower and upper) at the north
// star's tolerance.
#ifdef STAN_CUDA
#include <stan/math/rev/mat.hpp>
#include <gtest/gtest.h>
#include <random>
#include <stdexcept>
#include <vector>

using Eigen::MatrixXd;
using stan::math::var;
typedef Eigen::Matrix<var, -1, -1> matrix_v;

namespace {

struct threshold_override {
  int64_t saved;
  explicit threshold_override(int64_t v) : saved(stan::math::cuda::tri_size_worth_transfer()) {
    stan::math::cuda::tri_size_worth_transfer() = v;
  }
  ~threshold_override() { stan::math::cuda::tri_size_worth_transfer() = saved; }
};

MatrixXd uniform_matrix(int rows, int cols, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::uniform_real_distribution<double> u(-5, 5);
  MatrixXd m(rows, cols);
  for (int i = 0; i < rows * cols; ++i) m(i) = u(rng);
  return m;
}

// full square matrix whose relevant triangle is well conditioned (the reference's fixture: diagonal 20); the other
// triangle holds data that must not be read
MatrixXd tri_fixture(int n, unsigned seed) {
  MatrixXd a = uniform_matrix(n, n, seed);
  const double diag = std::max(20.0, 0.6 * n);
  for (int i = 0; i < n; ++i) a(i, i) = diag;
  return a;
}

double rel_err(const MatrixXd& got, const MatrixXd& ref) {
  const double scale = ref.cwiseAbs().maxCoeff();
  double worst = 0;
  for (int i = 0; i < ref.size(); ++i)
    worst = std::max(worst, std::fabs(got(i) - ref(i)) / std::max(std::fabs(ref(i)), 1e-6 * scale));
  return worst;
}

uint64_t device_calls() {
  stan_cuda_stats_t s;
  stan_cuda_stats(&s);
  return s.evals + s.kernel_launches;
}

matrix_v to_v(const MatrixXd& m) {
  matrix_v v(m.rows(), m.cols());
  for (int i = 0; i < m.size(); ++i) v(i) = m(i);
  return v;
}
MatrixXd adj_of(const matrix_v& v) {
  MatrixXd a(v.rows(), v.cols());
  for (int i = 0; i < v.size(); ++i) a(i) = v(i).adj();
  return a;
}

template <Eigen::UpLoType TriView>
void check_shape(int n, int m) {
  MatrixXd A = tri_fixture(n, 10 + n), B = uniform_matrix(n, m, 20 + m), W = uniform_matrix(n, m, 30 + n + m);
  {
    threshold_override force(0);
    const uint64_t before = device_calls();
    MatrixXd X = stan::math::mdivide_left_tri<TriView>(A, B);
    if (m >= 8) EXPECT_GT(device_calls(), before);
    EXPECT_LE(rel_err(X, stan::math::mdivide_left_tri_reference_cpu<TriView>(A, B)), 1e-10) << n << " x " << m;
  }
  for (int mask = 1; mask <= 3; ++mask) {
    MatrixXd val[2], da[2], db[2];
    for (int side = 0; side < 2; ++side) {
      threshold_override force(side == 0 ? 0 : INT64_MAX);
      matrix_v Av = to_v(A), Bv = to_v(B), Xv;
      if (mask == 3) Xv = stan::math::mdivide_left_tri<TriView>(Av, Bv);
      else if (mask == 1) Xv = stan::math::mdivide_left_tri<TriView>(Av, B);
      else Xv = stan::math::mdivide_left_tri<TriView>(A, Bv);
      var f = 0;
      for (int i = 0; i < Xv.size(); ++i) f += W(i) * Xv(i);
      f.grad();
      val[side].resize(n, m);
      for (int i = 0; i < n * m; ++i) val[side](i) = Xv(i).val();
      da[side] = adj_of(Av);
      db[side] = adj_of(Bv);
      stan::math::recover_memory();
    }
    EXPECT_LE(rel_err(val[0], val[1]), 1e-10) << "value, mask " << mask;
    if (mask & 1) {
      // adjA = -adjB X' (:100-101) is a product: the componentwise gate of a product, |got - ref| <= 1e-10 (|adjB| |X|')_ij
      // (its entries span many orders of magnitude and each one is a sum with cancellation)
      MatrixXd adjB = A.template triangularView<TriView>().transpose().solve(W);
      MatrixXd bound = adjB.cwiseAbs() * val[1].cwiseAbs().transpose();
      double worst = 0;
      for (int i = 0; i < n * n; ++i) worst = std::max(worst, std::fabs(da[0](i) - da[1](i)) / std::max(bound(i), 1e-300));
      EXPECT_LE(worst, 1e-10) << "adjoint of A, mask " << mask;
      // and nothing outside the triangle that is an operand (:45-62)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
          if (TriView == Eigen::Lower ? i < j : i > j) EXPECT_EQ(0.0, da[0](i, j));
    }
    if (mask & 2) EXPECT_LE(rel_err(db[0], db[1]), 1e-10) << "adjoint of B, mask " << mask;
  }
}

}  // namespace

class CudaTriSolve : public ::testing::TestWithParam<std::pair<int, int>> {};

TEST_P(CudaTriSolve, lower_matches_reference) { check_shape<Eigen::Lower>(GetParam().first, GetParam().second); }
TEST_P(CudaTriSolve, upper_matches_reference) { check_shape<Eigen::Upper>(GetParam().first, GetParam().second); }

INSTANTIATE_TEST_CASE_P(Shapes, CudaTriSolve,
                        ::testing::Values(std::make_pair(256, 256),  // the reference's OpenCL test size
                                          std::make_pair(1, 8), std::make_pair(63, 9), std::make_pair(64, 64), std::make_pair(65, 17),
                                          std::make_pair(130, 300), std::make_pair(513, 40), std::make_pair(700, 8)));

TEST(CudaTriSolveRouting, mdivide_right_tri_matches_reference) {
  // x = b A^-1 (prim/mat/fun/mdivide_right_tri.hpp:57-80), solved on the device as x' = A^-T b'
  for (int n : {256, 300, 517}) {
    MatrixXd A = tri_fixture(n, 40 + n), b = uniform_matrix(33, n, 50 + n);
    const uint64_t before = device_calls();
    MatrixXd xl = stan::math::mdivide_right_tri<Eigen::Lower>(b, A), xu = stan::math::mdivide_right_tri<Eigen::Upper>(b, A);
    EXPECT_GT(device_calls(), before);
    EXPECT_LE(rel_err(xl, stan::math::mdivide_right_tri_reference_cpu<Eigen::Lower>(b, A)), 1e-10) << n;
    EXPECT_LE(rel_err(xu, stan::math::mdivide_right_tri_reference_cpu<Eigen::Upper>(b, A)), 1e-10) << n;
    MatrixXd low = stan::math::mdivide_right_tri_low(b, A);
    EXPECT_LE(rel_err(low, xl), 1e-14);
  }
  MatrixXd A5 = tri_fixture(5, 1), b_bad = uniform_matrix(9, 4, 2);
  EXPECT_THROW(stan::math::mdivide_right_tri<Eigen::Lower>(b_bad, A5), std::invalid_argument);
  // var arguments stay on the reference's generic template
  matrix_v Av = to_v(tri_fixture(6, 3)), bv = to_v(uniform_matrix(2, 6, 4));
  matrix_v xv = stan::math::mdivide_right_tri<Eigen::Lower>(bv, Av);
  EXPECT_EQ(2, xv.rows());
  stan::math::recover_memory();
}

TEST(CudaTriSolveRouting, inverse_small_systems_and_exceptions) {
  MatrixXd A = tri_fixture(300, 1);
  MatrixXd inv = stan::math::mdivide_left_tri<Eigen::Lower>(A);  // 300 >= 256: device
  EXPECT_LE(rel_err(inv, stan::math::mdivide_left_tri_reference_cpu<Eigen::Lower>(A)), 1e-10);
  MatrixXd low = stan::math::mdivide_left_tri_low(A, uniform_matrix(300, 12, 2));
  EXPECT_LE(rel_err(low, stan::math::mdivide_left_tri_reference_cpu<Eigen::Lower>(A, uniform_matrix(300, 12, 2))), 1e-10);
  // below the threshold, and single right-hand sides, stay on the host
  MatrixXd small = tri_fixture(50, 3), bs = uniform_matrix(50, 50, 4);
  uint64_t before = device_calls();
  MatrixXd xs = stan::math::mdivide_left_tri<Eigen::Upper>(small, bs);
  Eigen::VectorXd v = Eigen::VectorXd::Random(300);
  Eigen::VectorXd xv = stan::math::mdivide_left_tri<Eigen::Lower>(A, v);
  EXPECT_EQ(before, device_calls());
  EXPECT_EQ(0.0, (xs - stan::math::mdivide_left_tri_reference_cpu<Eigen::Upper>(small, bs)).cwiseAbs().maxCoeff());
  // the reference's size checks (:61-62)
  threshold_override force(0);
  MatrixXd not_square(3, 4), b3 = uniform_matrix(3, 9, 5);
  EXPECT_THROW(stan::math::mdivide_left_tri<Eigen::Lower>(not_square, b3), std::invalid_argument);
  EXPECT_THROW(stan::math::mdivide_left_tri<Eigen::Lower>(tri_fixture(5, 6), b3), std::invalid_argument);
  matrix_v Av = to_v(tri_fixture(5, 6));
  EXPECT_THROW(stan::math::mdivide_left_tri<Eigen::Lower>(Av, b3), std::invalid_argument);
  stan::math::recover_memory();
}
#endif
