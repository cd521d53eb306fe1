// Dense `multiply` on the device through the reference's own C++ signatures, against the unmodified
// reference overloads (reachable in the same binary as multiply_reference_cpu, see overlay/).  Modelled on the
// reference's OpenCL-vs-CPU tests (test/unit/math/rev/mat/fun/multiply_test.cpp:15-107: the same 234 x 234 case with
// the transfer threshold forced to 0 / INT_MAX, test/unit/math/opencl/multiply_transpose_test.cpp,
// matrix_cl_test.cpp, copy_test.cpp), at the north star's tolerance.
#ifdef STAN_CUDA
#include <stan/math/rev/mat.hpp>
#include <gtest/gtest.h>
#include <climits>
#include <random>
#include <stdexcept>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::MatrixXd;
using stan::math::var;
typedef Matrix<var, Dynamic, Dynamic> matrix_v;

namespace {

MatrixXd random_matrix(int rows, int cols, unsigned seed, double lo = -5, double hi = 5) {
  std::mt19937_64 rng(seed);
  std::uniform_real_distribution<double> u(lo, hi);
  MatrixXd m(rows, cols);
  for (int j = 0; j < cols; ++j)
    for (int i = 0; i < rows; ++i) m(i, j) = u(rng);
  return m;
}

// Componentwise backward-error gate of a product: |got - ref| <= 1e-10 * (|A| |B|)_ij -- the scale any summation
// order of the same products is accurate to; with `bound` = |ref| it is the plain relative gate.
void expect_product(const MatrixXd& got, const MatrixXd& ref, const MatrixXd& bound, const char* what) {
  ASSERT_EQ(got.rows(), ref.rows()) << what;
  ASSERT_EQ(got.cols(), ref.cols()) << what;
  double worst = 0;
  for (int i = 0; i < ref.size(); ++i) {
    const double den = bound(i) > 0 ? bound(i) : 1.0;
    worst = std::max(worst, std::fabs(got(i) - ref(i)) / den);
  }
  EXPECT_LE(worst, 1e-10) << what;
}

struct threshold_override {
  int64_t saved;
  explicit threshold_override(int64_t v) : saved(stan::math::cuda::tuning_opts().multiply_dim_prod_worth_transfer) {
    stan::math::cuda::tuning_opts().multiply_dim_prod_worth_transfer = v;
  }
  ~threshold_override() { stan::math::cuda::tuning_opts().multiply_dim_prod_worth_transfer = saved; }
};

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
MatrixXd val_of(const matrix_v& v) {
  MatrixXd a(v.rows(), v.cols());
  for (int i = 0; i < v.size(); ++i) a(i) = v(i).val();
  return a;
}

uint64_t device_evals() {
  stan_cuda_stats_t s;
  stan_cuda_stats(&s);
  return s.evals + s.kernel_launches;
}

}  // namespace

class CudaMultiply : public ::testing::TestWithParam<std::vector<int>> {};

TEST_P(CudaMultiply, prim_matches_reference) {
  const int m = GetParam()[0], k = GetParam()[1], n = GetParam()[2];
  MatrixXd A = random_matrix(m, k, 1 + m), B = random_matrix(k, n, 2 + n);
  threshold_override force(0);
  const uint64_t before = device_evals();
  MatrixXd C = stan::math::multiply(A, B);
  if (static_cast<int64_t>(m) * k * n >= 8 * (static_cast<int64_t>(m) * k + k * n + m * n) && m * k * n > 0) {
    EXPECT_GT(device_evals(), before) << "the product did not go to the device";
  }
  MatrixXd Cr = stan::math::multiply_reference_cpu(A, B);
  expect_product(C, Cr, A.cwiseAbs() * B.cwiseAbs(), "prim multiply");
}

TEST_P(CudaMultiply, rev_matches_reference_for_every_mix_of_var_and_double) {
  const int m = GetParam()[0], k = GetParam()[1], n = GetParam()[2];
  MatrixXd A = random_matrix(m, k, 3 + m), B = random_matrix(k, n, 4 + n), W = random_matrix(m, n, 5 + k, -1, 1);
  const MatrixXd bound_c = A.cwiseAbs() * B.cwiseAbs(), bound_a = W.cwiseAbs() * B.cwiseAbs().transpose(),
                 bound_b = A.cwiseAbs().transpose() * W.cwiseAbs();
  for (int mask = 1; mask <= 3; ++mask) {
    MatrixXd c[2], da[2], db[2];
    for (int side = 0; side < 2; ++side) {  // 0: device, 1: reference
      threshold_override force(side == 0 ? 0 : INT64_MAX);
      matrix_v Av = to_v(A), Bv = to_v(B), Cv;
      if (mask == 3) Cv = stan::math::multiply(Av, Bv);
      else if (mask == 1) Cv = stan::math::multiply(Av, B);
      else Cv = stan::math::multiply(A, Bv);
      var f = 0;
      for (int i = 0; i < Cv.size(); ++i) f += W(i) * Cv(i);
      f.grad();
      c[side] = val_of(Cv);
      da[side] = adj_of(Av);
      db[side] = adj_of(Bv);
      stan::math::recover_memory();
    }
    expect_product(c[0], c[1], bound_c, "rev multiply value");
    if (mask & 1) expect_product(da[0], da[1], bound_a, "rev multiply adjoint of A");
    if (mask & 2) expect_product(db[0], db[1], bound_b, "rev multiply adjoint of B");
  }
}

INSTANTIATE_TEST_CASE_P(Shapes, CudaMultiply,
                        ::testing::Values(std::vector<int>{234, 234, 234},   // the reference's OpenCL test size
                                          std::vector<int>{64, 64, 64}, std::vector<int>{128, 8, 128},
                                          std::vector<int>{129, 9, 127}, std::vector<int>{301, 257, 263},
                                          std::vector<int>{1000, 37, 50}, std::vector<int>{33, 1000, 65},
                                          std::vector<int>{17, 19, 23}, std::vector<int>{512, 512, 16}));

TEST(CudaMultiplyRouting, small_and_skinny_products_stay_on_the_host) {
  MatrixXd A = random_matrix(50, 40, 1), B = random_matrix(40, 30, 2);
  uint64_t before = device_evals();
  MatrixXd C = stan::math::multiply(A, B);  // 60000 < 2000000
  EXPECT_EQ(before, device_evals());
  EXPECT_EQ(0.0, (C - A * B).cwiseAbs().maxCoeff());
  MatrixXd T = random_matrix(3000, 1000, 3);
  Eigen::VectorXd v = Eigen::VectorXd::Random(1000);
  before = device_evals();
  Eigen::VectorXd tv = stan::math::multiply(T, v);  // 3e6 products but one per transferred element: the copy costs more
  EXPECT_EQ(before, device_evals());
  EXPECT_EQ(0.0, (tv - T * v).cwiseAbs().maxCoeff());
  MatrixXd A2 = random_matrix(300, 200, 4), B2 = random_matrix(200, 100, 5);
  before = device_evals();
  MatrixXd C2 = stan::math::multiply(A2, B2);  // 6e6: device
  EXPECT_GT(device_evals(), before);
}

TEST(CudaMultiplyRouting, same_exceptions_as_the_reference) {
  threshold_override force(0);
  MatrixXd A = random_matrix(100, 90, 1), B = random_matrix(91, 80, 2);
  EXPECT_THROW(stan::math::multiply(A, B), std::invalid_argument);
  matrix_v Av = to_v(A);
  EXPECT_THROW(stan::math::multiply(Av, B), std::invalid_argument);
  MatrixXd B_ok = random_matrix(90, 80, 3);
  matrix_v A_nan = to_v(A);
  A_nan(5, 7) = std::numeric_limits<double>::quiet_NaN();
  EXPECT_THROW(stan::math::multiply(A_nan, B_ok), std::domain_error);  // check_not_nan, rev/mat/fun/multiply.hpp:624-625
  std::string m_gpu, m_cpu;
  try {
    stan::math::multiply(A, B);
  } catch (const std::invalid_argument& e) {
    m_gpu = e.what();
  }
  try {
    stan::math::multiply_reference_cpu(A, B);
  } catch (const std::invalid_argument& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // empty operands: check_multiplicable rejects them (prim/mat/err/check_multiplicable.hpp: "must have a positive size")
  MatrixXd E0(0, 5), E1(5, 0);
  EXPECT_THROW(stan::math::multiply(E0, random_matrix(5, 3, 1)), std::invalid_argument);
  EXPECT_THROW(stan::math::multiply(E1, E0), std::invalid_argument);
  EXPECT_THROW(stan::math::multiply_reference_cpu(E1, E0), std::invalid_argument);
  stan::math::recover_memory();
}

// matrix_cuda: the matrix_cl surface the hooks use (test/unit/math/opencl/matrix_cl_test.cpp, copy_test.cpp,
// multiply_transpose_test.cpp)
TEST(CudaMatrix, copy_round_trip_and_products_with_transposes) {
  using stan::math::cuda::from_matrix_cuda;
  using stan::math::cuda::matrix_cuda;
  using stan::math::cuda::transpose;
  MatrixXd A = random_matrix(67, 45, 1), B = random_matrix(45, 81, 2), Bt = B.transpose(), At = A.transpose();
  matrix_cuda a(A), b(B), bt(Bt), at(At);
  EXPECT_EQ(67, a.rows());
  EXPECT_EQ(45, a.cols());
  EXPECT_EQ(0.0, (from_matrix_cuda(a) - A).cwiseAbs().maxCoeff());
  const MatrixXd ref = A * B, bound = A.cwiseAbs() * B.cwiseAbs();
  expect_product(from_matrix_cuda(a * b), ref, bound, "A * B");
  expect_product(from_matrix_cuda(a * transpose(bt)), ref, bound, "A * (B')'");
  expect_product(from_matrix_cuda(transpose(at) * b), ref, bound, "(A')' * B");
  expect_product(from_matrix_cuda(transpose(at) * transpose(bt)), ref, bound, "(A')' * (B')'");
  EXPECT_THROW(a * a, std::invalid_argument);
  matrix_cuda moved = std::move(a);
  EXPECT_EQ(67, moved.rows());
  matrix_cuda empty(0, 0);
  EXPECT_EQ(0, from_matrix_cuda(empty).size());
  Eigen::VectorXd v = Eigen::VectorXd::LinSpaced(45, -1, 1);
  matrix_cuda vd(v);
  Eigen::VectorXd mv = from_matrix_cuda<Dynamic, 1>(moved * vd);
  expect_product(mv, A * v, A.cwiseAbs() * v.cwiseAbs(), "matrix x vector through the handle API");
}
#endif
