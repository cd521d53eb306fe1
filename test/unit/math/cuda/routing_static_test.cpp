// Instantiation routing (SURVEY 8b "Which instantiations take the GPU path"), checked at COMPILE time: which
// argument kinds of the reference's public templates the overlay sends to the device wrappers and which stay on
// the reference body (fvar, var data matrix, scalar operands of the elementwise densities, ...).  The binary has
// no device work: it runs on the CPU-only container as part of `pytest -m "not gpu"`.
#ifdef STAN_CUDA
#include <stan/math/mix/mat.hpp>
#include <cstdio>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using stan::math::fvar;
using stan::math::var;
namespace cu = stan::math::cuda;

typedef Matrix<double, Dynamic, Dynamic> matrix_d;
typedef Matrix<var, Dynamic, Dynamic> matrix_v;
typedef Matrix<double, Dynamic, 1> vector_d;
typedef Matrix<var, Dynamic, 1> vector_v;
typedef Matrix<var, 1, Dynamic> row_vector_v;
typedef Matrix<fvar<double>, Dynamic, 1> vector_fd;
typedef Matrix<int, Dynamic, 1> vector_i;
typedef std::vector<int> ints;

// bernoulli_logit_glm / poisson_log_glm (bernoulli_logit_glm_lpmf.hpp:46-49, interface test :342-382)
static_assert(cu::int_glm_eligible<ints, matrix_d, var, vector_v>::value, "var alpha, var beta vector");
static_assert(cu::int_glm_eligible<ints, matrix_d, double, vector_d>::value, "all double");
static_assert(cu::int_glm_eligible<int, matrix_d, var, row_vector_v>::value, "scalar y, row vector beta");
static_assert(cu::int_glm_eligible<vector_i, matrix_d, vector_v, std::vector<var>>::value, "Eigen int y, vector alpha");
static_assert(cu::int_glm_eligible<ints, cu::data_matrix_view, var, vector_v>::value, "zero-copy view");
static_assert(cu::int_glm_eligible<ints, matrix_v, var, vector_v>::value, "var x (column-major): device path, N x K partials");
static_assert(!cu::int_glm_eligible<ints, Matrix<var, Dynamic, Dynamic, Eigen::RowMajor>, var, vector_v>::value, "row-major var x stays on the reference body");
static_assert(!cu::int_glm_eligible<ints, Matrix<fvar<double>, Dynamic, Dynamic>, double, vector_d>::value, "fvar x stays on the reference body");
static_assert(!cu::int_glm_eligible<ints, matrix_d, fvar<double>, vector_fd>::value, "fvar stays on the reference body");
static_assert(!cu::int_glm_eligible<vector_d, matrix_d, var, vector_v>::value, "real outcomes are not int outcomes");
// normal_id_glm (normal_id_glm_lpdf.hpp:49-53, interface test :441-478)
static_assert(cu::normal_glm_eligible<vector_d, matrix_d, var, vector_v, var>::value, "C4 signature");
static_assert(cu::normal_glm_eligible<vector_v, matrix_d, vector_v, vector_v, vector_v>::value, "var y, vector alpha/sigma");
static_assert(cu::normal_glm_eligible<double, matrix_d, double, vector_d, double>::value, "scalar y");
static_assert(cu::normal_glm_eligible<vector_d, matrix_v, var, vector_v, var>::value, "var x: device path");
static_assert(!cu::normal_glm_eligible<vector_fd, matrix_d, double, vector_d, double>::value, "fvar y");
// neg_binomial_2_log_glm
static_assert(cu::neg_binomial_glm_eligible<ints, matrix_d, var, vector_v, var>::value, "scalar phi");
static_assert(cu::neg_binomial_glm_eligible<ints, matrix_d, vector_v, vector_v, vector_v>::value, "vector alpha / phi");
static_assert(cu::neg_binomial_glm_eligible<ints, matrix_v, var, vector_v, var>::value, "var x: device path");
// ordered_logistic_glm / categorical_logit_glm (std::vector<int> y, dynamic data matrix)
static_assert(cu::ordered_logistic_glm_eligible<ints, double, Dynamic, var, var>::value, "var beta, var cuts");
static_assert(cu::ordered_logistic_glm_eligible<ints, double, Dynamic, double, double>::value, "all double");
static_assert(!cu::ordered_logistic_glm_eligible<ints, var, Dynamic, var, var>::value, "var x");
static_assert(!cu::ordered_logistic_glm_eligible<int, double, Dynamic, var, var>::value, "scalar y stays on the reference body");
static_assert(!cu::ordered_logistic_glm_eligible<ints, double, 1, var, var>::value, "row-vector x stays on the reference body");
static_assert(cu::categorical_logit_glm_eligible<ints, double, Dynamic, var, var>::value, "var alpha, var beta");
static_assert(!cu::categorical_logit_glm_eligible<ints, double, Dynamic, fvar<double>, double>::value, "fvar alpha");
static_assert(!cu::categorical_logit_glm_eligible<ints, var, Dynamic, var, var>::value, "var x");
// vectorised normal_lpdf: at least one container operand (normal_lpdf.hpp:36-39)
static_assert(cu::normal_lpdf_eligible<vector_d, vector_v, var>::value, "C1 signature");
static_assert(cu::normal_lpdf_eligible<std::vector<double>, double, std::vector<var>>::value, "vector sigma");
static_assert(cu::normal_lpdf_eligible<vector_v, vector_v, vector_v>::value, "all var vectors");
static_assert(!cu::normal_lpdf_eligible<double, var, var>::value, "scalar call stays on the reference loop");
static_assert(!cu::normal_lpdf_eligible<vector_d, vector_fd, double>::value, "fvar");
// vectorised student_t_lpdf: at least one container operand (student_t_lpdf.hpp:52-55)
static_assert(cu::student_t_lpdf_eligible<vector_d, var, vector_v, var>::value, "the model statement");
static_assert(cu::student_t_lpdf_eligible<std::vector<double>, std::vector<var>, double, std::vector<var>>::value, "vector nu, sigma");
static_assert(cu::student_t_lpdf_eligible<vector_v, vector_v, vector_v, vector_v>::value, "all var vectors");
static_assert(cu::student_t_lpdf_eligible<double, double, vector_d, double>::value, "all data");
static_assert(!cu::student_t_lpdf_eligible<double, var, var, var>::value, "scalar call stays on the reference loops");
static_assert(!cu::student_t_lpdf_eligible<vector_d, vector_fd, double, double>::value, "fvar");
// vectorised cauchy_lpdf / lognormal_lpdf: the operand kinds of normal_lpdf, and something to compute
static_assert(cu::locscale_lpdf_eligible<false, vector_d, vector_v, var>::value, "the model statement");
static_assert(cu::locscale_lpdf_eligible<true, vector_d, vector_v, var>::value, "propto with variables");
static_assert(cu::locscale_lpdf_eligible<false, vector_d, vector_d, double>::value, "all data, not propto");
static_assert(!cu::locscale_lpdf_eligible<true, vector_d, vector_d, double>::value, "propto, all data: only the reference's checks run");
static_assert(!cu::locscale_lpdf_eligible<false, double, var, var>::value, "scalar call stays on the reference loop");
static_assert(!cu::locscale_lpdf_eligible<false, vector_d, vector_fd, double>::value, "fvar");
// vectorised bernoulli_logit_lpmf / poisson_log_lpmf / poisson_lpmf / bernoulli_lpmf: container operand of double / var
static_assert(cu::int_density_eligible<ints, vector_v>::value, "d1 signature");
static_assert(cu::int_density_eligible<int, std::vector<var>>::value, "scalar n");
static_assert(cu::int_density_eligible<vector_i, vector_d>::value, "Eigen int n, double operand");
static_assert(cu::int_density_eligible<ints, row_vector_v>::value, "row vector operand");
static_assert(!cu::int_density_eligible<ints, var>::value, "scalar operand stays on the reference loop");
static_assert(!cu::int_density_eligible<ints, vector_fd>::value, "fvar");
static_assert(!cu::int_density_eligible<vector_d, vector_v>::value, "real n is not an integer outcome");

// the public names keep their return types (return_type_t of the non-data arguments)
static_assert(std::is_same<decltype(stan::math::bernoulli_logit_glm_lpmf(ints(), matrix_d(), var(), vector_v())), var>::value, "");
static_assert(std::is_same<decltype(stan::math::poisson_log_glm_lpmf(ints(), matrix_d(), 0.0, vector_d())), double>::value, "");
static_assert(std::is_same<decltype(stan::math::normal_id_glm_lpdf(vector_d(), matrix_d(), 0.0, vector_d(), var())), var>::value, "");
static_assert(std::is_same<decltype(stan::math::normal_lpdf(vector_d(), vector_v(), 1.0)), var>::value, "");
static_assert(std::is_same<decltype(stan::math::bernoulli_logit_lpmf(ints(), vector_d())), double>::value, "");
static_assert(std::is_same<decltype(stan::math::poisson_log_lpmf(ints(), vector_fd())), fvar<double>>::value, "");
static_assert(std::is_same<decltype(stan::math::poisson_lpmf(ints(), vector_v())), var>::value, "");
static_assert(std::is_same<decltype(stan::math::cauchy_lpdf(vector_d(), vector_v(), 1.0)), var>::value, "");
static_assert(std::is_same<decltype(stan::math::gamma_lpdf(vector_d(), var(), vector_v())), var>::value, "");
static_assert(std::is_same<decltype(stan::math::beta_lpdf(vector_v(), 2.0, 3.0)), var>::value, "");
static_assert(std::is_same<decltype(stan::math::weibull_lpdf<true>(vector_d(), var(), vector_d())), var>::value, "");
static_assert(std::is_same<decltype(stan::math::exponential_lpdf(vector_d(), vector_fd())), fvar<double>>::value, "");
static_assert(std::is_same<decltype(stan::math::exponential_lpdf<true>(vector_v(), 1.5)), var>::value, "");
static_assert(std::is_same<decltype(stan::math::lognormal_lpdf<true>(vector_d(), vector_d(), 1.0)), double>::value, "");
static_assert(std::is_same<decltype(stan::math::lognormal_lpdf(vector_d(), vector_fd(), 1.0)), fvar<double>>::value, "");
static_assert(std::is_same<decltype(stan::math::poisson_lpmf<true>(ints(), vector_d())), double>::value, "");
static_assert(std::is_same<decltype(stan::math::bernoulli_lpmf(ints(), vector_fd())), fvar<double>>::value, "");
static_assert(std::is_same<decltype(stan::math::bernoulli_lpmf(1, std::vector<var>())), var>::value, "");

int main() {
  // anything not eligible runs the reference body in this very binary: no device is touched
  std::vector<int> y{1, 0, 1};
  const double lp = stan::math::bernoulli_logit_lpmf(y, 0.3);  // scalar operand -> reference loop
  vector_fd th(3);
  th << fvar<double>(0.1, 1.0), fvar<double>(0.2, 1.0), fvar<double>(-0.3, 1.0);
  const fvar<double> lf = stan::math::bernoulli_logit_lpmf(y, th);  // fvar -> reference loop
  const double lq = stan::math::bernoulli_lpmf(y, 0.3) + stan::math::poisson_lpmf(y, 1.5);  // scalar operands -> reference loops
  std::printf("routing_static_test PASSED (%.6f %.6f %.6f %.6f)\n", lp, lf.val_, lf.d_, lq);
  return (lp < 0 && lf.val_ < 0 && lq < 0) ? 0 : 1;
}
#else
int main() { return 1; }
#endif
