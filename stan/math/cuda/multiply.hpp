#ifndef STAN_MATH_CUDA_MULTIPLY_HPP
#define STAN_MATH_CUDA_MULTIPLY_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/matrix_cuda.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <cstdint>

namespace stan {
namespace math {
namespace cuda {

/**
 * When a dense product goes to the device.  The reference's rule (stan/math/opencl/opencl_context.hpp:179,
 * multiply_dim_prod_worth_transfer = 2000000, applied at prim/mat/fun/multiply.hpp:72-73 and
 * rev/mat/fun/multiply.hpp:88-89) plus one more condition: both operands and the result cross PCIe on every call
 * (8 (mk + kn + mn) bytes against 2 mnk flops), so shapes with fewer than `multiply_flops_per_element` products per
 * transferred element -- matrix x vector and the like -- stay on the host, where one pass of Eigen is faster than
 * the copy.  Both knobs are plain variables like the reference's tuning_opts().
 */
struct tuning_struct {
  int64_t multiply_dim_prod_worth_transfer = 2000000;
  int64_t multiply_flops_per_element = 8;
};
inline tuning_struct& tuning_opts() {
  static tuning_struct t;
  return t;
}
inline bool multiply_worth_transfer(int64_t m, int64_t k, int64_t n) {
  const tuning_struct& t = tuning_opts();
  return m * k * n > t.multiply_dim_prod_worth_transfer && m * k * n >= t.multiply_flops_per_element * (m * k + k * n + m * n);
}

/** host matrices in, host matrix out: C = op(A) * op(B) in one call of the C ABI (upload, product, download) */
inline void multiply_on_device(int64_t m, int64_t n, int64_t k, const double* a, int64_t lda, bool trans_a, const double* b,
                               int64_t ldb, bool trans_b, double* c, int64_t ldc) {
  stan_cuda_multiply_args g = stan_cuda_multiply_args();
  g.m = m;
  g.n = n;
  g.k = k;
  g.a = a;
  g.lda = lda;
  g.trans_a = trans_a ? 1 : 0;
  g.b = b;
  g.ldb = ldb;
  g.trans_b = trans_b ? 1 : 0;
  g.c = c;
  g.ldc = ldc;
  check_cuda_rc("multiply", stan_cuda_multiply(&g));
}

/** GPU body of multiply(m1, m2) for double matrices (stan/math/prim/mat/fun/multiply.hpp:71-79) */
template <int R1, int C1, int R2, int C2>
inline Eigen::Matrix<double, R1, C2> multiply_device(const Eigen::Matrix<double, R1, C1>& m1,
                                                     const Eigen::Matrix<double, R2, C2>& m2) {
  Eigen::Matrix<double, R1, C2> out(m1.rows(), m2.cols());
  multiply_on_device(m1.rows(), m2.cols(), m1.cols(), m1.data(), m1.rows(), false, m2.data(), m2.rows(), false, out.data(),
                     out.rows());
  return out;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
