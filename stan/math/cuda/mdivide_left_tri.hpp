#ifndef STAN_MATH_CUDA_MDIVIDE_LEFT_TRI_HPP
#define STAN_MATH_CUDA_MDIVIDE_LEFT_TRI_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <cstdint>

namespace stan {
namespace math {
namespace cuda {

/**
 * Smallest triangular system that goes to the device: the knob the reference calls tri_inverse_size_worth_transfer
 * (stan/math/opencl/opencl_context.hpp:182, 100 there).  Substitution is n^2 m flops on 8 (n^2 + 2 n m) transferred
 * bytes; the default keeps systems that one host core solves faster than the copy at home.
 */
inline int64_t& tri_size_worth_transfer() {
  static int64_t n = 256;
  return n;
}
inline bool tri_solve_worth_transfer(int64_t n, int64_t m) { return n >= tri_size_worth_transfer() && m >= 8; }

/** x = op(A)^-1 b on the device; A is n x n column-major, b and x are n x m */
inline void tri_solve_on_device(bool lower, bool trans, int64_t n, int64_t m, const double* a, const double* b, double* x) {
  check_cuda_rc("mdivide_left_tri", stan_cuda_tri_solve(lower ? 1 : 0, trans ? 1 : 0, n, m, a, n, b, n, x, n));
}

/** GPU body of mdivide_left_tri<TriView>(A, b) for double matrices (stan/math/prim/mat/fun/mdivide_left_tri.hpp:58-77) */
template <int TriView, int R1, int C1, int R2, int C2>
inline Eigen::Matrix<double, R1, C2> mdivide_left_tri_device(const Eigen::Matrix<double, R1, C1>& A,
                                                             const Eigen::Matrix<double, R2, C2>& b) {
  Eigen::Matrix<double, R1, C2> x(b.rows(), b.cols());
  tri_solve_on_device(TriView == Eigen::Lower, false, A.rows(), b.cols(), A.data(), b.data(), x.data());
  return x;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
