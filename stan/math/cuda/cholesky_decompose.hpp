#ifndef STAN_MATH_CUDA_CHOLESKY_DECOMPOSE_HPP
#define STAN_MATH_CUDA_CHOLESKY_DECOMPOSE_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/multiply.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/mat/err/constraint_tolerance.hpp>
#include <cstdint>

namespace stan {
namespace math {
namespace cuda {

/**
 * Smallest matrix whose Cholesky factor (and whose reverse sweep) goes to the device: the knob the reference calls
 * cholesky_size_worth_transfer (stan/math/opencl/opencl_context.hpp:168, 1250 there).  The factorisation is n^3 / 3
 * flops on 8 n^2 transferred bytes per direction, so the break-even against one host core sits far lower here.
 */
inline int64_t& cholesky_size_worth_transfer() {
  static int64_t n = 400;
  return n;
}

/**
 * L = chol(m) on the device (lower, zeros above the diagonal), with the reference's check_symmetric
 * (prim/mat/err/check_symmetric.hpp, CONSTRAINT_TOLERANCE) evaluated there as well.  Returns the code of the C ABI:
 * STAN_CUDA_OK, STAN_CUDA_NOT_SYMMETRIC (the caller runs the reference's check_symmetric on the host: it throws the
 * reference's exception) or STAN_CUDA_NOT_POS_DEF (the caller runs the reference's LLT + check_pos_definite).
 */
inline int cholesky_on_device(const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& m,
                              Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& L) {
  L.resize(m.rows(), m.cols());
  const int rc = stan_cuda_cholesky_decompose(m.rows(), m.data(), m.rows(), L.data(), L.rows(), CONSTRAINT_TOLERANCE);
  check_cuda_rc("cholesky_decompose", rc);
  return rc;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
