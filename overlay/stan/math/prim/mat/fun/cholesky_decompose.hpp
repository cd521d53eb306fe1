// Overlay of stan/math/prim/mat/fun/cholesky_decompose.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  The reference's own transparent GPU hook sits in
// the double specialisation (:54-58, #ifdef STAN_OPENCL: matrices of at least cholesky_size_worth_transfer rows go
// to opencl::cholesky_decompose); with -DSTAN_CUDA the same decision sends the factorisation to the B200.  Both
// reference functions are kept under the name cholesky_decompose_reference_cpu.
#ifndef STAN_MATH_CUDA_OVERLAY_PRIM_CHOLESKY_DECOMPOSE_HPP
#define STAN_MATH_CUDA_OVERLAY_PRIM_CHOLESKY_DECOMPOSE_HPP

#define cholesky_decompose cholesky_decompose_reference_cpu
#include_next <stan/math/prim/mat/fun/cholesky_decompose.hpp>
#undef cholesky_decompose

#ifdef STAN_CUDA
#include <stan/math/cuda/cholesky_decompose.hpp>
#include <stan/math/prim/mat/err/check_square.hpp>
#include <stan/math/prim/mat/err/check_symmetric.hpp>
#endif

namespace stan {
namespace math {

template <typename T>
inline Eigen::Matrix<T, Eigen::Dynamic, Eigen::Dynamic> cholesky_decompose(
    const Eigen::Matrix<T, Eigen::Dynamic, Eigen::Dynamic>& m) {
  return cholesky_decompose_reference_cpu(m);
}

#ifdef STAN_CUDA
template <>
inline Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> cholesky_decompose(
    const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& m) {
  if (m.rows() >= cuda::cholesky_size_worth_transfer() && m.rows() == m.cols()) {
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> L;
    if (cuda::cholesky_on_device(m, L) == STAN_CUDA_OK) {  // symmetric (:60, checked on the device) and positive definite
      return L;
    }
    // not symmetric or not positive definite: the reference's own checks, in its order, throw its exception
  }
  return cholesky_decompose_reference_cpu(m);
}
#endif

}  // namespace math
}  // namespace stan
#endif
