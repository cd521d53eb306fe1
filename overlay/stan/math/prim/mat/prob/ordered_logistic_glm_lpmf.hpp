// Overlay of stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  std::vector<int> classes with a
// dynamic data matrix x and double / var beta and cut-points go to the B200 path; scalar y, row-vector
// x, var x stay on the reference body.
#ifndef STAN_MATH_CUDA_OVERLAY_ORDERED_LOGISTIC_GLM_LPMF_HPP
#define STAN_MATH_CUDA_OVERLAY_ORDERED_LOGISTIC_GLM_LPMF_HPP

#define ordered_logistic_glm_lpmf ordered_logistic_glm_lpmf_reference_cpu
#include_next <stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp>
#undef ordered_logistic_glm_lpmf

#ifdef STAN_CUDA
#include <stan/math/cuda/ordered_logistic_glm_lpmf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_x, typename T_beta, typename T_cuts>
inline auto ordered_logistic_glm(std::true_type, const T_y& y, const T_x& x, const T_beta& beta, const T_cuts& cuts)
    -> decltype(cuda::ordered_logistic_glm_lpmf_device<propto>(y, x, beta, cuts)) {
  return cuda::ordered_logistic_glm_lpmf_device<propto>(y, x, beta, cuts);
}
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_beta_scalar, typename T_cuts_scalar>
using ordered_logistic_on_device
    = cuda::ordered_logistic_glm_eligible<T_y, T_x_scalar, T_x_rows, T_beta_scalar, T_cuts_scalar>;
#else
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_beta_scalar, typename T_cuts_scalar>
using ordered_logistic_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_x, typename T_beta, typename T_cuts>
inline auto ordered_logistic_glm(std::false_type, const T_y& y, const T_x& x, const T_beta& beta, const T_cuts& cuts)
    -> decltype(ordered_logistic_glm_lpmf_reference_cpu<propto>(y, x, beta, cuts)) {
  return ordered_logistic_glm_lpmf_reference_cpu<propto>(y, x, beta, cuts);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_x_scalar, int T_x_rows, typename T_beta_scalar,
          typename T_cuts_scalar>
typename stan::return_type_t<T_x_scalar, T_beta_scalar, T_cuts_scalar> ordered_logistic_glm_lpmf(
    const T_y& y, const Eigen::Matrix<T_x_scalar, T_x_rows, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, 1>& beta,
    const Eigen::Matrix<T_cuts_scalar, Eigen::Dynamic, 1>& cuts) {
  return cuda_dispatch::ordered_logistic_glm<propto>(
      typename cuda_dispatch::ordered_logistic_on_device<T_y, T_x_scalar, T_x_rows, T_beta_scalar,
                                                         T_cuts_scalar>::type(),
      y, x, beta, cuts);
}

template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_beta_scalar, typename T_cuts_scalar>
typename stan::return_type_t<T_x_scalar, T_beta_scalar, T_cuts_scalar> ordered_logistic_glm_lpmf(
    const T_y& y, const Eigen::Matrix<T_x_scalar, T_x_rows, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, 1>& beta,
    const Eigen::Matrix<T_cuts_scalar, Eigen::Dynamic, 1>& cuts) {
  return ordered_logistic_glm_lpmf<false>(y, x, beta, cuts);
}

}  // namespace math
}  // namespace stan
#endif
