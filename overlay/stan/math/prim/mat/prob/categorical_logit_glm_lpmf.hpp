// Overlay of stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  std::vector<int> classes with a
// dynamic data matrix x and double / var alpha and beta go to the B200 path; scalar y, row-vector x,
// var x stay on the reference body.
#ifndef STAN_MATH_CUDA_OVERLAY_CATEGORICAL_LOGIT_GLM_LPMF_HPP
#define STAN_MATH_CUDA_OVERLAY_CATEGORICAL_LOGIT_GLM_LPMF_HPP

#define categorical_logit_glm_lpmf categorical_logit_glm_lpmf_reference_cpu
#include_next <stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp>
#undef categorical_logit_glm_lpmf

#ifdef STAN_CUDA
#include <stan/math/cuda/categorical_logit_glm_lpmf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
inline auto categorical_logit_glm(std::true_type, const T_y& y, const T_x& x, const T_alpha& alpha, const T_beta& beta)
    -> decltype(cuda::categorical_logit_glm_lpmf_device<propto>(y, x, alpha, beta)) {
  return cuda::categorical_logit_glm_lpmf_device<propto>(y, x, alpha, beta);
}
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_alpha_scalar, typename T_beta_scalar>
using categorical_logit_on_device
    = cuda::categorical_logit_glm_eligible<T_y, T_x_scalar, T_x_rows, T_alpha_scalar, T_beta_scalar>;
#else
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_alpha_scalar, typename T_beta_scalar>
using categorical_logit_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
inline auto categorical_logit_glm(std::false_type, const T_y& y, const T_x& x, const T_alpha& alpha,
                                  const T_beta& beta)
    -> decltype(categorical_logit_glm_lpmf_reference_cpu<propto>(y, x, alpha, beta)) {
  return categorical_logit_glm_lpmf_reference_cpu<propto>(y, x, alpha, beta);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_x_scalar, int T_x_rows, typename T_alpha_scalar,
          typename T_beta_scalar>
return_type_t<T_x_scalar, T_alpha_scalar, T_beta_scalar> categorical_logit_glm_lpmf(
    const T_y& y, const Eigen::Matrix<T_x_scalar, T_x_rows, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_alpha_scalar, Eigen::Dynamic, 1>& alpha,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, Eigen::Dynamic>& beta) {
  return cuda_dispatch::categorical_logit_glm<propto>(
      typename cuda_dispatch::categorical_logit_on_device<T_y, T_x_scalar, T_x_rows, T_alpha_scalar,
                                                          T_beta_scalar>::type(),
      y, x, alpha, beta);
}

template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_alpha_scalar, typename T_beta_scalar>
return_type_t<T_x_scalar, T_alpha_scalar, T_beta_scalar> categorical_logit_glm_lpmf(
    const T_y& y, const Eigen::Matrix<T_x_scalar, T_x_rows, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_alpha_scalar, Eigen::Dynamic, 1>& alpha,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, Eigen::Dynamic>& beta) {
  return categorical_logit_glm_lpmf<false>(y, x, alpha, beta);
}

}  // namespace math
}  // namespace stan
#endif
