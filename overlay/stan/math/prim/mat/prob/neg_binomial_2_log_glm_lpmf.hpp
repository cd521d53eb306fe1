// Overlay of stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  Data matrix x, integer outcomes
// and double / var alpha, beta, phi go to the B200 path; everything else stays on the reference body.
#ifndef STAN_MATH_CUDA_OVERLAY_NEG_BINOMIAL_2_LOG_GLM_LPMF_HPP
#define STAN_MATH_CUDA_OVERLAY_NEG_BINOMIAL_2_LOG_GLM_LPMF_HPP

#define neg_binomial_2_log_glm_lpmf neg_binomial_2_log_glm_lpmf_reference_cpu
#include_next <stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp>
#undef neg_binomial_2_log_glm_lpmf

#ifdef STAN_CUDA
#include <stan/math/cuda/neg_binomial_2_log_glm_lpmf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
inline return_type_t<T_x, T_alpha, T_beta, T_precision> neg_binomial_2_log_glm(std::true_type, const T_y& y,
                                                                               const T_x& x, const T_alpha& alpha,
                                                                               const T_beta& beta,
                                                                               const T_precision& phi) {
  return cuda::neg_binomial_2_log_glm_lpmf_device<propto>(y, x, alpha, beta, phi);
}
template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
using neg_binomial_on_device = cuda::neg_binomial_glm_eligible<T_y, T_x, T_alpha, T_beta, T_precision>;
#else
template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
using neg_binomial_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
inline return_type_t<T_x, T_alpha, T_beta, T_precision> neg_binomial_2_log_glm(std::false_type, const T_y& y,
                                                                               const T_x& x, const T_alpha& alpha,
                                                                               const T_beta& beta,
                                                                               const T_precision& phi) {
  return neg_binomial_2_log_glm_lpmf_reference_cpu<propto>(y, x, alpha, beta, phi);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
return_type_t<T_x, T_alpha, T_beta, T_precision> neg_binomial_2_log_glm_lpmf(const T_y& y, const T_x& x,
                                                                             const T_alpha& alpha,
                                                                             const T_beta& beta,
                                                                             const T_precision& phi) {
  return cuda_dispatch::neg_binomial_2_log_glm<propto>(
      typename cuda_dispatch::neg_binomial_on_device<T_y, T_x, T_alpha, T_beta, T_precision>::type(), y, x, alpha,
      beta, phi);
}

template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
inline return_type_t<T_x, T_alpha, T_beta, T_precision> neg_binomial_2_log_glm_lpmf(const T_y& y, const T_x& x,
                                                                                    const T_alpha& alpha,
                                                                                    const T_beta& beta,
                                                                                    const T_precision& phi) {
  return neg_binomial_2_log_glm_lpmf<false>(y, x, alpha, beta, phi);
}

}  // namespace math
}  // namespace stan
#endif
