// Overlay of stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp.
//
// Put `-I <repo>/overlay -I <repo>` in front of the Stan Math include path: every
// `#include <stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp>` (stan/math/prim/mat.hpp:269)
// then finds this file first.  It pulls in the UNMODIFIED reference header with #include_next,
// under a different function name, and defines stan::math::bernoulli_logit_glm_lpmf with the
// reference's exact signature: instantiations with a data matrix x and integer outcomes go to the
// B200 path, every other instantiation (fvar, var x, real-valued y ...) is forwarded to the
// reference body.  Without -DSTAN_CUDA everything is forwarded, so the overlay is inert.
// The equivalent in-tree patch is shown in INTEGRATION.md.
#ifndef STAN_MATH_CUDA_OVERLAY_BERNOULLI_LOGIT_GLM_LPMF_HPP
#define STAN_MATH_CUDA_OVERLAY_BERNOULLI_LOGIT_GLM_LPMF_HPP

#define bernoulli_logit_glm_lpmf bernoulli_logit_glm_lpmf_reference_cpu
#include_next <stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp>
#undef bernoulli_logit_glm_lpmf

#ifdef STAN_CUDA
#include <stan/math/cuda/bernoulli_logit_glm_lpmf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
inline return_type_t<T_x, T_alpha, T_beta> bernoulli_logit_glm(std::true_type, const T_y& y, const T_x& x,
                                                               const T_alpha& alpha, const T_beta& beta) {
  return cuda::bernoulli_logit_glm_lpmf_device<propto>(y, x, alpha, beta);
}
template <typename T_y, typename T_x, typename T_alpha, typename T_beta>
using bernoulli_on_device = cuda::int_glm_eligible<T_y, T_x, T_alpha, T_beta>;
#else
template <typename T_y, typename T_x, typename T_alpha, typename T_beta>
using bernoulli_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
inline return_type_t<T_x, T_alpha, T_beta> bernoulli_logit_glm(std::false_type, const T_y& y, const T_x& x,
                                                               const T_alpha& alpha, const T_beta& beta) {
  return bernoulli_logit_glm_lpmf_reference_cpu<propto>(y, x, alpha, beta);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
return_type_t<T_x, T_alpha, T_beta> bernoulli_logit_glm_lpmf(const T_y& y, const T_x& x, const T_alpha& alpha,
                                                             const T_beta& beta) {
  return cuda_dispatch::bernoulli_logit_glm<propto>(
      typename cuda_dispatch::bernoulli_on_device<T_y, T_x, T_alpha, T_beta>::type(), y, x, alpha, beta);
}

template <typename T_y, typename T_x, typename T_alpha, typename T_beta>
inline return_type_t<T_x, T_beta, T_alpha> bernoulli_logit_glm_lpmf(const T_y& y, const T_x& x,
                                                                    const T_alpha& alpha, const T_beta& beta) {
  return bernoulli_logit_glm_lpmf<false>(y, x, alpha, beta);
}

#ifdef STAN_CUDA
/** Extension of the CUDA subsystem: x as a zero-copy view of column-major data (cuda::data_matrix_view). */
template <bool propto, typename T_y, typename T_alpha, typename T_beta>
inline return_type_t<T_alpha, T_beta> bernoulli_logit_glm_lpmf(const T_y& y, const cuda::data_matrix_view& x, const T_alpha& alpha,
                                         const T_beta& beta) {
  static_assert(cuda::int_glm_eligible<T_y, cuda::data_matrix_view, T_alpha, T_beta>::value,
                "a data_matrix_view is accepted on the device path only");
  return cuda::bernoulli_logit_glm_lpmf_device<propto>(y, x, alpha, beta);
}
template <typename T_y, typename T_alpha, typename T_beta>
inline return_type_t<T_alpha, T_beta> bernoulli_logit_glm_lpmf(const T_y& y, const cuda::data_matrix_view& x, const T_alpha& alpha,
                                         const T_beta& beta) {
  return cuda::bernoulli_logit_glm_lpmf_device<false>(y, x, alpha, beta);
}
#endif

}  // namespace math
}  // namespace stan
#endif
