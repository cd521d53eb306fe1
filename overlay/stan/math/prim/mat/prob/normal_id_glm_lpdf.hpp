// Overlay of stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp (see the note at the top of
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp for how the overlay works).
// Data matrix x with real outcomes (double or var) and double/var alpha, beta, sigma -> B200 path;
// anything else -> the reference body (stan/math/prim/mat.hpp:320 includes this name).
#ifndef STAN_MATH_CUDA_OVERLAY_NORMAL_ID_GLM_LPDF_HPP
#define STAN_MATH_CUDA_OVERLAY_NORMAL_ID_GLM_LPDF_HPP

#define normal_id_glm_lpdf normal_id_glm_lpdf_reference_cpu
#include_next <stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp>
#undef normal_id_glm_lpdf

#ifdef STAN_CUDA
#include <stan/math/cuda/normal_id_glm_lpdf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
inline return_type_t<T_y, T_x, T_alpha, T_beta, T_scale> normal_id_glm(std::true_type, const T_y& y, const T_x& x,
                                                                       const T_alpha& alpha, const T_beta& beta,
                                                                       const T_scale& sigma) {
  return cuda::normal_id_glm_lpdf_device<propto>(y, x, alpha, beta, sigma);
}
template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
using normal_id_on_device = cuda::normal_glm_eligible<T_y, T_x, T_alpha, T_beta, T_scale>;
#else
template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
using normal_id_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
inline return_type_t<T_y, T_x, T_alpha, T_beta, T_scale> normal_id_glm(std::false_type, const T_y& y, const T_x& x,
                                                                       const T_alpha& alpha, const T_beta& beta,
                                                                       const T_scale& sigma) {
  return normal_id_glm_lpdf_reference_cpu<propto>(y, x, alpha, beta, sigma);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
return_type_t<T_y, T_x, T_alpha, T_beta, T_scale> normal_id_glm_lpdf(const T_y& y, const T_x& x,
                                                                     const T_alpha& alpha, const T_beta& beta,
                                                                     const T_scale& sigma) {
  return cuda_dispatch::normal_id_glm<propto>(
      typename cuda_dispatch::normal_id_on_device<T_y, T_x, T_alpha, T_beta, T_scale>::type(), y, x, alpha, beta,
      sigma);
}

template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
inline return_type_t<T_y, T_x, T_alpha, T_beta, T_scale> normal_id_glm_lpdf(const T_y& y, const T_x& x,
                                                                            const T_alpha& alpha,
                                                                            const T_beta& beta,
                                                                            const T_scale& sigma) {
  return normal_id_glm_lpdf<false>(y, x, alpha, beta, sigma);
}

#ifdef STAN_CUDA
/** Extension of the CUDA subsystem: x as a zero-copy view of column-major data (cuda::data_matrix_view). */
template <bool propto, typename T_y, typename T_alpha, typename T_beta, typename T_scale>
inline return_type_t<T_y, T_alpha, T_beta, T_scale> normal_id_glm_lpdf(const T_y& y, const cuda::data_matrix_view& x,
                                                                       const T_alpha& alpha, const T_beta& beta,
                                                                       const T_scale& sigma) {
  static_assert(cuda::normal_glm_eligible<T_y, cuda::data_matrix_view, T_alpha, T_beta, T_scale>::value,
                "a data_matrix_view is accepted on the device path only");
  return cuda::normal_id_glm_lpdf_device<propto>(y, x, alpha, beta, sigma);
}
template <typename T_y, typename T_alpha, typename T_beta, typename T_scale>
inline return_type_t<T_y, T_alpha, T_beta, T_scale> normal_id_glm_lpdf(const T_y& y, const cuda::data_matrix_view& x,
                                                                       const T_alpha& alpha, const T_beta& beta,
                                                                       const T_scale& sigma) {
  return cuda::normal_id_glm_lpdf_device<false>(y, x, alpha, beta, sigma);
}
#endif

}  // namespace math
}  // namespace stan
#endif
