// Overlay of stan/math/prim/scal/prob/exponential_lpdf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  Vectorised calls -- y or beta is a std::vector or
// Eigen vector of double / var -- go to the B200 kernel; scalar calls, fvar, calls with nothing to compute under propto
// and anything else stay on the reference loop.
#ifndef STAN_MATH_CUDA_OVERLAY_EXPONENTIAL_LPDF_HPP
#define STAN_MATH_CUDA_OVERLAY_EXPONENTIAL_LPDF_HPP

#define exponential_lpdf exponential_lpdf_reference_cpu
#include_next <stan/math/prim/scal/prob/exponential_lpdf.hpp>
#undef exponential_lpdf

#ifdef STAN_CUDA
#include <stan/math/cuda/locscale_lpdf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_inv_scale>
inline return_type_t<T_y, T_inv_scale> exponential(std::true_type, const T_y& y, const T_inv_scale& beta) {
  return cuda::exponential_lpdf_device<propto>(y, beta);
}
template <bool propto, typename T_y, typename T_inv_scale>
using exponential_on_device = cuda::locscale_lpdf_eligible<propto, T_y, double, T_inv_scale>;
#else
template <bool propto, typename T_y, typename T_inv_scale>
using exponential_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_inv_scale>
inline return_type_t<T_y, T_inv_scale> exponential(std::false_type, const T_y& y, const T_inv_scale& beta) {
  return exponential_lpdf_reference_cpu<propto>(y, beta);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_inv_scale>
inline return_type_t<T_y, T_inv_scale> exponential_lpdf(const T_y& y, const T_inv_scale& beta) {
  return cuda_dispatch::exponential<propto>(typename cuda_dispatch::exponential_on_device<propto, T_y, T_inv_scale>::type(), y,
                                            beta);
}

template <typename T_y, typename T_inv_scale>
inline return_type_t<T_y, T_inv_scale> exponential_lpdf(const T_y& y, const T_inv_scale& beta) {
  return exponential_lpdf<false>(y, beta);
}

}  // namespace math
}  // namespace stan
#endif
