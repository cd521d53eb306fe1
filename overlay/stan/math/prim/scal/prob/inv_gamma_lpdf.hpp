// Overlay of stan/math/prim/scal/prob/inv_gamma_lpdf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  Vectorised calls -- at least one
// of y, alpha, beta is a std::vector or Eigen vector of double / var -- go to the B200 kernel; scalar
// calls, fvar, calls with nothing to compute under propto and anything else stay on the reference loop.
#ifndef STAN_MATH_CUDA_OVERLAY_INV_GAMMA_LPDF_HPP
#define STAN_MATH_CUDA_OVERLAY_INV_GAMMA_LPDF_HPP

#define inv_gamma_lpdf inv_gamma_lpdf_reference_cpu
#include_next <stan/math/prim/scal/prob/inv_gamma_lpdf.hpp>
#undef inv_gamma_lpdf

#ifdef STAN_CUDA
// the device wrapper needs the Eigen/std::vector specialisations of operands_and_partials and the
// container meta helpers, which the scal-level reference header does not pull in by itself
#include <stan/math/cuda/locscale_lpdf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_loc, T_scale> inv_gamma(std::true_type, const T_y& y, const T_loc& mu,
                                                 const T_scale& sigma) {
  return cuda::inv_gamma_lpdf_device<propto>(y, mu, sigma);
}
template <bool propto, typename T_y, typename T_loc, typename T_scale>
using inv_gamma_on_device = cuda::locscale_lpdf_eligible<propto, T_y, T_loc, T_scale>;
#else
template <bool propto, typename T_y, typename T_loc, typename T_scale>
using inv_gamma_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_loc, T_scale> inv_gamma(std::false_type, const T_y& y, const T_loc& mu,
                                                 const T_scale& sigma) {
  return inv_gamma_lpdf_reference_cpu<propto>(y, mu, sigma);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_loc, T_scale> inv_gamma_lpdf(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  return cuda_dispatch::inv_gamma<propto>(typename cuda_dispatch::inv_gamma_on_device<propto, T_y, T_loc, T_scale>::type(), y, mu,
                                       sigma);
}

template <typename T_y, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_loc, T_scale> inv_gamma_lpdf(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  return inv_gamma_lpdf<false>(y, mu, sigma);
}

}  // namespace math
}  // namespace stan
#endif
