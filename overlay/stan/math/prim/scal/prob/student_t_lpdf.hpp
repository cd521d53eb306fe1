// Overlay of stan/math/prim/scal/prob/student_t_lpdf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  Vectorised calls -- at least one
// of y, nu, mu, sigma is a std::vector or Eigen vector of double / var -- go to the B200 kernel; scalar
// calls, fvar and anything else stay on the reference loops.
#ifndef STAN_MATH_CUDA_OVERLAY_STUDENT_T_LPDF_HPP
#define STAN_MATH_CUDA_OVERLAY_STUDENT_T_LPDF_HPP

#define student_t_lpdf student_t_lpdf_reference_cpu
#include_next <stan/math/prim/scal/prob/student_t_lpdf.hpp>
#undef student_t_lpdf

#ifdef STAN_CUDA
#include <stan/math/cuda/student_t_lpdf.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_y, typename T_dof, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_dof, T_loc, T_scale> student_t(std::true_type, const T_y& y, const T_dof& nu,
                                                           const T_loc& mu, const T_scale& sigma) {
  return cuda::student_t_lpdf_device<propto>(y, nu, mu, sigma);
}
template <typename T_y, typename T_dof, typename T_loc, typename T_scale>
using student_t_on_device = cuda::student_t_lpdf_eligible<T_y, T_dof, T_loc, T_scale>;
#else
template <typename T_y, typename T_dof, typename T_loc, typename T_scale>
using student_t_on_device = std::false_type;
#endif
template <bool propto, typename T_y, typename T_dof, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_dof, T_loc, T_scale> student_t(std::false_type, const T_y& y, const T_dof& nu,
                                                           const T_loc& mu, const T_scale& sigma) {
  return student_t_lpdf_reference_cpu<propto>(y, nu, mu, sigma);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_y, typename T_dof, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_dof, T_loc, T_scale> student_t_lpdf(const T_y& y, const T_dof& nu, const T_loc& mu,
                                                                const T_scale& sigma) {
  return cuda_dispatch::student_t<propto>(typename cuda_dispatch::student_t_on_device<T_y, T_dof, T_loc, T_scale>::type(),
                                          y, nu, mu, sigma);
}

template <typename T_y, typename T_dof, typename T_loc, typename T_scale>
inline return_type_t<T_y, T_dof, T_loc, T_scale> student_t_lpdf(const T_y& y, const T_dof& nu, const T_loc& mu,
                                                                const T_scale& sigma) {
  return student_t_lpdf<false>(y, nu, mu, sigma);
}

}  // namespace math
}  // namespace stan
#endif
