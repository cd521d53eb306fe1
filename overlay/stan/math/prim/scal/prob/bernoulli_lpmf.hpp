// Overlay of stan/math/prim/scal/prob/bernoulli_lpmf.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  Calls whose real operand is a std::vector or
// Eigen vector of double / var (and whose outcome is int / std::vector<int> / Eigen int vector) go to the B200
// kernel (stan/math/cuda/int_density.hpp); scalar operands, fvar and anything else stay on the reference loop.
#ifndef STAN_MATH_CUDA_OVERLAY_BERNOULLI_LPMF_HPP
#define STAN_MATH_CUDA_OVERLAY_BERNOULLI_LPMF_HPP

#define bernoulli_lpmf bernoulli_lpmf_reference_cpu
#include_next <stan/math/prim/scal/prob/bernoulli_lpmf.hpp>
#undef bernoulli_lpmf

#ifdef STAN_CUDA
#include <stan/math/cuda/int_density.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {
#ifdef STAN_CUDA
template <bool propto, typename T_n, typename T_op>
inline return_type_t<T_op> bernoulli_prob(std::true_type, const T_n& n, const T_op& theta) {
  return cuda::bernoulli_lpmf_device<propto>(n, theta);
}
template <typename T_n, typename T_op>
using bernoulli_prob_on_device = cuda::int_density_eligible<T_n, T_op>;
#else
template <typename T_n, typename T_op>
using bernoulli_prob_on_device = std::false_type;
#endif
template <bool propto, typename T_n, typename T_op>
inline return_type_t<T_op> bernoulli_prob(std::false_type, const T_n& n, const T_op& theta) {
  return bernoulli_lpmf_reference_cpu<propto>(n, theta);
}
}  // namespace cuda_dispatch

template <bool propto, typename T_n, typename T_op>
inline return_type_t<T_op> bernoulli_lpmf(const T_n& n, const T_op& theta) {
  return cuda_dispatch::bernoulli_prob<propto>(typename cuda_dispatch::bernoulli_prob_on_device<T_n, T_op>::type(), n, theta);
}

template <typename T_n, typename T_op>
inline return_type_t<T_op> bernoulli_lpmf(const T_n& n, const T_op& theta) {
  return bernoulli_lpmf<false>(n, theta);
}

}  // namespace math
}  // namespace stan
#endif
