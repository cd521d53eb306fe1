#ifndef STAN_MATH_CUDA_NORMAL_LPDF_HPP
#define STAN_MATH_CUDA_NORMAL_LPDF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/scal/err/check_positive.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>

namespace stan {
namespace math {
namespace cuda {

/** vectorised calls only: at least one operand is a container, all are double / int / var */
template <typename T_y, typename T_loc, typename T_scale>
struct normal_lpdf_eligible
    : std::integral_constant<bool, is_param_arg<T_y>::value && is_param_arg<T_loc>::value
                                       && is_param_arg<T_scale>::value
                                       && (is_vector<T_y>::value || is_vector<T_loc>::value
                                           || is_vector<T_scale>::value)> {};

/**
 * GPU body of the vectorised normal_lpdf (reference: stan/math/prim/scal/prob/normal_lpdf.hpp).
 * Checks :45-58 unchanged and eager; the loop :78-110 becomes one kernel that returns the
 * log-density and the partials of every non-constant operand (per element for containers, summed
 * for scalars -- what the reference accumulates through broadcast_array).
 */
template <bool propto, typename T_y, typename T_loc, typename T_scale>
return_type_t<T_y, T_loc, T_scale> normal_lpdf_device(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  static const char* function = "normal_lpdf";
  if (size_zero(y, mu, sigma)) {
    return 0.0;
  }
  check_not_nan(function, "Random variable", y);
  check_finite(function, "Location parameter", mu);
  check_positive(function, "Scale parameter", sigma);
  check_consistent_sizes(function, "Random variable", y, "Location parameter", mu, "Scale parameter", sigma);
  if (!include_summand<propto, T_y, T_loc, T_scale>::value) {
    return 0.0;
  }

  operands_and_partials<T_y, T_loc, T_scale> ops_partials(y, mu, sigma);
  values<T_y> y_val(y);
  values<T_loc> mu_val(mu);
  values<T_scale> sigma_val(sigma);
  double logp = 0;
  stan_cuda_normal_args a = stan_cuda_normal_args();
  a.n = max_size(y, mu, sigma);
  a.y = y_val.data();
  a.y_len = y_val.size();
  a.y_is_data = (is_constant_all<T_y>::value && y_val.borrowed()) ? 1 : 0;
  a.mu = mu_val.data();
  a.mu_len = mu_val.size();
  a.sigma = sigma_val.data();
  a.sigma_len = sigma_val.size();
  a.lp = &logp;
  a.flags = STAN_CUDA_LP_MAIN;
  if (include_summand<propto>::value) {
    a.flags |= STAN_CUDA_LP_CONST;
  }
  if (include_summand<propto, T_scale>::value) {
    a.flags |= STAN_CUDA_LP_SCALE;
  }
  if (!is_constant_all<T_y>::value) {
    a.flags |= STAN_CUDA_NEED_DY;
    a.dy = partials_data(ops_partials.edge1_);
  }
  if (!is_constant_all<T_loc>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dmu = partials_data(ops_partials.edge2_);
  }
  if (!is_constant_all<T_scale>::value) {
    a.flags |= STAN_CUDA_NEED_DSIGMA;
    a.dsigma = partials_data(ops_partials.edge3_);
  }
  check_cuda_rc(function, stan_cuda_normal_lpdf_eval(&a));
  return ops_partials.build(logp);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
