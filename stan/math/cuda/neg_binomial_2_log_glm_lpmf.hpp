#ifndef STAN_MATH_CUDA_NEG_BINOMIAL_2_LOG_GLM_LPMF_HPP
#define STAN_MATH_CUDA_NEG_BINOMIAL_2_LOG_GLM_LPMF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/arena_partials.hpp>
#include <stan/math/cuda/glm_call.hpp>
#include <stan/math/prim/scal/err/check_consistent_size.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_nonnegative.hpp>
#include <stan/math/prim/scal/err/check_positive_finite.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>

namespace stan {
namespace math {
namespace cuda {

template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
struct neg_binomial_glm_eligible
    : std::integral_constant<bool, is_int_outcome<T_y>::value && is_design_matrix<T_x>::value
                                       && is_param_arg<T_alpha>::value && is_param_arg<T_beta>::value
                                       && is_param_arg<T_precision>::value> {};

/**
 * GPU body of neg_binomial_2_log_glm_lpmf (reference:
 * stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp).  Size checks :79-88 and the O(K) eager
 * parameter checks :90-92 unchanged; check_nonnegative(y) :89 is answered by the device from the
 * range of y computed at cache fill; check_finite(theta) :120 by the finiteness witness of the
 * evaluation (the reference's own check then runs on the host to throw).  Summands per
 * include_summand exactly as :130-153.
 */
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_precision>
return_type_t<typename x_operand<T_x>::scalar, T_alpha, T_beta, T_precision> neg_binomial_2_log_glm_lpmf_device(const T_y& y, const T_x& x,
                                                                               const T_alpha& alpha,
                                                                               const T_beta& beta,
                                                                               const T_precision& phi) {
  static const char* function = "neg_binomial_2_log_glm_lpmf";
  const size_t N = x.rows();
  const size_t M = x.cols();

  check_consistent_size(function, "Vector of dependent variables", y, N);
  check_consistent_size(function, "Weight vector", beta, M);
  if (is_vector<T_precision>::value) {
    check_consistent_sizes(function, "Vector of precision parameters", phi, "Vector of dependent variables", y);
  }
  if (is_vector<T_alpha>::value) {
    check_consistent_sizes(function, "Vector of intercepts", alpha, "Vector of dependent variables", y);
  }
  if (size_zero(y, phi) || !include_summand<propto, typename x_operand<T_x>::scalar, T_alpha, T_beta, T_precision>::value) {
    check_nonnegative(function, "Failures variables", y);  // nothing goes to the device
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    check_positive_finite(function, "Precision parameter", phi);
    return 0;
  }

  int_outcome<T_y> y_val(y, N);
  values<T_precision> phi_val(phi);
  glm_call<T_x, T_alpha, T_beta> call(STAN_CUDA_NEG_BINOMIAL_2_LOG, x, alpha, beta);
  arena_partials<typename x_operand<T_x>::type, T_alpha, T_beta, T_precision> ops_partials(x_operand<T_x>::get(x), alpha, beta,
                                                                                           phi);
  stan_cuda_glm_args& a = call.args();
  a.y = y_val.data();
  a.y_per_call = y_val.per_call() ? 1 : 0;
  a.y_is_double = 0;
  a.sigma = phi_val.data();
  a.sigma_len = phi_val.size();
  a.flags = STAN_CUDA_LP_MAIN;  // include_summand<propto, x, alpha, beta, phi> holds past the early return
  if (include_summand<propto>::value) {
    a.flags |= STAN_CUDA_LP_CONST;
  }
  if (include_summand<propto, T_precision>::value) {
    a.flags |= STAN_CUDA_LP_SCALE;
  }
  if (include_summand<propto, typename x_operand<T_x>::scalar, T_alpha, T_beta>::value) {
    a.flags |= STAN_CUDA_LP_LINEAR;
  }
  if (!is_constant_all<T_beta>::value) {
    a.flags |= STAN_CUDA_NEED_DBETA;
    a.dbeta = ops_partials.partials(2);
  }
  if (!is_constant_all<T_alpha>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dalpha = ops_partials.partials(1);
  }
  if (!is_constant_all<T_precision>::value) {
    a.flags |= STAN_CUDA_NEED_DSIGMA;
    a.dsigma = ops_partials.partials(3);
  }
  if (!is_constant_all<typename x_operand<T_x>::scalar>::value) {  // var x: theta_derivative * beta^T into the x edge (:163-168)
    a.flags |= STAN_CUDA_NEED_DX;
    a.dx = ops_partials.partials(0);
    a.lddx = N;
  }
  // eager O(K) parameter checks, after the outcome check in the reference's order
  int rc = STAN_CUDA_OK;
  bool params_ok = true;
  try {
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    check_positive_finite(function, "Precision parameter", phi);
  } catch (const std::domain_error&) {
    params_ok = false;
  }
  if (params_ok) {
    rc = call.run(function);
  }
  if (!params_ok || rc == STAN_CUDA_BAD_OUTCOME) {
    check_nonnegative(function, "Failures variables", y);  // throws first, as in the reference
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    check_positive_finite(function, "Precision parameter", phi);
    rc = call.rerun_after_stale_outcomes(function);
  }
  if (rc == STAN_CUDA_NONFINITE) {
    Eigen::Array<double, Eigen::Dynamic, 1> theta = call.host_linear_predictor(x).array();
    check_finite(function, "Matrix of independent variables", theta);
  }
  return ops_partials.build(call.logp());
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
