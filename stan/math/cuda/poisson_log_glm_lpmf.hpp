#ifndef STAN_MATH_CUDA_POISSON_LOG_GLM_LPMF_HPP
#define STAN_MATH_CUDA_POISSON_LOG_GLM_LPMF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/arena_partials.hpp>
#include <stan/math/cuda/glm_call.hpp>
#include <stan/math/prim/scal/err/check_consistent_size.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_nonnegative.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>

namespace stan {
namespace math {
namespace cuda {

/**
 * GPU body of poisson_log_glm_lpmf (reference: stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp).
 * Checks :66-80 unchanged; summands per include_summand exactly as :102-112 -- including the
 * reference's use of include_summand<propto, T_partials_return>, which is false whenever propto
 * is true, so the propto variant returns 0 with a non-zero gradient.
 */
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
return_type_t<typename x_operand<T_x>::scalar, T_alpha, T_beta> poisson_log_glm_lpmf_device(const T_y& y, const T_x& x, const T_alpha& alpha,
                                                                const T_beta& beta) {
  static const char* function = "poisson_log_glm_lpmf";
  using T_partials_return = partials_return_t<T_y, double, T_alpha, T_beta>;
  const size_t N = x.rows();
  const size_t M = x.cols();

  check_consistent_size(function, "Vector of dependent variables", y, N);
  check_consistent_size(function, "Weight vector", beta, M);
  if (is_vector<T_alpha>::value) {
    check_consistent_sizes(function, "Vector of intercepts", alpha, "Vector of dependent variables", y);
  }
  // check_nonnegative(y) (:70) is answered by the device from the range computed when y was
  // cached -- no O(N) host loop per evaluation; see STAN_CUDA_BAD_OUTCOME below.
  if (size_zero(y)) {
    return 0;
  }
  if (!include_summand<propto, typename x_operand<T_x>::scalar, T_alpha, T_beta>::value) {
    check_nonnegative(function, "Vector of dependent variables", y);  // nothing goes to the device
    return 0;
  }

  int_outcome<T_y> y_val(y, N);
  glm_call<T_x, T_alpha, T_beta> call(STAN_CUDA_POISSON_LOG, x, alpha, beta);
  // x is data on this path: its (empty) edge is declared on a double so that views of x work too
  arena_partials<typename x_operand<T_x>::type, T_alpha, T_beta> ops_partials(x_operand<T_x>::get(x), alpha, beta);
  stan_cuda_glm_args& a = call.args();
  a.y = y_val.data();
  a.y_per_call = y_val.per_call() ? 1 : 0;
  a.y_is_double = 0;
  a.flags = 0;
  if (include_summand<propto>::value) {
    a.flags |= STAN_CUDA_LP_CONST;
  }
  if (include_summand<propto, T_partials_return>::value) {
    a.flags |= STAN_CUDA_LP_MAIN;
  }
  if (!is_constant_all<T_beta>::value) {
    a.flags |= STAN_CUDA_NEED_DBETA;
    a.dbeta = ops_partials.partials(2);
  }
  if (!is_constant_all<T_alpha>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dalpha = ops_partials.partials(1);
  }
  if (!is_constant_all<typename x_operand<T_x>::scalar>::value) {  // var x: d * beta^T into the x edge (N x K, column-major)
    a.flags |= STAN_CUDA_NEED_DX;
    a.dx = ops_partials.partials(0);
    a.lddx = N;
  }
  int rc = call.run(function);
  if (rc == STAN_CUDA_BAD_OUTCOME) {
    check_nonnegative(function, "Vector of dependent variables", y);  // throws as the reference does
    rc = call.rerun_after_stale_outcomes(function);
  }
  if (rc == STAN_CUDA_NONFINITE) {
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    Eigen::Matrix<double, Eigen::Dynamic, 1> theta = call.host_linear_predictor(x);
    check_finite(function, "Matrix of independent variables", theta);
  }
  return ops_partials.build(call.logp());
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
