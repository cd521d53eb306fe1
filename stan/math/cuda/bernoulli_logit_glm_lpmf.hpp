#ifndef STAN_MATH_CUDA_BERNOULLI_LOGIT_GLM_LPMF_HPP
#define STAN_MATH_CUDA_BERNOULLI_LOGIT_GLM_LPMF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/arena_partials.hpp>
#include <stan/math/cuda/glm_call.hpp>
#include <stan/math/prim/scal/err/check_bounded.hpp>
#include <stan/math/prim/scal/err/check_consistent_size.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>

namespace stan {
namespace math {
namespace cuda {

/**
 * GPU body of bernoulli_logit_glm_lpmf for data x and int y (the instantiations selected by
 * int_glm_eligible).  Same checks in the same order as the reference
 * (stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:61-78), the numeric body replaced by one
 * call into the C ABI, the lazy finiteness checks re-derived on the host only when the device
 * reports a non-finite log-density (:106-110).
 */
template <bool propto, typename T_y, typename T_x, typename T_alpha, typename T_beta>
return_type_t<typename x_operand<T_x>::scalar, T_alpha, T_beta> bernoulli_logit_glm_lpmf_device(const T_y& y, const T_x& x,
                                                                    const T_alpha& alpha, const T_beta& beta) {
  static const char* function = "bernoulli_logit_glm_lpmf";
  const size_t N = x.rows();
  const size_t M = x.cols();

  check_consistent_size(function, "Vector of dependent variables", y, N);
  check_consistent_size(function, "Weight vector", beta, M);
  if (is_vector<T_alpha>::value) {
    check_consistent_sizes(function, "Vector of intercepts", alpha, "Vector of dependent variables", y);
  }
  // check_bounded(y, 0, 1) (:70) is answered by the device from the range computed when y was
  // cached -- no O(N) host loop per evaluation; see STAN_CUDA_BAD_OUTCOME below.
  if (size_zero(y)) {
    return 0;
  }
  if (!include_summand<propto, typename x_operand<T_x>::scalar, T_alpha, T_beta>::value) {
    check_bounded(function, "Vector of dependent variables", y, 0, 1);  // nothing goes to the device
    return 0;
  }

  int_outcome<T_y> y_val(y, N);
  glm_call<T_x, T_alpha, T_beta> call(STAN_CUDA_BERNOULLI_LOGIT, x, alpha, beta);
  // x is data on this path: its (empty) edge is declared on a double so that views of x work too
  arena_partials<typename x_operand<T_x>::type, T_alpha, T_beta> ops_partials(x_operand<T_x>::get(x), alpha, beta);
  stan_cuda_glm_args& a = call.args();
  a.y = y_val.data();
  a.y_per_call = y_val.per_call() ? 1 : 0;
  a.y_is_double = 0;
  a.flags = STAN_CUDA_LP_MAIN;
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
    check_bounded(function, "Vector of dependent variables", y, 0, 1);  // throws as the reference does
    rc = call.rerun_after_stale_outcomes(function);
  }
  if (rc == STAN_CUDA_NONFINITE) {
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    Eigen::Array<double, Eigen::Dynamic, 1> ytheta = call.host_linear_predictor(x).array();
    for (size_t i = 0; i < N; ++i) {
      ytheta[i] *= 2.0 * y_val.data()[i] - 1.0;
    }
    check_finite(function, "Matrix of independent variables", ytheta);
  }
  return ops_partials.build(call.logp());
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
