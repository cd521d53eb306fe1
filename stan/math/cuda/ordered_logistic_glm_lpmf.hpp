#ifndef STAN_MATH_CUDA_ORDERED_LOGISTIC_GLM_LPMF_HPP
#define STAN_MATH_CUDA_ORDERED_LOGISTIC_GLM_LPMF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/glm_call.hpp>
#include <stan/math/prim/scal/err/check_consistent_size.hpp>
#include <stan/math/prim/scal/err/check_bounded.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/mat/err/check_ordered.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/** std::vector<int> classes, dynamic data matrix x, double / var weights and cut-points. */
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_beta_scalar, typename T_cuts_scalar>
struct ordered_logistic_glm_eligible
    : std::integral_constant<bool, std::is_same<T_y, std::vector<int>>::value
                                       && std::is_same<T_x_scalar, double>::value && T_x_rows == Eigen::Dynamic
                                       && is_double_or_var<T_beta_scalar>::value
                                       && is_double_or_var<T_cuts_scalar>::value> {};

/**
 * GPU body of ordered_logistic_glm_lpmf (reference:
 * stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:46-199).  Size checks :71-77 and the O(C)
 * checks of the cut-points :79-85 unchanged and in the reference's order; check_bounded(y, 1,
 * N_classes) :78 is answered by the device from the range of y computed at cache fill; the lazy
 * finiteness checks :119-122 run on the host only when the device reports a non-finite sum of
 * locations.  Any number of cut-points (beyond 1024 the device runs its general path, glm_ordered_wide.cu).
 */
template <bool propto, typename T_beta_scalar, typename T_cuts_scalar>
return_type_t<T_beta_scalar, T_cuts_scalar> ordered_logistic_glm_lpmf_device(
    const std::vector<int>& y, const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, 1>& beta,
    const Eigen::Matrix<T_cuts_scalar, Eigen::Dynamic, 1>& cuts) {
  typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> T_x;
  typedef Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, 1> T_beta;
  typedef Eigen::Matrix<T_cuts_scalar, Eigen::Dynamic, 1> T_cuts;
  static const char* function = "ordered_logistic_glm_lpmf";
  const size_t N_instances = x.rows();
  const size_t N_attributes = x.cols();
  const size_t N_classes = length(cuts) + 1;

  check_consistent_size(function, "Vector of dependent variables", y, N_instances);
  check_consistent_size(function, "Weight vector", beta, N_attributes);
  // everything the reference checks after the outcomes, in its order (:79-85)
  auto check_cuts = [&]() {
    check_ordered(function, "Cut-points", cuts);
    if (N_classes > 1) {
      if (N_classes > 2) {
        check_finite(function, "Final cut-point", cuts[N_classes - 2]);
      }
      check_finite(function, "First cut-point", cuts[0]);
    }
  };
  const bool nothing_to_do
      = size_zero(y, cuts) || !include_summand<propto, double, T_beta_scalar, T_cuts_scalar>::value;
  bool cuts_ok = true;
  try {
    check_cuts();
  } catch (const std::domain_error&) {
    cuts_ok = false;
  }
  if (nothing_to_do || !cuts_ok) {  // nothing goes to the device: the reference's checks, then 0 (:88-92)
    check_bounded(function, "Vector of dependent variables", y, 1, N_classes);
    check_cuts();
    return 0;
  }

  int_outcome<std::vector<int>> y_val(y, N_instances);
  values<T_cuts> cuts_val(cuts);
  const double no_intercept = 0;
  glm_call<T_x, double, T_beta> call(STAN_CUDA_ORDERED_LOGISTIC, x, no_intercept, beta);
  operands_and_partials<T_x, T_beta, T_cuts> ops_partials(x, beta, cuts);
  stan_cuda_glm_args& a = call.args();
  a.y = y_val.data();
  a.y_is_double = 0;
  a.cuts = cuts_val.data();
  a.cuts_len = cuts_val.size();
  a.flags = STAN_CUDA_LP_MAIN;
  if (!is_constant_all<T_beta_scalar>::value) {
    a.flags |= STAN_CUDA_NEED_DBETA;
    a.dbeta = partials_data(ops_partials.edge2_);
  }
  if (!is_constant_all<T_cuts_scalar>::value) {
    a.flags |= STAN_CUDA_NEED_DCUTS;
    a.dcuts = partials_data(ops_partials.edge3_);
  }
  int rc = call.run(function);
  if (rc == STAN_CUDA_BAD_OUTCOME) {
    check_bounded(function, "Vector of dependent variables", y, 1, N_classes);  // throws, as in the reference
    rc = call.rerun_after_stale_outcomes(function);
  }
  if (rc == STAN_CUDA_NONFINITE) {  // :119-122
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Matrix of independent variables", x);
  }
  return ops_partials.build(call.logp());
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
