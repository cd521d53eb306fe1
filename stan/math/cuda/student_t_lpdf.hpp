#ifndef STAN_MATH_CUDA_STUDENT_T_LPDF_HPP
#define STAN_MATH_CUDA_STUDENT_T_LPDF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/cuda/arena_partials.hpp>
#include <stan/math/cuda/normal_lpdf.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/scal/err/check_positive_finite.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>
#include <stdexcept>
#include <string>

namespace stan {
namespace math {
namespace cuda {

/** vectorised calls only: at least one operand is a container, all are double / int / var */
template <typename T_y, typename T_dof, typename T_loc, typename T_scale>
struct student_t_lpdf_eligible
    : std::integral_constant<bool, is_param_arg<T_y>::value && is_param_arg<T_dof>::value
                                       && is_param_arg<T_loc>::value && is_param_arg<T_scale>::value
                                       && (is_vector<T_y>::value || is_vector<T_dof>::value
                                           || is_vector<T_loc>::value || is_vector<T_scale>::value)> {};

/** the reference's eager checks, in its order (stan/math/prim/scal/prob/student_t_lpdf.hpp:67-73) */
template <typename T_y, typename T_dof, typename T_loc, typename T_scale>
inline void student_t_checks(const char* function, const T_y& y, const T_dof& nu, const T_loc& mu,
                             const T_scale& sigma, bool sizes_too) {
  check_not_nan(function, "Random variable", y);
  check_positive_finite(function, "Degrees of freedom parameter", nu);
  check_finite(function, "Location parameter", mu);
  check_positive_finite(function, "Scale parameter", sigma);
  if (sizes_too) {
    check_consistent_sizes(function, "Random variable", y, "Degrees of freedom parameter", nu, "Location parameter",
                           mu, "Scale parameter", sigma);
  }
}

/**
 * GPU body of the vectorised student_t_lpdf (reference: stan/math/prim/scal/prob/student_t_lpdf.hpp).
 * The precomputation loops :85-149 and the accumulation loop :153-199 become one kernel that returns the
 * log-density and the partials of every non-constant operand (per element for containers, summed for scalars --
 * what the reference accumulates through broadcast_array).  The eager domain checks :67-70 are answered by the
 * same kernel (STAN_CUDA_BAD_OPERAND); the reference's check functions then run on the host only to throw the
 * identical exception.  The size check :71-73 stays in front.
 */
template <bool propto, typename T_y, typename T_dof, typename T_loc, typename T_scale>
return_type_t<T_y, T_dof, T_loc, T_scale> student_t_lpdf_device(const T_y& y, const T_dof& nu, const T_loc& mu,
                                                                const T_scale& sigma) {
  static const char* function = "student_t_lpdf";
  if (size_zero(y, nu, mu, sigma)) {
    return 0.0;
  }
  if (!include_summand<propto, T_y, T_dof, T_loc, T_scale>::value) {  // nothing goes to the device (:75-77)
    student_t_checks(function, y, nu, mu, sigma, true);
    return 0.0;
  }
  // the reference reports a domain error before a size error (:67-73): if the sizes disagree, run its checks in
  // its order and let the first one throw
  const size_t N = max_size(y, nu, mu, sigma);
  if ((is_vector<T_y>::value && length(y) != N) || (is_vector<T_dof>::value && length(nu) != N)
      || (is_vector<T_loc>::value && length(mu) != N) || (is_vector<T_scale>::value && length(sigma) != N)) {
    student_t_checks(function, y, nu, mu, sigma, true);
  }

  const size_t by = staged_operand<T_y>::bytes(y), bn = staged_operand<T_dof>::bytes(nu),
               bm = staged_operand<T_loc>::bytes(mu), bs = staged_operand<T_scale>::bytes(sigma);
  double* stage = nullptr;
  if (by + bn + bm + bs > 0) {
    stage = static_cast<double*>(stan_cuda_host_staging(by + bn + bm + bs));
    if (!stage) {
      check_cuda_rc(function, STAN_CUDA_ERR_CUDA);
    }
  }
  double* const stage_y = stage;
  double* const stage_nu = stage ? stage + by / sizeof(double) : nullptr;
  double* const stage_mu = stage ? stage + (by + bn) / sizeof(double) : nullptr;
  double* const stage_sigma = stage ? stage + (by + bn + bm) / sizeof(double) : nullptr;
  // operands_and_partials, built in place; containers of var leave their values in the staging area in the same pass
  double* const sinks[5] = {gathered_by_node<T_y>::value ? stage_y : nullptr, gathered_by_node<T_dof>::value ? stage_nu : nullptr,
                            gathered_by_node<T_loc>::value ? stage_mu : nullptr,
                            gathered_by_node<T_scale>::value ? stage_sigma : nullptr, nullptr};
  arena_partials<T_y, T_dof, T_loc, T_scale> ops_partials(y, nu, mu, sigma, double(), sinks);
  double logp = 0;
  stan_cuda_student_t_args a = stan_cuda_student_t_args();
  a.n = N;
  a.y = gathered_by_node<T_y>::value ? stage_y : staged_operand<T_y>::fill(y, stage_y);
  a.y_len = length(y);
  a.y_is_data = (is_constant_all<T_y>::value && staged_operand<T_y>::borrowed) ? 1 : 0;
  a.nu = gathered_by_node<T_dof>::value ? stage_nu : staged_operand<T_dof>::fill(nu, stage_nu);
  a.nu_len = length(nu);
  a.mu = gathered_by_node<T_loc>::value ? stage_mu : staged_operand<T_loc>::fill(mu, stage_mu);
  a.mu_len = length(mu);
  a.sigma = gathered_by_node<T_scale>::value ? stage_sigma : staged_operand<T_scale>::fill(sigma, stage_sigma);
  a.sigma_len = length(sigma);
  a.lp = &logp;
  a.flags = STAN_CUDA_LP_MAIN;                      // include_summand<propto, T_y, T_dof, T_loc, T_scale> (:164-166)
  if (include_summand<propto>::value) {
    a.flags |= STAN_CUDA_LP_CONST;                  // :155-157
  }
  if (include_summand<propto, T_dof>::value) {
    a.flags |= STAN_CUDA_LP_DOF;                    // :158-160
  }
  if (include_summand<propto, T_scale>::value) {
    a.flags |= STAN_CUDA_LP_SCALE;                  // :161-163
  }
  if (!is_constant_all<T_y>::value) {
    a.flags |= STAN_CUDA_NEED_DY;
    a.dy = ops_partials.partials(0);
  }
  if (!is_constant_all<T_dof>::value) {
    a.flags |= STAN_CUDA_NEED_DNU;
    a.dnu = ops_partials.partials(1);
  }
  if (!is_constant_all<T_loc>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dmu = ops_partials.partials(2);
  }
  if (!is_constant_all<T_scale>::value) {
    a.flags |= STAN_CUDA_NEED_DSIGMA;
    a.dsigma = ops_partials.partials(3);
  }
  const int rc = stan_cuda_student_t_lpdf_eval(&a);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    student_t_checks(function, y, nu, mu, sigma, false);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  return ops_partials.build(logp);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
