#ifndef STAN_MATH_CUDA_INT_DENSITY_HPP
#define STAN_MATH_CUDA_INT_DENSITY_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/cuda/normal_lpdf.hpp>
#include <stan/math/prim/scal/err/check_bounded.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_nonnegative.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/scal/fun/constants.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>
#include <stdexcept>
#include <string>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/**
 * Vectorised lpmf's with an integer outcome and one real operand.  GPU path when the
 * operand is a container of double / var (the work is O(N) special functions and an N-operand tape
 * node) and the outcome is an int, std::vector<int> or Eigen int vector; scalar operands, fvar and
 * anything else stay on the reference loop.
 */
template <typename T_n, typename T_prob>
struct int_density_eligible
    : std::integral_constant<bool, is_int_outcome<T_n>::value && is_param_arg<T_prob>::value
                                       && is_vector<T_prob>::value> {};

inline const int* int_outcome_ptr(const int& n) { return &n; }
inline const int* int_outcome_ptr(const std::vector<int>& n) { return n.data(); }
template <int R, int C, int O, int MR, int MC>
inline const int* int_outcome_ptr(const Eigen::Matrix<int, R, C, O, MR, MC>& n) {
  return n.data();
}

/** Fills the argument block shared by the two densities and calls the C ABI. */
/** Page-locked staging for the operand's values (nullptr: the operand is borrowed, or the allocation failed -- the
 * call then reports it). */
template <typename T_prob>
inline double* int_density_stage(const T_prob& theta) {
  const size_t bt = staged_operand<T_prob>::bytes(theta);
  return bt > 0 ? static_cast<double*>(stan_cuda_host_staging(bt)) : nullptr;
}

template <typename T_n, typename T_prob>
inline int int_density_call(int density, bool lp_const, const T_n& n, const T_prob& theta, size_t N,
                            arena_partials<T_prob>& ops_partials, double* stage, double* logp) {
  if (staged_operand<T_prob>::bytes(theta) > 0 && !stage) {
    return STAN_CUDA_ERR_CUDA;
  }
  stan_cuda_density_args a = stan_cuda_density_args();
  a.density = density;
  a.n = N;
  a.y = int_outcome_ptr(n);
  a.y_len = length(n);
  // (a container of var left its values in `stage` while the tape node was assembled: one pass over the operands)
  a.theta = gathered_by_node<T_prob>::value ? stage : staged_operand<T_prob>::fill(theta, stage);
  a.theta_len = length(theta);
  a.lp = logp;
  a.flags = STAN_CUDA_LP_MAIN;
  if (lp_const) {
    a.flags |= STAN_CUDA_LP_CONST;
  }
  if (!is_constant_all<T_prob>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dtheta = ops_partials.partials(0);
  }
  return stan_cuda_density_eval(&a);
}

/**
 * GPU body of bernoulli_logit_lpmf (reference: stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:28-82).
 * The loop :54-80 becomes one kernel; the eager checks :42-43 are answered by the same kernel
 * (STAN_CUDA_BAD_OPERAND), after which the reference's check functions run on the host only to throw the
 * identical exception.
 */
template <bool propto, typename T_n, typename T_prob>
return_type_t<T_prob> bernoulli_logit_lpmf_device(const T_n& n, const T_prob& theta) {
  static const char* function = "bernoulli_logit_lpmf";
  if (size_zero(n, theta)) {
    return 0.0;
  }
  const size_t N = max_size(n, theta);
  const bool sizes_ok = (!is_vector<T_n>::value || length(n) == N) && length(theta) == N;
  if (!sizes_ok || !include_summand<propto, T_prob>::value) {  // nothing goes to the device
    check_bounded(function, "n", n, 0, 1);
    check_not_nan(function, "Logit transformed probability parameter", theta);
    check_consistent_sizes(function, "Random variable", n, "Probability parameter", theta);
    return 0.0;
  }
  double* const stage = int_density_stage(theta);
  double* const sinks[5] = {gathered_by_node<T_prob>::value ? stage : nullptr, nullptr, nullptr, nullptr, nullptr};
  arena_partials<T_prob> ops_partials(theta, double(), double(), double(), double(), sinks);
  double logp = 0;
  const int rc = int_density_call(STAN_CUDA_DENSITY_BERNOULLI_LOGIT, false, n, theta, N, ops_partials, stage, &logp);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    check_bounded(function, "n", n, 0, 1);
    check_not_nan(function, "Logit transformed probability parameter", theta);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  return ops_partials.build(logp);
}

/**
 * GPU body of poisson_log_lpmf (reference: stan/math/prim/scal/prob/poisson_log_lpmf.hpp:22-97), including
 * its LOG_ZERO early returns (:51-62, STAN_CUDA_LOG_ZERO) and the skipped term of a -inf log rate next to a
 * zero count (:77).
 */
template <bool propto, typename T_n, typename T_log_rate>
return_type_t<T_log_rate> poisson_log_lpmf_device(const T_n& n, const T_log_rate& alpha) {
  static const char* function = "poisson_log_lpmf";
  if (size_zero(n, alpha)) {
    return 0.0;
  }
  const size_t N = max_size(n, alpha);
  const bool sizes_ok = (!is_vector<T_n>::value || length(n) == N) && length(alpha) == N;
  if (!sizes_ok || !include_summand<propto, T_log_rate>::value) {
    check_nonnegative(function, "Random variable", n);
    check_not_nan(function, "Log rate parameter", alpha);
    check_consistent_sizes(function, "Random variable", n, "Log rate parameter", alpha);
    return 0.0;
  }
  double* const stage = int_density_stage(alpha);
  double* const sinks[5] = {gathered_by_node<T_log_rate>::value ? stage : nullptr, nullptr, nullptr, nullptr, nullptr};
  arena_partials<T_log_rate> ops_partials(alpha, double(), double(), double(), double(), sinks);
  double logp = 0;
  const int rc = int_density_call(STAN_CUDA_DENSITY_POISSON_LOG, include_summand<propto>::value, n, alpha, N,
                                  ops_partials, stage, &logp);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    check_nonnegative(function, "Random variable", n);
    check_not_nan(function, "Log rate parameter", alpha);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  if (rc == STAN_CUDA_LOG_ZERO) {
    return LOG_ZERO;
  }
  return ops_partials.build(logp);
}

/**
 * GPU body of poisson_lpmf (reference: stan/math/prim/scal/prob/poisson_lpmf.hpp:20-77): the loop :57-73 becomes one
 * kernel, which also answers the eager checks :31-33 (STAN_CUDA_BAD_OPERAND) and the two LOG_ZERO scans :45-54
 * (STAN_CUDA_LOG_ZERO).  A zero rate next to a zero count contributes no term (:59) and the partial 0/0 - 1, as there.
 */
template <bool propto, typename T_n, typename T_rate>
return_type_t<T_rate> poisson_lpmf_device(const T_n& n, const T_rate& lambda) {
  static const char* function = "poisson_lpmf";
  if (size_zero(n, lambda)) {
    return 0.0;
  }
  const size_t N = max_size(n, lambda);
  const bool sizes_ok = (!is_vector<T_n>::value || length(n) == N) && length(lambda) == N;
  if (!sizes_ok || !include_summand<propto, T_rate>::value) {
    check_nonnegative(function, "Random variable", n);
    check_not_nan(function, "Rate parameter", lambda);
    check_nonnegative(function, "Rate parameter", lambda);
    check_consistent_sizes(function, "Random variable", n, "Rate parameter", lambda);
    return 0.0;
  }
  double* const stage = int_density_stage(lambda);
  double* const sinks[5] = {gathered_by_node<T_rate>::value ? stage : nullptr, nullptr, nullptr, nullptr, nullptr};
  arena_partials<T_rate> ops_partials(lambda, double(), double(), double(), double(), sinks);
  double logp = 0;
  const int rc = int_density_call(STAN_CUDA_DENSITY_POISSON, include_summand<propto>::value, n, lambda, N, ops_partials,
                                  stage, &logp);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    check_nonnegative(function, "Random variable", n);
    check_not_nan(function, "Rate parameter", lambda);
    check_nonnegative(function, "Rate parameter", lambda);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  if (rc == STAN_CUDA_LOG_ZERO) {
    return LOG_ZERO;
  }
  return ops_partials.build(logp);
}

/**
 * GPU body of bernoulli_lpmf (reference: stan/math/prim/scal/prob/bernoulli_lpmf.hpp:28-112): the per-element loop
 * :92-107 becomes one kernel, which also answers the eager checks :41-43; a probability container of length one takes
 * the reference's closed form from the number of successes (:58-91) inside the library.
 */
template <bool propto, typename T_n, typename T_prob>
return_type_t<T_prob> bernoulli_lpmf_device(const T_n& n, const T_prob& theta) {
  static const char* function = "bernoulli_lpmf";
  if (size_zero(n, theta)) {
    return 0.0;
  }
  const size_t N = max_size(n, theta);
  const bool sizes_ok = (!is_vector<T_n>::value || length(n) == N) && length(theta) == N;
  if (!sizes_ok || !include_summand<propto, T_prob>::value) {
    check_bounded(function, "n", n, 0, 1);
    check_finite(function, "Probability parameter", theta);
    check_bounded(function, "Probability parameter", theta, 0.0, 1.0);
    check_consistent_sizes(function, "Random variable", n, "Probability parameter", theta);
    return 0.0;
  }
  double* const stage = int_density_stage(theta);
  double* const sinks[5] = {gathered_by_node<T_prob>::value ? stage : nullptr, nullptr, nullptr, nullptr, nullptr};
  arena_partials<T_prob> ops_partials(theta, double(), double(), double(), double(), sinks);
  double logp = 0;
  const int rc = int_density_call(STAN_CUDA_DENSITY_BERNOULLI, false, n, theta, N, ops_partials, stage, &logp);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    check_bounded(function, "n", n, 0, 1);
    check_finite(function, "Probability parameter", theta);
    check_bounded(function, "Probability parameter", theta, 0.0, 1.0);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  return ops_partials.build(logp);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
