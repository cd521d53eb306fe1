#ifndef STAN_MATH_CUDA_LOCSCALE_LPDF_HPP
#define STAN_MATH_CUDA_LOCSCALE_LPDF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/normal_lpdf.hpp>
#include <stan/math/prim/scal/err/check_less_or_equal.hpp>
#include <stan/math/prim/scal/err/check_nonnegative.hpp>
#include <stan/math/prim/scal/err/check_positive_finite.hpp>
#include <stan/math/prim/scal/fun/constants.hpp>
#include <stdexcept>
#include <string>

namespace stan {
namespace math {
namespace cuda {

/**
 * cauchy_lpdf, lognormal_lpdf, gamma_lpdf, exponential_lpdf, beta_lpdf, weibull_lpdf and inv_gamma_lpdf on the device: the operands, their staging and the tape node are
 * those of normal_lpdf_device (stan/math/cuda/normal_lpdf.hpp); the density id selects the per-element term of the
 * kernel (stan_cuda_locscale_lpdf_eval).  Eligible calls: the vectorised ones with something to compute -- when
 * include_summand<propto, T_y, T_loc, T_scale> is false the reference body only runs its checks, and stays in charge.
 */
template <bool propto, typename T_y, typename T_loc, typename T_scale>
struct locscale_lpdf_eligible
    : std::integral_constant<bool, normal_lpdf_eligible<T_y, T_loc, T_scale>::value
                                       && include_summand<propto, T_y, T_loc, T_scale>::value> {};

namespace detail {

/** the eager checks of cauchy_lpdf.hpp:47-49, in the reference's order */
struct cauchy_checks {
  static const char* function() { return "cauchy_lpdf"; }
  template <typename T_y, typename T_loc, typename T_scale>
  static void run(const T_y& y, const T_loc& mu, const T_scale& sigma) {
    check_not_nan(function(), "Random variable", y);
    check_finite(function(), "Location parameter", mu);
    check_positive_finite(function(), "Scale parameter", sigma);
  }
  template <typename T_y, typename T_loc, typename T_scale>
  static void sizes(const T_y& y, const T_loc& mu, const T_scale& sigma) {
    check_consistent_sizes(function(), "Random variable", y, "Location parameter", mu, "Scale parameter", sigma);
  }
};
/** the eager checks of lognormal_lpdf.hpp:24-27 */
struct lognormal_checks {
  static const char* function() { return "lognormal_lpdf"; }
  template <typename T_y, typename T_loc, typename T_scale>
  static void run(const T_y& y, const T_loc& mu, const T_scale& sigma) {
    check_not_nan(function(), "Random variable", y);
    check_nonnegative(function(), "Random variable", y);
    check_finite(function(), "Location parameter", mu);
    check_positive_finite(function(), "Scale parameter", sigma);
  }
  template <typename T_y, typename T_loc, typename T_scale>
  static void sizes(const T_y& y, const T_loc& mu, const T_scale& sigma) {
    check_consistent_sizes(function(), "Random variable", y, "Location parameter", mu, "Scale parameter", sigma);
  }
};

/** the eager checks of gamma_lpdf.hpp:55-57 (alpha in the location slot, beta in the scale slot) */
struct gamma_checks {
  static const char* function() { return "gamma_lpdf"; }
  template <typename T_y, typename T_shape, typename T_inv_scale>
  static void run(const T_y& y, const T_shape& alpha, const T_inv_scale& beta) {
    check_not_nan(function(), "Random variable", y);
    check_positive_finite(function(), "Shape parameter", alpha);
    check_positive_finite(function(), "Inverse scale parameter", beta);
  }
  template <typename T_y, typename T_shape, typename T_inv_scale>
  static void sizes(const T_y& y, const T_shape& alpha, const T_inv_scale& beta) {
    check_consistent_sizes(function(), "Random variable", y, "Shape parameter", alpha, "Inverse scale parameter", beta);
  }
};
/** the eager checks of exponential_lpdf.hpp:55-56 (beta in the scale slot; the location slot holds one zero) */
struct exponential_checks {
  static const char* function() { return "exponential_lpdf"; }
  template <typename T_y, typename T_unused, typename T_inv_scale>
  static void run(const T_y& y, const T_unused&, const T_inv_scale& beta) {
    check_nonnegative(function(), "Random variable", y);
    check_positive_finite(function(), "Inverse scale parameter", beta);
  }
  template <typename T_y, typename T_unused, typename T_inv_scale>
  static void sizes(const T_y& y, const T_unused&, const T_inv_scale& beta) {
    check_consistent_sizes(function(), "Random variable", y, "Inverse scale parameter", beta);
  }
};

/** the eager checks of beta_lpdf.hpp:48-54: the shapes and NaN outcomes, then (behind the size check) the range of y */
struct beta_checks {
  static const char* function() { return "beta_lpdf"; }
  template <typename T_y, typename T_a, typename T_b>
  static void run(const T_y& y, const T_a& alpha, const T_b& beta) {
    check_positive_finite(function(), "First shape parameter", alpha);
    check_positive_finite(function(), "Second shape parameter", beta);
    check_not_nan(function(), "Random variable", y);
  }
  template <typename T_y, typename T_a, typename T_b>
  static void sizes(const T_y& y, const T_a& alpha, const T_b& beta) {
    check_consistent_sizes(function(), "Random variable", y, "First shape parameter", alpha, "Second shape parameter", beta);
    check_nonnegative(function(), "Random variable", y);
    check_less_or_equal(function(), "Random variable", y, 1);
  }
};
/** the eager checks of weibull_lpdf.hpp:38-40 */
struct weibull_checks {
  static const char* function() { return "weibull_lpdf"; }
  template <typename T_y, typename T_a, typename T_s>
  static void run(const T_y& y, const T_a& alpha, const T_s& sigma) {
    check_finite(function(), "Random variable", y);
    check_positive_finite(function(), "Shape parameter", alpha);
    check_positive_finite(function(), "Scale parameter", sigma);
  }
  template <typename T_y, typename T_a, typename T_s>
  static void sizes(const T_y& y, const T_a& alpha, const T_s& sigma) {
    check_consistent_sizes(function(), "Random variable", y, "Shape parameter", alpha, "Scale parameter", sigma);
  }
};

/** the eager checks of inv_gamma_lpdf.hpp:41-43 */
struct inv_gamma_checks {
  static const char* function() { return "inv_gamma_lpdf"; }
  template <typename T_y, typename T_a, typename T_b>
  static void run(const T_y& y, const T_a& alpha, const T_b& beta) {
    check_not_nan(function(), "Random variable", y);
    check_positive_finite(function(), "Shape parameter", alpha);
    check_positive_finite(function(), "Scale parameter", beta);
  }
  template <typename T_y, typename T_a, typename T_b>
  static void sizes(const T_y& y, const T_a& alpha, const T_b& beta) {
    check_consistent_sizes(function(), "Random variable", y, "Shape parameter", alpha, "Scale parameter", beta);
  }
};

/**
 * One vectorised evaluation: stages the operands, builds the tape node in place (arena_partials), calls the device.
 * summands: the STAN_CUDA_LP_* bits of the terms the reference keeps for these operand types under this propto.
 */
template <typename Checks, typename T_y, typename T_loc, typename T_scale>
return_type_t<T_y, T_loc, T_scale> locscale_lpdf_call(int density, uint32_t summands, const T_y& y, const T_loc& mu,
                                                      const T_scale& sigma) {
  const char* function = Checks::function();
  // the reference reports a domain error before a size error: if the sizes disagree, its checks run in its order
  const size_t N = max_size(y, mu, sigma);
  if ((is_vector<T_y>::value && length(y) != N) || (is_vector<T_loc>::value && length(mu) != N)
      || (is_vector<T_scale>::value && length(sigma) != N)) {
    Checks::run(y, mu, sigma);
    Checks::sizes(y, mu, sigma);
  }
  const size_t by = staged_operand<T_y>::bytes(y), bm = staged_operand<T_loc>::bytes(mu),
               bs = staged_operand<T_scale>::bytes(sigma);
  double* stage = nullptr;
  if (by + bm + bs > 0) {
    stage = static_cast<double*>(stan_cuda_host_staging(by + bm + bs));
    if (!stage) {
      check_cuda_rc(function, STAN_CUDA_ERR_CUDA);
    }
  }
  double* const stage_y = stage;
  double* const stage_mu = stage ? stage + by / sizeof(double) : nullptr;
  double* const stage_sigma = stage ? stage + (by + bm) / sizeof(double) : nullptr;
  double* const sinks[5] = {gathered_by_node<T_y>::value ? stage_y : nullptr, gathered_by_node<T_loc>::value ? stage_mu : nullptr,
                            gathered_by_node<T_scale>::value ? stage_sigma : nullptr, nullptr, nullptr};
  arena_partials<T_y, T_loc, T_scale> ops_partials(y, mu, sigma, double(), double(), sinks);
  double logp = 0;
  stan_cuda_normal_args a = stan_cuda_normal_args();
  a.n = N;
  a.y = gathered_by_node<T_y>::value ? stage_y : staged_operand<T_y>::fill(y, stage_y);
  a.y_len = length(y);
  a.y_is_data = (is_constant_all<T_y>::value && staged_operand<T_y>::borrowed) ? 1 : 0;
  a.mu = gathered_by_node<T_loc>::value ? stage_mu : staged_operand<T_loc>::fill(mu, stage_mu);
  a.mu_len = length(mu);
  a.sigma = gathered_by_node<T_scale>::value ? stage_sigma : staged_operand<T_scale>::fill(sigma, stage_sigma);
  a.sigma_len = length(sigma);
  a.lp = &logp;
  a.flags = summands;
  if (!is_constant_all<T_y>::value) {
    a.flags |= STAN_CUDA_NEED_DY;
    a.dy = ops_partials.partials(0);
  }
  if (!is_constant_all<T_loc>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dmu = ops_partials.partials(1);
  }
  if (!is_constant_all<T_scale>::value) {
    a.flags |= STAN_CUDA_NEED_DSIGMA;
    a.dsigma = ops_partials.partials(2);
  }
  const int rc = stan_cuda_locscale_lpdf_eval(density, &a);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    Checks::run(y, mu, sigma);
    Checks::sizes(y, mu, sigma);  // (the sizes agree here; beta_lpdf tests the range of y behind its size check)
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  if (rc == STAN_CUDA_LOG_ZERO) {
    return LOG_ZERO;
  }
  return ops_partials.build(logp);
}

}  // namespace detail

/**
 * GPU body of the vectorised cauchy_lpdf (reference: stan/math/prim/scal/prob/cauchy_lpdf.hpp:35-113): the loop :77-110
 * becomes one kernel; the eager checks :47-49 are answered by the same kernel (STAN_CUDA_BAD_OPERAND), after which the
 * reference's check functions run on the host only to throw the identical exception.
 */
template <bool propto, typename T_y, typename T_loc, typename T_scale>
return_type_t<T_y, T_loc, T_scale> cauchy_lpdf_device(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  if (size_zero(y, mu, sigma)) {
    return 0.0;
  }
  const uint32_t summands = STAN_CUDA_LP_MAIN | (include_summand<propto>::value ? STAN_CUDA_LP_CONST : 0u)   // :87-95
                            | (include_summand<propto, T_scale>::value ? STAN_CUDA_LP_SCALE : 0u);
  return detail::locscale_lpdf_call<detail::cauchy_checks>(STAN_CUDA_LOCSCALE_CAUCHY, summands, y, mu, sigma);
}

/**
 * GPU body of the vectorised lognormal_lpdf (reference: stan/math/prim/scal/prob/lognormal_lpdf.hpp:19-133): the loops
 * :41-131 become one kernel, including the LOG_ZERO scan of the outcomes :41-45 (STAN_CUDA_LOG_ZERO) and the -log(y)
 * summand the reference ties to include_summand<propto, T_y> (:113-115, STAN_CUDA_LP_Y).  The reference runs its checks
 * before it looks at the sizes (:24-31): with an empty operand they have nothing to reject.
 */
template <bool propto, typename T_y, typename T_loc, typename T_scale>
return_type_t<T_y, T_loc, T_scale> lognormal_lpdf_device(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  if (size_zero(y, mu, sigma)) {
    detail::lognormal_checks::run(y, mu, sigma);
    detail::lognormal_checks::sizes(y, mu, sigma);
    return 0.0;
  }
  const uint32_t summands = STAN_CUDA_LP_MAIN | (include_summand<propto>::value ? STAN_CUDA_LP_CONST : 0u)   // :92-118
                            | (include_summand<propto, T_scale>::value ? STAN_CUDA_LP_SCALE : 0u)
                            | (include_summand<propto, T_y>::value ? STAN_CUDA_LP_Y : 0u);
  return detail::locscale_lpdf_call<detail::lognormal_checks>(STAN_CUDA_LOCSCALE_LOGNORMAL, summands, y, mu, sigma);
}

/**
 * GPU body of the vectorised gamma_lpdf (reference: stan/math/prim/scal/prob/gamma_lpdf.hpp:41-145): the loops :69-141
 * become one kernel -- lgamma and digamma of the shape included --, with the reference's four summand conditions
 * (:118-129) as four flags, its LOG_ZERO for a negative outcome (:69-74) and the zero it leaves in log_y for an outcome of
 * zero (:86-90).
 */
template <bool propto, typename T_y, typename T_shape, typename T_inv_scale>
return_type_t<T_y, T_shape, T_inv_scale> gamma_lpdf_device(const T_y& y, const T_shape& alpha, const T_inv_scale& beta) {
  if (size_zero(y, alpha, beta)) {
    return 0.0;
  }
  const uint32_t summands = (include_summand<propto, T_shape>::value ? STAN_CUDA_LP_DOF : 0u)
                            | (include_summand<propto, T_shape, T_inv_scale>::value ? STAN_CUDA_LP_SCALE : 0u)
                            | (include_summand<propto, T_y, T_shape>::value ? STAN_CUDA_LP_Y : 0u)
                            | (include_summand<propto, T_y, T_inv_scale>::value ? STAN_CUDA_LP_MAIN : 0u);
  return detail::locscale_lpdf_call<detail::gamma_checks>(STAN_CUDA_LOCSCALE_GAMMA, summands, y, alpha, beta);
}

/**
 * GPU body of the vectorised exponential_lpdf (reference: stan/math/prim/scal/prob/exponential_lpdf.hpp:42-95).  The
 * argument block's location slot carries one constant zero: a double operand adds no edge to the tape node, which
 * therefore holds the same operands in the same order as the reference's operands_and_partials<T_y, T_inv_scale>.
 */
template <bool propto, typename T_y, typename T_inv_scale>
return_type_t<T_y, T_inv_scale> exponential_lpdf_device(const T_y& y, const T_inv_scale& beta) {
  if (size_zero(y, beta)) {
    return 0.0;
  }
  const uint32_t summands = (include_summand<propto, T_inv_scale>::value ? STAN_CUDA_LP_SCALE : 0u)           // :76-81
                            | (include_summand<propto, T_y, T_inv_scale>::value ? STAN_CUDA_LP_MAIN : 0u);
  const double unused = 0.0;
  return detail::locscale_lpdf_call<detail::exponential_checks>(STAN_CUDA_LOCSCALE_EXPONENTIAL, summands, y, unused, beta);
}

/**
 * GPU body of the vectorised beta_lpdf (reference: stan/math/prim/scal/prob/beta_lpdf.hpp:40-181): five summands, each
 * under its own include_summand condition (:150-164), three digammas for the shape partials (:170-177).  The reference runs
 * its checks before it looks at the sizes (:48-56).
 */
template <bool propto, typename T_y, typename T_scale_succ, typename T_scale_fail>
return_type_t<T_y, T_scale_succ, T_scale_fail> beta_lpdf_device(const T_y& y, const T_scale_succ& alpha,
                                                               const T_scale_fail& beta) {
  if (size_zero(y, alpha, beta)) {
    detail::beta_checks::run(y, alpha, beta);
    detail::beta_checks::sizes(y, alpha, beta);
    return 0.0;
  }
  const uint32_t summands = (include_summand<propto, T_scale_succ, T_scale_fail>::value ? STAN_CUDA_LP_SCALE : 0u)
                            | (include_summand<propto, T_scale_succ>::value ? STAN_CUDA_LP_DOF : 0u)
                            | (include_summand<propto, T_scale_fail>::value ? STAN_CUDA_LP_LINEAR : 0u)
                            | (include_summand<propto, T_y, T_scale_succ>::value ? STAN_CUDA_LP_Y : 0u)
                            | (include_summand<propto, T_y, T_scale_fail>::value ? STAN_CUDA_LP_MAIN : 0u);
  return detail::locscale_lpdf_call<detail::beta_checks>(STAN_CUDA_LOCSCALE_BETA, summands, y, alpha, beta);
}

/**
 * GPU body of the vectorised weibull_lpdf (reference: stan/math/prim/scal/prob/weibull_lpdf.hpp:29-143), including its
 * LOG_ZERO for a negative outcome (:56-61).  Checks before sizes, as there (:38-45).
 */
template <bool propto, typename T_y, typename T_shape, typename T_scale>
return_type_t<T_y, T_shape, T_scale> weibull_lpdf_device(const T_y& y, const T_shape& alpha, const T_scale& sigma) {
  if (size_zero(y, alpha, sigma)) {
    detail::weibull_checks::run(y, alpha, sigma);
    detail::weibull_checks::sizes(y, alpha, sigma);
    return 0.0;
  }
  const uint32_t summands = (include_summand<propto, T_shape>::value ? STAN_CUDA_LP_DOF : 0u)
                            | (include_summand<propto, T_y, T_shape>::value ? STAN_CUDA_LP_Y : 0u)
                            | (include_summand<propto, T_shape, T_scale>::value ? STAN_CUDA_LP_SCALE : 0u) | STAN_CUDA_LP_MAIN;
  return detail::locscale_lpdf_call<detail::weibull_checks>(STAN_CUDA_LOCSCALE_WEIBULL, summands, y, alpha, sigma);
}

/**
 * GPU body of the vectorised inv_gamma_lpdf (reference: stan/math/prim/scal/prob/inv_gamma_lpdf.hpp:33-146): gamma_lpdf's
 * slots and flags with the terms of :115-138 and LOG_ZERO for an outcome that is not positive (:58-63).  Checks before
 * sizes, as there (:41-48).
 */
template <bool propto, typename T_y, typename T_shape, typename T_scale>
return_type_t<T_y, T_shape, T_scale> inv_gamma_lpdf_device(const T_y& y, const T_shape& alpha, const T_scale& beta) {
  if (size_zero(y, alpha, beta)) {
    detail::inv_gamma_checks::run(y, alpha, beta);
    detail::inv_gamma_checks::sizes(y, alpha, beta);
    return 0.0;
  }
  const uint32_t summands = (include_summand<propto, T_shape>::value ? STAN_CUDA_LP_DOF : 0u)
                            | (include_summand<propto, T_shape, T_scale>::value ? STAN_CUDA_LP_SCALE : 0u)
                            | (include_summand<propto, T_y, T_shape>::value ? STAN_CUDA_LP_Y : 0u)
                            | (include_summand<propto, T_y, T_scale>::value ? STAN_CUDA_LP_MAIN : 0u);
  return detail::locscale_lpdf_call<detail::inv_gamma_checks>(STAN_CUDA_LOCSCALE_INV_GAMMA, summands, y, alpha, beta);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
