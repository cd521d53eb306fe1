#ifndef STAN_MATH_CUDA_NORMAL_LPDF_HPP
#define STAN_MATH_CUDA_NORMAL_LPDF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/cuda/arena_partials.hpp>
#include <stan/math/prim/scal/err/check_consistent_sizes.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/scal/err/check_positive.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>
#include <stdexcept>
#include <string>

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
 * Values of one operand as contiguous doubles for the device.  Double containers are borrowed;
 * var / int containers are gathered -- one pass over the caller's container (seq_access: no scalar_seq_view
 * copy), in parallel over TBB workers when STAN_THREADS is on -- straight into the thread's page-locked staging
 * area (stan_cuda_host_staging), so the upload is a DMA without a bounce buffer.  No domain checks here: the kernel tests every element.
 */
template <typename T>
struct staged_operand {
  static constexpr bool borrowed = false;
  static size_t bytes(const T& v) { return ((length(v) + 1) & ~size_t(1)) * sizeof(double); }
  static const double* fill(const T& v, double* dst) {
    for_each_chunk(length(v), [&](size_t begin, size_t end) {
      for (size_t i = begin; i != end; ++i) {
        dst[i] = value_of_rec(seq_access<T>::at(v, i));
      }
    });
    return dst;
  }
};
template <>
struct staged_operand<std::vector<double>> {
  static constexpr bool borrowed = true;
  static size_t bytes(const std::vector<double>&) { return 0; }
  static const double* fill(const std::vector<double>& v, double*) { return v.data(); }
};
template <int R, int C, int O, int MR, int MC>
struct staged_operand<Eigen::Matrix<double, R, C, O, MR, MC>> {
  static constexpr bool borrowed = true;
  static size_t bytes(const Eigen::Matrix<double, R, C, O, MR, MC>&) { return 0; }
  static const double* fill(const Eigen::Matrix<double, R, C, O, MR, MC>& v, double*) { return v.data(); }
};

/**
 * GPU body of the vectorised normal_lpdf (reference: stan/math/prim/scal/prob/normal_lpdf.hpp).
 * The loop :78-110 becomes one kernel that returns the log-density and the partials of every
 * non-constant operand (per element for containers, summed for scalars -- what the reference
 * accumulates through broadcast_array).  The eager domain checks :51-53 are answered by the same
 * kernel (STAN_CUDA_BAD_OPERAND); the reference's check functions then run on the host only to
 * throw the identical exception.  Size checks :54-55 stay in front.
 */
template <bool propto, typename T_y, typename T_loc, typename T_scale>
return_type_t<T_y, T_loc, T_scale> normal_lpdf_device(const T_y& y, const T_loc& mu, const T_scale& sigma) {
  static const char* function = "normal_lpdf";
  if (size_zero(y, mu, sigma)) {
    return 0.0;
  }
  if (!include_summand<propto, T_y, T_loc, T_scale>::value) {  // nothing goes to the device
    check_not_nan(function, "Random variable", y);
    check_finite(function, "Location parameter", mu);
    check_positive(function, "Scale parameter", sigma);
    check_consistent_sizes(function, "Random variable", y, "Location parameter", mu, "Scale parameter", sigma);
    return 0.0;
  }
  // the reference reports a domain error before a size error (:51-55): if the sizes disagree,
  // run its checks in its order and let the first one throw
  const size_t N = max_size(y, mu, sigma);
  if ((is_vector<T_y>::value && length(y) != N) || (is_vector<T_loc>::value && length(mu) != N)
      || (is_vector<T_scale>::value && length(sigma) != N)) {
    check_not_nan(function, "Random variable", y);
    check_finite(function, "Location parameter", mu);
    check_positive(function, "Scale parameter", sigma);
    check_consistent_sizes(function, "Random variable", y, "Location parameter", mu, "Scale parameter", sigma);
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
  // operands_and_partials, built in place; containers of var leave their values in the staging area in the same pass
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
  a.flags = STAN_CUDA_LP_MAIN;
  if (include_summand<propto>::value) {
    a.flags |= STAN_CUDA_LP_CONST;
  }
  if (include_summand<propto, T_scale>::value) {
    a.flags |= STAN_CUDA_LP_SCALE;
  }
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
  const int rc = stan_cuda_normal_lpdf_eval(&a);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OPERAND) {
    check_not_nan(function, "Random variable", y);
    check_finite(function, "Location parameter", mu);
    check_positive(function, "Scale parameter", sigma);
    throw std::domain_error(std::string(function) + ": operand outside the domain (device check)");
  }
  return ops_partials.build(logp);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
