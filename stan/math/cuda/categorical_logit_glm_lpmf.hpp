#ifndef STAN_MATH_CUDA_CATEGORICAL_LOGIT_GLM_LPMF_HPP
#define STAN_MATH_CUDA_CATEGORICAL_LOGIT_GLM_LPMF_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/prim/scal/err/check_consistent_size.hpp>
#include <stan/math/prim/scal/err/check_bounded.hpp>
#include <stan/math/prim/scal/err/check_finite.hpp>
#include <stan/math/prim/scal/err/check_size_match.hpp>
#include <stan/math/prim/scal/fun/size_zero.hpp>
#include <string>
#include <system_error>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/** std::vector<int> classes, dynamic data matrix x, double / var intercepts and weight matrix. */
template <typename T_y, typename T_x_scalar, int T_x_rows, typename T_alpha_scalar, typename T_beta_scalar>
struct categorical_logit_glm_eligible
    : std::integral_constant<bool, std::is_same<T_y, std::vector<int>>::value
                                       && std::is_same<T_x_scalar, double>::value && T_x_rows == Eigen::Dynamic
                                       && is_double_or_var<T_alpha_scalar>::value
                                       && is_double_or_var<T_beta_scalar>::value> {};

/**
 * GPU body of categorical_logit_glm_lpmf (reference:
 * stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:38-190).  Size checks :56-62 unchanged;
 * check_bounded(y, 1, N_classes) :63-64 is answered by the device from the range of y computed at
 * cache fill; the lazy finiteness checks :109-113 run on the host only when the device reports a
 * non-finite log-density.  Any shape is served: up to 16 classes with classes * padded attributes <= 4096 and at most
 * 512 attributes by the fused single-pass kernels, everything else by the general three-pass path
 * (csrc/glm_categorical_wide.cu).
 */
template <bool propto, typename T_alpha_scalar, typename T_beta_scalar>
return_type_t<T_alpha_scalar, T_beta_scalar> categorical_logit_glm_lpmf_device(
    const std::vector<int>& y, const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& x,
    const Eigen::Matrix<T_alpha_scalar, Eigen::Dynamic, 1>& alpha,
    const Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, Eigen::Dynamic>& beta) {
  typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> T_x;
  typedef Eigen::Matrix<T_alpha_scalar, Eigen::Dynamic, 1> T_alpha;
  typedef Eigen::Matrix<T_beta_scalar, Eigen::Dynamic, Eigen::Dynamic> T_beta;
  static const char* function = "categorical_logit_glm_lpmf";
  const size_t N_instances = x.rows();
  const size_t N_attributes = x.cols();
  const size_t N_classes = beta.cols();

  check_consistent_size(function, "Vector of dependent variables", y, N_instances);
  check_consistent_size(function, "Intercept vector", alpha, N_classes);
  check_size_match(function, "x.cols()", N_attributes, "beta.rows()", beta.rows());
  if (size_zero(y) || N_classes == 1 || !include_summand<propto, double, T_alpha_scalar, T_beta_scalar>::value) {
    check_bounded(function, "categorical outcome out of support", y, 1, N_classes);  // nothing goes to the device
    return 0;
  }

  std::vector<double> alpha_val(N_classes), beta_val(N_attributes * N_classes);
  for (size_t c = 0; c < N_classes; ++c) {
    alpha_val[c] = value_of_rec(alpha.coeff(c));
  }
  for (size_t q = 0; q < beta_val.size(); ++q) {
    beta_val[q] = value_of_rec(beta.coeff(q));  // column-major K x C, as the C ABI expects
  }
  operands_and_partials<T_x, T_alpha, T_beta> ops_partials(x, alpha, beta);
  double logp = 0;
  stan_cuda_glm_args a = stan_cuda_glm_args();
  a.family = STAN_CUDA_CATEGORICAL_LOGIT;
  a.n = N_instances;
  a.k = N_attributes;
  a.n_classes = N_classes;
  matrix_layout(x, &a.x, &a.ldx, &a.x_row_major);
  a.y = y.data();
  a.y_is_double = 0;
  a.alpha = alpha_val.data();
  a.alpha_len = N_classes;
  a.beta = beta_val.data();
  a.lp = &logp;
  a.flags = STAN_CUDA_LP_MAIN;
  if (!is_constant_all<T_alpha_scalar>::value) {
    a.flags |= STAN_CUDA_NEED_DALPHA;
    a.dalpha = partials_data(ops_partials.edge2_);
  }
  if (!is_constant_all<T_beta_scalar>::value) {
    a.flags |= STAN_CUDA_NEED_DBETA;
    a.dbeta = partials_data(ops_partials.edge3_);
  }
  int rc = stan_cuda_glm_eval(&a);
  check_cuda_rc(function, rc);
  if (rc == STAN_CUDA_BAD_OUTCOME) {
    check_bounded(function, "categorical outcome out of support", y, 1, N_classes);  // throws, as in the reference
    stan_cuda_cache_invalidate(a.y);  // the host check passed: the device copy is stale
    stan_cuda_cache_invalidate(a.x);
    rc = stan_cuda_glm_eval(&a);
    check_cuda_rc(function, rc);
    if (rc == STAN_CUDA_BAD_OUTCOME) {
      throw std::system_error(STAN_CUDA_ERR_STATE, std::generic_category(),
                              std::string(function) + ": device copy of the outcome vector disagrees with the host");
    }
  }
  if (rc == STAN_CUDA_NONFINITE) {  // :109-113
    check_finite(function, "Weight vector", beta);
    check_finite(function, "Intercept", alpha);
    check_finite(function, "Matrix of independent variables", x);
  }
  return ops_partials.build(logp);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
