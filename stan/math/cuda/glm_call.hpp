#ifndef STAN_MATH_CUDA_GLM_CALL_HPP
#define STAN_MATH_CUDA_GLM_CALL_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <string>
#include <utility>
#include <system_error>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/**
 * Argument block of one GLM evaluation, filled from the reference's argument kinds.
 * The outputs point straight into the partials_ storage of the caller's operands_and_partials
 * edges, so the device results land where build() reads them.
 */
template <typename T_x, typename T_alpha, typename T_beta>
class glm_call {
 public:
  glm_call(int family, const T_x& x, const T_alpha& alpha, const T_beta& beta)
      : x_val_(x), alpha_val_(alpha), beta_val_(beta), logp_(0) {
    a_ = stan_cuda_glm_args();
    a_.family = family;
    a_.n = x.rows();
    a_.k = x.cols();
    matrix_layout(x_val_.matrix(), &a_.x, &a_.ldx, &a_.x_row_major);
    a_.x_per_call = x_val_.per_call() ? 1 : 0;
    a_.alpha = alpha_val_.data();
    a_.alpha_len = alpha_val_.size();
    if (beta_val_.size() == static_cast<size_t>(a_.k)) {
      a_.beta = beta_val_.data();
    } else {  // scalar weight with K != 1: broadcast
      beta_bcast_.assign(a_.k, beta_val_.size() ? beta_val_.data()[0] : 0.0);
      a_.beta = beta_bcast_.data();
    }
    a_.lp = &logp_;
    a_.aux = aux_;
    aux_[0] = aux_[1] = aux_[2] = aux_[3] = 0;
  }

  stan_cuda_glm_args& args() { return a_; }
  double logp() const { return logp_; }
  /** family-specific finiteness witness the reference tests (logp | sum d | sum s^2) */
  double witness() const { return aux_[1]; }
  /** sum over rows of dL/d(linear predictor) */
  double sum_d() const { return aux_[0]; }
  const double* beta_data() const { return a_.beta; }
  const double* alpha_data() const { return a_.alpha; }

  /**
   * Runs the evaluation.  Returns STAN_CUDA_OK, STAN_CUDA_NONFINITE (outputs written, the caller
   * runs the reference's lazy check_finite sequence) or STAN_CUDA_BAD_OUTCOME (the device-resident
   * outcomes violate the family's domain; nothing evaluated; the caller runs the reference's eager
   * outcome check, which throws).  Failures of the device layer throw std::system_error.
   */
  int run(const char* function) {
    // A scalar weight broadcast to K != 1 columns has ONE partials slot (the caller pointed dbeta at it): the
    // device's K components are received in a temporary and summed into that slot, instead of being copied
    // over the neighbouring arena memory.
    double* const caller_dbeta = a_.dbeta;
    std::vector<double> dbeta_wide;
    if (!beta_bcast_.empty() && caller_dbeta != nullptr && a_.k != 1) {
      dbeta_wide.assign(a_.k, 0.0);
      a_.dbeta = dbeta_wide.data();
    }
    const int rc = stan_cuda_glm_eval(&a_);
    if (!dbeta_wide.empty()) {
      double sum = 0.0;
      for (size_t j = 0; j < dbeta_wide.size(); ++j) {
        sum += dbeta_wide[j];
      }
      *caller_dbeta = sum;
      a_.dbeta = caller_dbeta;
    }
    check_cuda_rc(function, rc);
    return rc;
  }

  /**
   * The device said the outcomes are out of range but the reference's host check (run by the
   * caller just before this) accepted them: the device copy is stale (host y was rewritten in
   * place).  Drop it, refill from the host data and evaluate again.
   */
  int rerun_after_stale_outcomes(const char* function) {
    stan_cuda_cache_invalidate(a_.y);
    stan_cuda_cache_invalidate(a_.x);
    const int rc = run(function);
    if (rc == STAN_CUDA_BAD_OUTCOME) {
      throw std::system_error(STAN_CUDA_ERR_STATE, std::generic_category(),
                              std::string(function) + ": device copy of the outcome vector disagrees with the host");
    }
    return rc;
  }

  /** Host recomputation of X * beta + alpha, used only on the error path so that the reference's
   * own check_finite call sees the same object it would have seen. */
  Eigen::Matrix<double, Eigen::Dynamic, 1> host_linear_predictor(const T_x& /*x*/) const {
    Eigen::Map<const Eigen::Matrix<double, Eigen::Dynamic, 1>> b(a_.beta, a_.k);
    Eigen::Matrix<double, Eigen::Dynamic, 1> theta = x_val_.matrix() * b;
    if (a_.alpha_len == 1) {
      theta.array() += a_.alpha[0];
    } else {
      theta += Eigen::Map<const Eigen::Matrix<double, Eigen::Dynamic, 1>>(a_.alpha, a_.n);
    }
    return theta;
  }

 private:
  design_values<T_x> x_val_;
  values<T_alpha> alpha_val_;
  values<T_beta> beta_val_;
  std::vector<double> beta_bcast_;
  stan_cuda_glm_args a_;
  double logp_;
  double aux_[4];
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
