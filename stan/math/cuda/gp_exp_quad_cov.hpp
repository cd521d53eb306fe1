#ifndef STAN_MATH_CUDA_GP_EXP_QUAD_COV_HPP
#define STAN_MATH_CUDA_GP_EXP_QUAD_COV_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/scal/err/check_positive.hpp>
#include <stan/math/prim/scal/err/check_positive_finite.hpp>
#include <stan/math/prim/scal/err/check_size_match.hpp>
#include <stan/math/prim/scal/fun/square.hpp>
#include <cstdint>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/**
 * Smallest covariance matrix (elements) that is built on the device: the reference's knobs are
 * gp_exp_quad_cov_simple / _complex (stan/math/opencl/opencl_context.hpp:175-176), counted on inputs + outputs.
 */
inline int64_t& gp_exp_quad_cov_worth_transfer() {
  static int64_t elements = 65536;
  return elements;
}

/** Points of a covariance function, dimension-major (coordinate d of point i at data[d * n + i]). */
struct gp_points {
  int64_t n, dim;
  const double* data;
  std::vector<double> packed;
};
inline void pack_points(const std::vector<double>& x, const std::vector<double>* /*length_scale*/, gp_points& p) {
  p.n = x.size();
  p.dim = 1;
  p.data = x.data();
}
/** with `length_scale`: x[i][d] / length_scale[d], what divide_columns computes (prim/mat/fun/divide_columns.hpp) */
inline void pack_points(const std::vector<Eigen::VectorXd>& x, const std::vector<double>* length_scale, gp_points& p) {
  p.n = x.size();
  p.dim = x.empty() ? 1 : x[0].size();
  p.packed.resize(static_cast<size_t>(p.n) * p.dim);
  for (int64_t i = 0; i < p.n; ++i) {
    for (int64_t d = 0; d < p.dim; ++d) {
      p.packed[d * p.n + i] = length_scale ? x[i][d] / (*length_scale)[d] : x[i][d];
    }
  }
  p.data = p.packed.data();
}

inline Eigen::MatrixXd gp_exp_quad_cov_on_device(const gp_points& x1, const gp_points* x2, double sigma_sq, double factor) {
  const int64_t n2 = x2 ? x2->n : x1.n;
  Eigen::MatrixXd cov(x1.n, n2);
  check_cuda_rc("gp_exp_quad_cov", stan_cuda_gp_exp_quad_cov(x1.n, n2, x1.dim, x1.data, x2 ? x2->data : nullptr, sigma_sq, factor,
                                                             cov.data(), x1.n > 0 ? x1.n : 1));
  return cov;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
