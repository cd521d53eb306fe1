// Overlay of stan/math/prim/mat/fun/gp_exp_quad_cov.hpp (how overlays work: see
// overlay/stan/math/prim/mat/fun/multiply.hpp).  The reference's transparent GPU hooks are six full specialisations for
// double arguments (:344-655, #ifdef STAN_OPENCL); with -DSTAN_CUDA the same six argument kinds build the covariance
// matrix on the B200 once it has gp_exp_quad_cov_worth_transfer elements.  Checks and their order are those of the
// reference's generic templates (:175-342), which remain reachable as gp_exp_quad_cov_reference_cpu.
#ifndef STAN_MATH_CUDA_OVERLAY_PRIM_GP_EXP_QUAD_COV_HPP
#define STAN_MATH_CUDA_OVERLAY_PRIM_GP_EXP_QUAD_COV_HPP

#define gp_exp_quad_cov gp_exp_quad_cov_reference_cpu
#include_next <stan/math/prim/mat/fun/gp_exp_quad_cov.hpp>
#undef gp_exp_quad_cov

#ifdef STAN_CUDA
#include <stan/math/cuda/gp_exp_quad_cov.hpp>
#endif
#include <type_traits>
#include <vector>

namespace stan {
namespace math {
namespace cuda_dispatch {

template <typename TX, typename TS, typename TL, typename Enable = void>
struct gp_cov_route {
  template <typename X_, typename S_, typename L_>
  static auto run(const std::vector<X_>& x, const S_& sigma, const L_& l) -> decltype(gp_exp_quad_cov_reference_cpu(x, sigma, l)) {
    return gp_exp_quad_cov_reference_cpu(x, sigma, l);
  }
};
template <typename TX1, typename TX2, typename TS, typename TL, typename Enable = void>
struct gp_cross_cov_route {
  template <typename X1_, typename X2_, typename S_, typename L_>
  static auto run(const std::vector<X1_>& x1, const std::vector<X2_>& x2, const S_& sigma, const L_& l)
      -> decltype(gp_exp_quad_cov_reference_cpu(x1, x2, sigma, l)) {
    return gp_exp_quad_cov_reference_cpu(x1, x2, sigma, l);
  }
};

#ifdef STAN_CUDA
template <typename T>
struct is_gp_point : std::integral_constant<bool, std::is_same<T, double>::value || std::is_same<T, Eigen::VectorXd>::value> {};

/** one sample, scalar length scale (:175-198): x of double or of VectorXd */
template <typename TX>
struct gp_cov_route<TX, double, double, typename std::enable_if<is_gp_point<TX>::value>::type> {
  static Eigen::MatrixXd run(const std::vector<TX>& x, const double& sigma, const double& length_scale) {
    if (static_cast<int64_t>(x.size()) * static_cast<int64_t>(x.size()) < cuda::gp_exp_quad_cov_worth_transfer()) {
      return gp_exp_quad_cov_reference_cpu(x, sigma, length_scale);
    }
    check_positive("gp_exp_quad_cov", "magnitude", sigma);
    check_positive("gp_exp_quad_cov", "length scale", length_scale);
    for (size_t n = 0; n < x.size(); ++n) {
      check_not_nan("gp_exp_quad_cov", "x", x[n]);
    }
    cuda::gp_points p;
    cuda::pack_points(x, nullptr, p);
    return cuda::gp_exp_quad_cov_on_device(p, nullptr, square(sigma), -0.5 / square(length_scale));
  }
};

/** one sample, one length scale per dimension (:215-236) */
template <>
struct gp_cov_route<Eigen::VectorXd, double, std::vector<double>, void> {
  static Eigen::MatrixXd run(const std::vector<Eigen::VectorXd>& x, const double& sigma, const std::vector<double>& length_scale) {
    if (static_cast<int64_t>(x.size()) * static_cast<int64_t>(x.size()) < cuda::gp_exp_quad_cov_worth_transfer()) {
      return gp_exp_quad_cov_reference_cpu(x, sigma, length_scale);
    }
    check_positive_finite("gp_exp_quad_cov", "magnitude", sigma);
    check_positive_finite("gp_exp_quad_cov", "length scale", length_scale);
    check_size_match("gp_exp_quad_cov", "x dimension", x[0].size(), "number of length scales", length_scale.size());
    cuda::gp_points p;
    cuda::pack_points(x, &length_scale, p);
    return cuda::gp_exp_quad_cov_on_device(p, nullptr, square(sigma), -0.5);
  }
};

/** two samples, scalar length scale (:258-286) */
template <typename TX>
struct gp_cross_cov_route<TX, TX, double, double, typename std::enable_if<is_gp_point<TX>::value>::type> {
  static Eigen::MatrixXd run(const std::vector<TX>& x1, const std::vector<TX>& x2, const double& sigma, const double& length_scale) {
    if (static_cast<int64_t>(x1.size()) * static_cast<int64_t>(x2.size()) < cuda::gp_exp_quad_cov_worth_transfer()) {
      return gp_exp_quad_cov_reference_cpu(x1, x2, sigma, length_scale);
    }
    const char* function_name = "gp_exp_quad_cov";
    check_positive(function_name, "magnitude", sigma);
    check_positive(function_name, "length scale", length_scale);
    for (size_t i = 0; i < x1.size(); ++i) {
      check_not_nan(function_name, "x1", x1[i]);
    }
    for (size_t i = 0; i < x2.size(); ++i) {
      check_not_nan(function_name, "x2", x2[i]);
    }
    cuda::gp_points p1, p2;
    cuda::pack_points(x1, nullptr, p1);
    cuda::pack_points(x2, nullptr, p2);
    if (p1.dim != p2.dim) {
      return gp_exp_quad_cov_reference_cpu(x1, x2, sigma, length_scale);
    }
    return cuda::gp_exp_quad_cov_on_device(p1, &p2, square(sigma), -0.5 / square(length_scale));
  }
};

/** two samples, one length scale per dimension (:307-342) */
template <>
struct gp_cross_cov_route<Eigen::VectorXd, Eigen::VectorXd, double, std::vector<double>, void> {
  static Eigen::MatrixXd run(const std::vector<Eigen::VectorXd>& x1, const std::vector<Eigen::VectorXd>& x2, const double& sigma,
                             const std::vector<double>& length_scale) {
    if (static_cast<int64_t>(x1.size()) * static_cast<int64_t>(x2.size()) < cuda::gp_exp_quad_cov_worth_transfer()) {
      return gp_exp_quad_cov_reference_cpu(x1, x2, sigma, length_scale);
    }
    const char* function_name = "gp_exp_quad_cov";
    for (size_t i = 0; i < x1.size(); ++i) {
      check_not_nan(function_name, "x1", x1[i]);
    }
    for (size_t i = 0; i < x2.size(); ++i) {
      check_not_nan(function_name, "x2", x2[i]);
    }
    check_positive_finite(function_name, "magnitude", sigma);
    check_positive_finite(function_name, "length scale", length_scale);
    check_size_match(function_name, "x dimension", x1[0].size(), "number of length scales", length_scale.size());
    check_size_match(function_name, "x dimension", x2[0].size(), "number of length scales", length_scale.size());
    cuda::gp_points p1, p2;
    cuda::pack_points(x1, &length_scale, p1);
    cuda::pack_points(x2, &length_scale, p2);
    return cuda::gp_exp_quad_cov_on_device(p1, &p2, square(sigma), -0.5);
  }
};
#endif
}  // namespace cuda_dispatch

template <typename TX, typename TS, typename TL>
inline auto gp_exp_quad_cov(const std::vector<TX>& x, const TS& sigma, const TL& length_scale)
    -> decltype(gp_exp_quad_cov_reference_cpu(x, sigma, length_scale)) {
  return cuda_dispatch::gp_cov_route<TX, TS, TL>::run(x, sigma, length_scale);
}

template <typename TX1, typename TX2, typename TS, typename TL>
inline auto gp_exp_quad_cov(const std::vector<TX1>& x1, const std::vector<TX2>& x2, const TS& sigma, const TL& length_scale)
    -> decltype(gp_exp_quad_cov_reference_cpu(x1, x2, sigma, length_scale)) {
  return cuda_dispatch::gp_cross_cov_route<TX1, TX2, TS, TL>::run(x1, x2, sigma, length_scale);
}

}  // namespace math
}  // namespace stan
#endif
