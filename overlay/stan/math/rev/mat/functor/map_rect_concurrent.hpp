// Overlay of stan/math/rev/mat/functor/map_rect_concurrent.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  With STAN_CUDA the job loop runs
// every shard inside a device scope (stan/math/cuda/map_rect_devices.hpp); without it the
// reference implementation is used unchanged.
#ifndef STAN_MATH_CUDA_OVERLAY_MAP_RECT_CONCURRENT_HPP
#define STAN_MATH_CUDA_OVERLAY_MAP_RECT_CONCURRENT_HPP

#define map_rect_concurrent map_rect_concurrent_reference_cpu
#include_next <stan/math/rev/mat/functor/map_rect_concurrent.hpp>
#undef map_rect_concurrent

#ifdef STAN_CUDA
#include <stan/math/cuda/map_rect_devices.hpp>
#endif

namespace stan {
namespace math {
namespace internal {

// definition of the template declared in stan/math/prim/mat/functor/map_rect_concurrent.hpp:14-23
template <int call_id, typename F, typename T_shared_param, typename T_job_param>
Eigen::Matrix<typename stan::return_type<T_shared_param, T_job_param>::type, Eigen::Dynamic, 1>
map_rect_concurrent(const Eigen::Matrix<T_shared_param, Eigen::Dynamic, 1>& shared_params,
                    const std::vector<Eigen::Matrix<T_job_param, Eigen::Dynamic, 1>>& job_params,
                    const std::vector<std::vector<double>>& x_r, const std::vector<std::vector<int>>& x_i,
                    std::ostream* msgs) {
#ifdef STAN_CUDA
  return cuda::map_rect_on_devices<call_id, F>(shared_params, job_params, x_r, x_i, msgs);
#else
  return map_rect_concurrent_reference_cpu<call_id, F>(shared_params, job_params, x_r, x_i, msgs);
#endif
}

}  // namespace internal
}  // namespace math
}  // namespace stan
#endif
