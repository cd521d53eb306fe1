#ifndef STAN_MATH_CUDA_CONTEXT_HPP
#define STAN_MATH_CUDA_CONTEXT_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <string>

namespace stan {
namespace math {
namespace cuda {

/** Select the devices of this process (0 = all visible; env STAN_CUDA_DEVICES). Optional: the first
 * evaluation initialises implicitly.  Counterpart of opencl_context (opencl/opencl_context.hpp). */
inline void init(int n_devices = 0) { check_cuda_rc("stan::math::cuda::init", stan_cuda_init(n_devices)); }
inline int device_count() { return stan_cuda_device_count(); }
inline std::string description() { return stan_cuda_describe(); }

/** Drop the device copy of a data matrix / outcome vector whose host contents changed in place. */
template <typename T>
inline void invalidate(const T* host_ptr) {
  stan_cuda_cache_invalidate(static_cast<const void*>(host_ptr));
}
inline void clear_cache() { stan_cuda_cache_clear(); }

/** RAII form of the map_rect shard scope. */
class shard_scope {
 public:
  shard_scope(int call_id, int job, int device) { stan_cuda_shard_scope_enter(call_id, job, device); }
  ~shard_scope() { stan_cuda_shard_scope_exit(); }
  shard_scope(const shard_scope&) = delete;
  shard_scope& operator=(const shard_scope&) = delete;
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
