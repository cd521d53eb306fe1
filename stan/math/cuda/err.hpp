#ifndef STAN_MATH_CUDA_ERR_HPP
#define STAN_MATH_CUDA_ERR_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <string>
#include <system_error>

namespace stan {
namespace math {
namespace cuda {

/**
 * Maps a negative return code of the C ABI onto std::system_error, the way the OpenCL backend
 * maps cl::Error (stan/math/opencl/err/check_opencl.hpp).  STAN_CUDA_NONFINITE (1) is not an
 * error here: the caller runs the reference's lazy check_finite sequence.
 */
inline void check_cuda_rc(const char* function, int rc) {
  if (rc >= 0) {
    return;
  }
  std::string msg = std::string(function) + ": " + stan_cuda_last_error();
  throw std::system_error(rc, std::generic_category(), msg);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
