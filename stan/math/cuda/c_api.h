/* stan/math/cuda/c_api.h -- the C ABI of the CUDA subsystem as seen from the Stan headers.
 * The declarations live in include/stan_cuda.h (one copy, shared with the ctypes binding). */
#ifndef STAN_MATH_CUDA_C_API_REDIRECT_H
#define STAN_MATH_CUDA_C_API_REDIRECT_H
#include "../../../include/stan_cuda.h"
#endif
