#ifndef STAN_MATH_CUDA_HPP
#define STAN_MATH_CUDA_HPP
// Umbrella of the CUDA subsystem (sibling of stan/math/opencl/).  Everything is inert unless the
// translation unit is compiled with -DSTAN_CUDA (cf. STAN_OPENCL, make/compiler_flags:165-178)
// and linked against libstanmath_cuda.so.
#ifdef STAN_CUDA
#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/cuda/context.hpp>
#include <stan/math/cuda/marshal.hpp>
#include <stan/math/cuda/glm_call.hpp>
#include <stan/math/cuda/bernoulli_logit_glm_lpmf.hpp>
#include <stan/math/cuda/poisson_log_glm_lpmf.hpp>
#include <stan/math/cuda/normal_id_glm_lpdf.hpp>
#include <stan/math/cuda/neg_binomial_2_log_glm_lpmf.hpp>
#include <stan/math/cuda/ordered_logistic_glm_lpmf.hpp>
#include <stan/math/cuda/categorical_logit_glm_lpmf.hpp>
#include <stan/math/cuda/normal_lpdf.hpp>
#include <stan/math/cuda/locscale_lpdf.hpp>
#include <stan/math/cuda/student_t_lpdf.hpp>
#include <stan/math/cuda/int_density.hpp>
#include <stan/math/cuda/matrix_cuda.hpp>
#include <stan/math/cuda/multiply.hpp>
#include <stan/math/cuda/rev_multiply.hpp>
#include <stan/math/cuda/cholesky_decompose.hpp>
#include <stan/math/cuda/rev_cholesky_decompose.hpp>
#include <stan/math/cuda/mdivide_left_tri.hpp>
#include <stan/math/cuda/rev_mdivide_left_tri.hpp>
#include <stan/math/cuda/gp_exp_quad_cov.hpp>
#include <stan/math/cuda/map_rect_devices.hpp>
#endif
#endif
