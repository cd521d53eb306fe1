/* stan_cuda.h -- the C ABI of libstanmath_cuda.so (B200 / sm_100a).
 *
 * This is the drop-in boundary of the stan/math/cuda/ subsystem: plain C, plain pointers and
 * sizes, no Eigen / Stan / torch types.  The C++ host layer (the headers under stan/math/cuda/) binds these
 * entry points from inside the reference's own function signatures; tests and bench.py bind
 * them through ctypes.  No exception crosses this boundary.
 *
 * What each entry point replaces in the reference (paths relative to the Stan Math 3.0.0 tree):
 *
 *   stan_cuda_glm_eval           the numeric body of
 *                                  bernoulli_logit_glm_lpmf  stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:80-135
 *                                  poisson_log_glm_lpmf      stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp:82-129
 *                                  normal_id_glm_lpdf        stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp:89-172
 *                                and the OpenCL overloads + kernels that did the same on matrix_cl
 *                                  stan/math/opencl/prim/{bernoulli_logit,poisson_log,normal_id}_glm_*.hpp
 *                                  stan/math/opencl/kernels/{bernoulli_logit,poisson_log,normal_id}_glm_*.hpp
 *                                  stan/math/opencl/kernels/matrix_multiply.hpp:354-386 (row_vector_matrix_multiply)
 *                                  neg_binomial_2_log_glm_lpmf stan/math/prim/mat/prob/neg_binomial_2_log_glm_lpmf.hpp:108-197
 *                                  ordered_logistic_glm_lpmf   stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:93-199
 *                                  categorical_logit_glm_lpmf  stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:74-190
 *   stan_cuda_glm_eval_batch     the jobs of one map_rect call whose functor is a single GLM call, one launch per
 *                                device; the per-job loop of
 *                                  stan/math/rev/mat/functor/map_rect_concurrent.hpp:39-54
 *   stan_cuda_normal_lpdf_eval   the loop of normal_lpdf      stan/math/prim/scal/prob/normal_lpdf.hpp:67-110
 *   stan_cuda_student_t_lpdf_eval the loops of student_t_lpdf stan/math/prim/scal/prob/student_t_lpdf.hpp:85-199
 *   stan_cuda_density_eval       the loops of the vectorised elementwise densities
 *                                  bernoulli_logit_lpmf      stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:51-80
 *                                  poisson_log_lpmf          stan/math/prim/scal/prob/poisson_log_lpmf.hpp:51-93
 *   stan_cuda_data_* / cache_*   matrix_cl + to_matrix_cl      stan/math/opencl/matrix_cl.hpp:32-454, stan/math/opencl/copy.hpp:37-93
 *                                and the cache-by-call_id of the MPI backend
 *                                  stan/math/prim/mat/functor/mpi_parallel_call.hpp:40-89
 *   stan_cuda_init / describe    opencl_context               stan/math/opencl/opencl_context.hpp:54-383
 *   stan_cuda_shard_scope_*      the per-job loop of map_rect_concurrent
 *                                  stan/math/rev/mat/functor/map_rect_concurrent.hpp:39-54
 *   stan_cuda_comm_*             (no counterpart: the reference is single-device; the closest is
 *                                  the Boost.MPI cluster, stan/math/prim/arr/functor/mpi_cluster.hpp:126-241)
 *
 * Return convention (every int-returning function):
 *    0  STAN_CUDA_OK
 *    1  STAN_CUDA_NONFINITE   the evaluation ran but its result is not finite; outputs are written;
 *                             the C++ wrapper then runs the reference's lazy check_finite sequence so
 *                             the same std::domain_error is thrown (bernoulli_logit_glm_lpmf.hpp:106-110)
 *    2  STAN_CUDA_BAD_OUTCOME the device-resident outcome vector violates the family's domain (bernoulli:
 *                             outside [0, 1], bernoulli_logit_glm_lpmf.hpp:70; poisson: negative,
 *                             poisson_log_glm_lpmf.hpp:70; ordered_logistic: outside [1, cuts_len + 1],
 *                             ordered_logistic_glm_lpmf.hpp:78; categorical_logit: outside [1, n_classes],
 *                             categorical_logit_glm_lpmf.hpp:63).  The range of y is computed ONCE, on the device,
 *                             when y is cached (the data is immutable), so the per-call O(N) host loop of
 *                             check_bounded / check_nonnegative disappears; nothing was evaluated; the C++
 *                             wrapper runs the reference's own check to throw the same std::domain_error
 *    3  STAN_CUDA_BAD_OPERAND (normal_lpdf) an operand violates the domain the reference checks eagerly
 *                             (y NaN, mu not finite, sigma not > 0: normal_lpdf.hpp:51-53).  The kernel
 *                             tests every element on its way through, so the wrapper runs no O(N) check
 *                             loops; on this code it runs the reference's checks, which throw
 *    4  STAN_CUDA_LOG_ZERO    (poisson_log_lpmf) some log rate is +inf, or -inf next to a non-zero count: the
 *                             reference returns LOG_ZERO before building any partial (poisson_log_lpmf.hpp:51-62);
 *                             *lp is -inf, no partial was written; the C++ wrapper returns LOG_ZERO
 *   <0  failure (bad argument, CUDA or NCCL error); stan_cuda_last_error() has the text;
 *       the C++ wrapper throws std::system_error (cf. stan/math/opencl/err/check_opencl.hpp)
 *
 * Threading: every function is re-entrant.  Each host thread gets its own stream, pinned staging
 * and device scratch per device; the data cache is guarded by a mutex.  Results are deterministic:
 * all reductions run in a fixed order, no floating-point atomics.
 */
#ifndef STAN_MATH_CUDA_C_API_H
#define STAN_MATH_CUDA_C_API_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define STAN_CUDA_API __attribute__((visibility("default")))
#else
#define STAN_CUDA_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define STAN_CUDA_OK 0
#define STAN_CUDA_NONFINITE 1
#define STAN_CUDA_BAD_OUTCOME 2
#define STAN_CUDA_BAD_OPERAND 3
#define STAN_CUDA_LOG_ZERO 4
#define STAN_CUDA_NOT_POS_DEF 5 /* cholesky_decompose: a pivot is not positive (matrix not positive definite, or NaN) */
#define STAN_CUDA_NOT_SYMMETRIC 6 /* cholesky_decompose: |a(m, n) - a(n, m)| <= tolerance fails for some pair */
#define STAN_CUDA_ERR_ARG (-1)
#define STAN_CUDA_ERR_CUDA (-2)
#define STAN_CUDA_ERR_NCCL (-3)
#define STAN_CUDA_ERR_STATE (-4)

/* GLM families */
#define STAN_CUDA_BERNOULLI_LOGIT 0
#define STAN_CUDA_POISSON_LOG 1
#define STAN_CUDA_NORMAL_ID 2
#define STAN_CUDA_NEG_BINOMIAL_2_LOG 3 /* precision phi travels in the sigma slot, d/dphi in dsigma */
#define STAN_CUDA_ORDERED_LOGISTIC 4   /* no intercept; cut-points in cuts[cuts_len], d/dcuts in dcuts; y in 1..cuts_len+1 */
#define STAN_CUDA_CATEGORICAL_LOGIT 5  /* n_classes columns: beta is k x n_classes (column-major), alpha[n_classes]; y in 1..n_classes */

/* flags: which summands of the log-density to include (the reference decides this at compile
 * time with include_summand<propto, ...>, stan/math/prim/scal/meta/include_summand.hpp) and
 * which partials the caller's operands_and_partials edges need. */
#define STAN_CUDA_LP_CONST 0x01u   /* data-only terms: -sum lgamma(y+1) | N*NEG_LOG_SQRT_TWO_PI   */
#define STAN_CUDA_LP_SCALE 0x02u   /* -sum log(sigma)  (normal_id_glm, normal_lpdf)                */
#define STAN_CUDA_LP_MAIN 0x04u    /* the parameter-dependent part                                */
#define STAN_CUDA_LP_ALL 0x07u
/* neg_binomial_2_log only: the sum(y * theta) term, which the reference ties to
 * include_summand<propto, T_x, T_alpha, T_beta> (neg_binomial_2_log_glm_lpmf.hpp:148-150) while
 * LP_MAIN is the -sum((y + phi) * logsumexp(theta, log phi)) term (:145-147) */
#define STAN_CUDA_LP_LINEAR 0x200u
#define STAN_CUDA_NEED_DBETA 0x08u
#define STAN_CUDA_NEED_DALPHA 0x10u /* GLM: d/dalpha ; normal_lpdf: d/dmu                          */
#define STAN_CUDA_NEED_DSIGMA 0x20u
#define STAN_CUDA_NEED_DCUTS 0x20u  /* ordered_logistic: d/dcuts (shares the bit of NEED_DSIGMA)  */
#define STAN_CUDA_NEED_DY 0x40u
#define STAN_CUDA_NEED_ALL 0x78u
#define STAN_CUDA_NEED_DX 0x400u    /* d/dx = d * beta^T (n x k), var x */
/* stan_cuda_glm_launch only: sum the result block over the ranks of the communicator INSIDE the
 * fused kernel (peer exchange over NVLink, see stan_cuda_comm_fused_available) instead of a
 * separate stan_cuda_allreduce.  Every rank must launch with the flag, in the same order. */
#define STAN_CUDA_FUSED_ALLREDUCE 0x100u

typedef struct stan_cuda_glm_args {
  int32_t family;      /* STAN_CUDA_BERNOULLI_LOGIT | POISSON_LOG | NORMAL_ID | NEG_BINOMIAL_2_LOG | ORDERED_LOGISTIC | CATEGORICAL_LOGIT */
  uint32_t flags;      /* STAN_CUDA_LP_* | STAN_CUDA_NEED_* */
  int64_t n;           /* rows of x (instances) */
  int64_t k;           /* columns of x (attributes) */
  const double* x;     /* host, borrowed; cached on the device the first time it is seen */
  int64_t ldx;         /* leading dimension in elements: >= n if column-major, >= k if row-major */
  int32_t x_row_major; /* 0 = column-major (Eigen::MatrixXd), 1 = row-major */
  int32_t y_is_double; /* 0: y is int32[n] (bernoulli, poisson); 1: y is double[n] */
  const void* y;       /* host, borrowed; cached with x */
  const double* alpha; /* host; alpha_len doubles (ordered_logistic: ignored, may be NULL) */
  int64_t alpha_len;   /* 1 (scalar intercept) or n */
  const double* beta;  /* host; k doubles */
  const double* sigma; /* host; normal_id: scale sigma; neg_binomial_2_log: precision phi */
  int64_t sigma_len;   /* 0 (bernoulli, poisson), 1 or n */
  /* outputs: host memory owned by the caller, NULL when the matching NEED_ flag is off */
  double* lp;          /* [1] */
  double* dbeta;       /* [k] */
  double* dalpha;      /* [alpha_len] */
  double* dsigma;      /* [sigma_len] */
  double* dy;          /* [n], normal_id only */
  double* aux;         /* optional [4]: sum d, sum s^2, sum log sigma, sum lgamma(y+1) */
  /* ordered_logistic only (stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:46-52) */
  const double* cuts;  /* host; cuts_len ascending cut-points (number of classes - 1) */
  int64_t cuts_len;
  double* dcuts;       /* [cuts_len] */
  /* categorical_logit only (stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:38-42): alpha has
   * n_classes entries (alpha_len = n_classes), beta is k x n_classes column-major, dalpha[n_classes],
   * dbeta[k * n_classes] */
  int64_t n_classes;
  /* 1: y is not immutable data -- a var outcome vector (normal_id_glm_lpdf with y a parameter), or a scalar
   * outcome the wrapper broadcasts into a temporary.  The cache is then keyed on x ALONE and y travels to the
   * device with every call (its range and sum lgamma(y + 1) are recomputed per call), so x is never refilled. */
  int32_t y_per_call;
  /* 1: x is not immutable data either (a var design matrix whose values the wrapper gathered into a temporary):
   * nothing is looked up by address or content; x (and y) are uploaded and re-tiled on every call into device
   * buffers the calling thread reuses from call to call. */
  int32_t x_per_call;
  /* var x (bernoulli_logit_glm_lpmf.hpp:125-128, poisson_log_glm_lpmf.hpp:119-122,
   * normal_id_glm_lpdf.hpp:121-124): with STAN_CUDA_NEED_DX the kernel also streams out
   * d/dx[i][j] = d_i * beta_j, column-major n x k with leading dimension lddx (>= n), host memory. */
  double* dx;
  int64_t lddx;
} stan_cuda_glm_args;

typedef struct stan_cuda_normal_args {
  uint32_t flags;      /* STAN_CUDA_LP_* | NEED_DALPHA (= d/dmu) | NEED_DSIGMA | NEED_DY */
  int32_t y_is_data;   /* 1: y is immutable data, cached on the device; 0: uploaded every call */
  int64_t n;           /* max_size(y, mu, sigma) */
  const double* y;
  int64_t y_len;       /* 1 or n */
  const double* mu;
  int64_t mu_len;      /* 1 or n */
  const double* sigma;
  int64_t sigma_len;   /* 1 or n */
  double* lp;          /* [1] */
  double* dmu;         /* [mu_len] */
  double* dsigma;      /* [sigma_len] */
  double* dy;          /* [y_len] */
} stan_cuda_normal_args;

/* Location-scale densities that share stan_cuda_normal_args: the density id selects the
 * per-element term of stan_cuda_locscale_lpdf_eval; stan_cuda_normal_lpdf_eval is id 0. */
#define STAN_CUDA_LOCSCALE_NORMAL 0    /* normal_lpdf(y | mu, sigma)     prim/scal/prob/normal_lpdf.hpp:38-113 */
#define STAN_CUDA_LOCSCALE_CAUCHY 1    /* cauchy_lpdf(y | mu, sigma)     prim/scal/prob/cauchy_lpdf.hpp:35-113 */
#define STAN_CUDA_LOCSCALE_LOGNORMAL 2 /* lognormal_lpdf(y | mu, sigma)  prim/scal/prob/lognormal_lpdf.hpp:19-133 */
/* two more densities of a real outcome on the same argument block (the names of its fields are normal_lpdf's):
 *   gamma_lpdf(y | alpha, beta): alpha travels in mu / dmu, beta in sigma / dsigma.  Summands (gamma_lpdf.hpp:118-129):
 *     LP_DOF = -lgamma(alpha), LP_SCALE = alpha * log(beta), LP_Y = (alpha - 1) * log(y), LP_MAIN = -beta * y.
 *   exponential_lpdf(y | beta): beta in sigma / dsigma, mu is ignored (pass one zero).  Summands
 *     (exponential_lpdf.hpp:76-81): LP_SCALE = log(beta), LP_MAIN = -beta * y. */
#define STAN_CUDA_LOCSCALE_GAMMA 3       /* prim/scal/prob/gamma_lpdf.hpp:41-145 */
#define STAN_CUDA_LOCSCALE_EXPONENTIAL 4 /* prim/scal/prob/exponential_lpdf.hpp:42-95 */
/*   beta_lpdf(y | alpha, beta): alpha in mu / dmu, beta in sigma / dsigma.  Summands (beta_lpdf.hpp:150-164):
 *     LP_SCALE = lgamma(alpha + beta), LP_DOF = -lgamma(alpha), LP_LINEAR = -lgamma(beta), LP_Y = (alpha - 1) * log(y),
 *     LP_MAIN = (beta - 1) * log1m(y).
 *   weibull_lpdf(y | alpha, sigma): alpha in mu / dmu.  Summands (weibull_lpdf.hpp:111-123): LP_DOF = log(alpha),
 *     LP_Y = (alpha - 1) * log(y), LP_SCALE = -alpha * log(sigma), LP_MAIN = -(y / sigma)^alpha. */
#define STAN_CUDA_LOCSCALE_BETA 5        /* prim/scal/prob/beta_lpdf.hpp:40-181 */
#define STAN_CUDA_LOCSCALE_WEIBULL 6     /* prim/scal/prob/weibull_lpdf.hpp:29-143 */
/*   inv_gamma_lpdf(y | alpha, beta): the slots and flags of gamma_lpdf.  Summands (inv_gamma_lpdf.hpp:115-126):
 *     LP_DOF = -lgamma(alpha), LP_SCALE = alpha * log(beta), LP_Y = -(alpha + 1) * log(y), LP_MAIN = -beta / y;
 *     LOG_ZERO for an outcome that is not positive (:58-63). */
#define STAN_CUDA_LOCSCALE_INV_GAMMA 7   /* prim/scal/prob/inv_gamma_lpdf.hpp:33-146 */
/* lognormal only: the -sum log(y) term, which the reference ties to include_summand<propto, T_y>
 * (lognormal_lpdf.hpp:113-115) */
#define STAN_CUDA_LP_Y 0x1000u

/* Vectorised elementwise densities with an integer outcome and one real operand: value + partials of the
 * operand in one pass (the pattern of stan_cuda_normal_lpdf_eval for the lpmf's whose
 * per-element terms are the epilogues of the GLM kernels). */
#define STAN_CUDA_DENSITY_BERNOULLI_LOGIT 1 /* bernoulli_logit_lpmf(n | theta)  prim/scal/prob/bernoulli_logit_lpmf.hpp:28-89 */
#define STAN_CUDA_DENSITY_POISSON_LOG 2     /* poisson_log_lpmf(n | alpha)      prim/scal/prob/poisson_log_lpmf.hpp:22-97 */
#define STAN_CUDA_DENSITY_POISSON 3         /* poisson_lpmf(n | lambda)         prim/scal/prob/poisson_lpmf.hpp:20-77 */
#define STAN_CUDA_DENSITY_BERNOULLI 4       /* bernoulli_lpmf(n | theta)        prim/scal/prob/bernoulli_lpmf.hpp:28-112 */

typedef struct stan_cuda_density_args {
  int32_t density;      /* STAN_CUDA_DENSITY_* */
  uint32_t flags;       /* STAN_CUDA_LP_CONST (poisson: -sum lgamma(n+1)) | LP_MAIN | NEED_DALPHA (= d/dtheta) */
  int64_t n;            /* max_size(y, theta) */
  const int32_t* y;     /* host; integer outcomes: immutable data, cached on the device when y_len == n */
  int64_t y_len;        /* 1 or n */
  const double* theta;  /* host; logit (bernoulli_logit), log rate (poisson_log), rate (poisson) or probability (bernoulli) */
  int64_t theta_len;    /* 1 or n */
  double* lp;           /* [1] */
  double* dtheta;       /* [theta_len]: per element, or the sum over n for a scalar operand */
} stan_cuda_density_args;

/* student_t_lpdf(y | nu, mu, sigma), vectorised over any mix of scalar / length-n operands: value and the
 * partials of all four operands in one pass (stan/math/prim/scal/prob/student_t_lpdf.hpp:55-201).
 * The summands the reference drops under propto (:117-146) are selected by flags:
 *   STAN_CUDA_LP_CONST  NEG_LOG_SQRT_PI per element                                              (:134-136)
 *   STAN_CUDA_LP_DOF    lgamma(nu/2 + 1/2) - lgamma(nu/2) - log(nu)/2                            (:137-140)
 *   STAN_CUDA_LP_SCALE  -log(sigma)                                                              (:141-143)
 *   STAN_CUDA_LP_MAIN   -(nu/2 + 1/2) * log1p(((y - mu) / sigma)^2 / nu)                         (:144-147)
 * Partials are written per element for a length-n operand and as the sum over n for a scalar one. */
#define STAN_CUDA_LP_DOF 0x800u
#define STAN_CUDA_NEED_DNU 0x08u /* shares the bit of NEED_DBETA */

typedef struct stan_cuda_student_t_args {
  uint32_t flags;      /* STAN_CUDA_LP_{CONST,DOF,SCALE,MAIN} | NEED_DY | NEED_DNU | NEED_DALPHA (= d/dmu) | NEED_DSIGMA */
  int32_t y_is_data;   /* 1: y is immutable data, cached on the device; 0: uploaded every call */
  int64_t n;           /* max_size(y, nu, mu, sigma) */
  const double* y;
  int64_t y_len;       /* 1 or n */
  const double* nu;
  int64_t nu_len;      /* 1 or n */
  const double* mu;
  int64_t mu_len;      /* 1 or n */
  const double* sigma;
  int64_t sigma_len;   /* 1 or n */
  double* lp;          /* [1] */
  double* dy;          /* [y_len] */
  double* dnu;         /* [nu_len] */
  double* dmu;         /* [mu_len] */
  double* dsigma;      /* [sigma_len] */
} stan_cuda_student_t_args;

/* ---- dense FP64 matrices on the device and their product ------------------------------
 * What the reference's OpenCL backend provides to its `multiply` hooks (stan/math/prim/mat/fun/multiply.hpp:71-79,
 * stan/math/rev/mat/fun/multiply.hpp:88-99, 111-123, 274-303, 444-473):
 *   stan_cuda_matrix            matrix_cl<double>            stan/math/opencl/matrix_cl.hpp:32-454
 *   stan_cuda_matrix_upload     to_matrix_cl / matrix_cl(A)  stan/math/opencl/copy.hpp:37-58
 *   stan_cuda_matrix_download   from_matrix_cl               stan/math/opencl/copy.hpp:69-93
 *   stan_cuda_matrix_multiply   A * B, transpose(A) * B ...  stan/math/opencl/multiply.hpp, transpose.hpp
 * Column-major like Eigen.  Every call returns with nothing in flight, so a handle may be used from any host
 * thread afterwards (not from two threads at once).  A matrix lives on the device of the calling thread's shard
 * scope (stan_cuda_shard_scope_enter), else on the first device. */
typedef struct stan_cuda_matrix stan_cuda_matrix;

/* C (m x n) = op(A) * op(B) with host matrices: upload, product, download in one call (the prim hook and each
 * product of the rev hook when nothing stays resident).  op(A) is m x k, op(B) is k x n; trans_* != 0 means the
 * host array holds the transpose (k x m resp. n x k, column-major). */
typedef struct stan_cuda_multiply_args {
  int64_t m, n, k;
  const double* a;
  int64_t lda;
  int32_t trans_a;
  int32_t trans_b;
  const double* b;
  int64_t ldb;
  double* c;
  int64_t ldc;
} stan_cuda_multiply_args;

typedef struct stan_cuda_stats_t {
  uint64_t cache_hits;
  uint64_t cache_misses;
  uint64_t bytes_uploaded;   /* cache-fill traffic, host -> device */
  uint64_t evals;            /* GLM + normal_lpdf evaluations */
  uint64_t kernel_launches;  /* launches of this library's own kernels */
  uint64_t allreduces;       /* NCCL all-reduce calls */
  double last_fill_ms;       /* wall time of the most recent cache fill */
  uint64_t fused_allreduces; /* all-reduces done inside the fused kernel over peer memory */
  uint64_t cache_evictions;  /* entries dropped to stay inside the cache budget */
} stan_cuda_stats_t;

/* ---- context -------------------------------------------------------------------------- */
/* Use the first n_devices visible devices (0 = all; env STAN_CUDA_DEVICES overrides 0).
 * Idempotent; implied by the first evaluation. */
STAN_CUDA_API int stan_cuda_init(int n_devices);
/* Bind this PROCESS to one device (one process per GPU: torchrun LOCAL_RANK). */
STAN_CUDA_API int stan_cuda_init_device(int device_ordinal);
STAN_CUDA_API int stan_cuda_device_count(void);
/* Human-readable description of the devices in use (cf. opencl_context.description()). */
STAN_CUDA_API const char* stan_cuda_describe(void);
STAN_CUDA_API const char* stan_cuda_last_error(void); /* thread-local */
STAN_CUDA_API int stan_cuda_stats(stan_cuda_stats_t* out);
STAN_CUDA_API void stan_cuda_stats_reset(void);
/* Rows are sharded over all devices of the process only when n*k reaches this many elements
 * (default 1<<24); smaller problems stay on one device. */
STAN_CUDA_API void stan_cuda_set_shard_min_elems(int64_t elems);
/* Kernel tuning knobs (cf. opencl_context_base::tuning_opts_, stan/math/opencl/opencl_context.hpp:164-183):
 * consumer threads per CTA (128|256|512; fixes the tile shape of data created afterwards), TMA
 * pipeline stages and resident CTAs per SM.  0 = automatic, negative = leave unchanged.
 * Environment: STAN_CUDA_THREADS / STAN_CUDA_STAGES / STAN_CUDA_CTAS_PER_SM. */
STAN_CUDA_API void stan_cuda_set_tuning(int threads, int stages, int ctas_per_sm);
/* Kernel variant for data created afterwards: 1 (default) = warp-synchronous kernel (csrc/glm_warp.cu), 0 = the
 * CTA-synchronous kernel it superseded (csrc/glm_fused.cu), present only in a library built with
 * STAN_CUDA_WITH_CTA_KERNEL=1 (A/B measurements); in the product build the call is accepted and ignored.
 * Environment: STAN_CUDA_KERNEL=warp|cta. */
STAN_CUDA_API void stan_cuda_set_kernel(int variant);
/* categorical_logit kernel for data created afterwards: 1 (default) = FP64 tensor-core kernel (mma.m8n8k4,
 * SASS DMMA) on a tile layout stored in the instruction's fragment order (csrc/glm_categorical_mma.cu;
 * k <= 256, wider x uses variant 0), 0 = vector FP64 kernel on the warp tile layout
 * (csrc/glm_categorical.cu).  Same results to rounding.  Environment: STAN_CUDA_CATEGORICAL=mma|vector. */
STAN_CUDA_API void stan_cuda_set_categorical_kernel(int variant);

/* ---- host-level evaluation (what the Stan wrappers call) -------------------------------- */
STAN_CUDA_API int stan_cuda_glm_eval(const stan_cuda_glm_args* a);
/* `count` independent GLMs of one family (bernoulli_logit | poisson_log | normal_id) and one k, scalar alpha /
 * sigma, in ONE kernel launch per device: item j runs on device j % device_count and its data is cached under
 * (call_id, j), exactly like the jobs of map_rect (stan/math/rev/mat/functor/map_rect_concurrent.hpp:39-54 and
 * stan_cuda_shard_scope_enter).  item_rc (optional, [count]) receives the per-item code; the return value is the
 * worst of them (or a negative error, after which no output is valid). */
STAN_CUDA_API int stan_cuda_glm_eval_batch(const stan_cuda_glm_args* items, int count, int call_id, int* item_rc);
STAN_CUDA_API int stan_cuda_normal_lpdf_eval(const stan_cuda_normal_args* a);
/* The same call for the other location-scale densities (density: STAN_CUDA_LOCSCALE_*).  Flags as for normal_lpdf:
 * LP_CONST = the constant (N * NEG_LOG_SQRT_TWO_PI | N * NEG_LOG_PI), LP_SCALE = -sum log(sigma), LP_MAIN = the
 * term every operand enters, LP_Y (lognormal) = -sum log(y).  Returns STAN_CUDA_BAD_OPERAND where the reference's
 * eager checks throw (cauchy_lpdf.hpp:47-49: y NaN, mu not finite, sigma not positive and finite; lognormal_lpdf.hpp:24-27:
 * also y negative; gamma_lpdf.hpp:55-57: y NaN, alpha or beta not positive and finite; exponential_lpdf.hpp:55-56: y NaN or
 * negative, beta not positive and finite) and STAN_CUDA_LOG_ZERO for lognormal_lpdf with an outcome equal to zero (:41-45)
 * and gamma_lpdf / weibull_lpdf with a negative outcome (gamma_lpdf.hpp:69-74, weibull_lpdf.hpp:56-61); beta_lpdf: either shape
 * not positive and finite, y NaN or outside [0, 1] (beta_lpdf.hpp:48-54) and weibull_lpdf: y not finite, alpha or sigma not
 * positive and finite (weibull_lpdf.hpp:38-40) are STAN_CUDA_BAD_OPERAND. */
STAN_CUDA_API int stan_cuda_locscale_lpdf_eval(int32_t density, const stan_cuda_normal_args* a);
/* Returns STAN_CUDA_BAD_OPERAND when an outcome is outside the density's support (bernoulli_logit: not in
 * {0, 1}, bernoulli_logit_lpmf.hpp:42; poisson_log: negative, poisson_log_lpmf.hpp:38) or the operand is NaN
 * (:43 / :39; poisson: NaN or negative rate, poisson_lpmf.hpp:32-33; bernoulli: probability not finite or outside
 * [0, 1], bernoulli_lpmf.hpp:42-43) -- every element is tested inside the kernel, the wrapper then runs the reference's own checks
 * to throw the identical exception -- and STAN_CUDA_LOG_ZERO as described above. */
STAN_CUDA_API int stan_cuda_density_eval(const stan_cuda_density_args* a);
/* Returns STAN_CUDA_BAD_OPERAND when y is NaN, nu or sigma is not positive and finite, or mu is not finite
 * (student_t_lpdf.hpp:67-70; tested per element inside the kernel, the wrapper then runs the reference's own
 * checks to throw the identical exception). */
STAN_CUDA_API int stan_cuda_student_t_lpdf_eval(const stan_cuda_student_t_args* a);
/* dense matrices (see stan_cuda_matrix above).  create: rows x cols of zeros. */
STAN_CUDA_API int stan_cuda_matrix_create(int64_t rows, int64_t cols, stan_cuda_matrix** out);
STAN_CUDA_API void stan_cuda_matrix_free(stan_cuda_matrix* m);
STAN_CUDA_API int64_t stan_cuda_matrix_rows(const stan_cuda_matrix* m);
STAN_CUDA_API int64_t stan_cuda_matrix_cols(const stan_cuda_matrix* m);
STAN_CUDA_API int stan_cuda_matrix_upload(stan_cuda_matrix* m, const double* host, int64_t ld);
STAN_CUDA_API int stan_cuda_matrix_download(const stan_cuda_matrix* m, double* host, int64_t ld);
/* c = op(a) * op(b); c must have been created with the shape of the product and must not be an operand */
STAN_CUDA_API int stan_cuda_matrix_multiply(const stan_cuda_matrix* a, int trans_a, const stan_cuda_matrix* b, int trans_b,
                                            stan_cuda_matrix* c);
STAN_CUDA_API int stan_cuda_multiply(const stan_cuda_multiply_args* a);
/* Cholesky factor and its reverse-mode adjoint (the cholesky_decompose hooks of
 * stan/math/prim/mat/fun/cholesky_decompose.hpp:34-62 and stan/math/rev/mat/fun/cholesky_decompose.hpp:255-440, for
 * which the OpenCL backend has opencl::cholesky_decompose, stan/math/opencl/cholesky_decompose.hpp:36-107, and
 * cholesky_opencl::chain()).  Host matrices, column-major.
 * decompose: l (n x n) = lower Cholesky factor of a (its lower triangle is read), zeros above the diagonal.  Returns
 *   STAN_CUDA_NOT_POS_DEF when a pivot is not positive -- the wrapper then runs the reference's LLT +
 *   check_pos_definite on the host to throw the identical exception.  symmetry_tolerance >= 0: the reference's
 *   check_symmetric (prim/mat/err/check_symmetric.hpp:39-46, |a(m, n) - a(n, m)| <= tolerance, NaN fails) runs on
 *   the uploaded matrix instead of as a strided O(n^2) host loop (30 ms at n = 2048, four times the factorisation);
 *   STAN_CUDA_NOT_SYMMETRIC, reported before NOT_POS_DEF as the reference orders its checks, makes the wrapper run
 *   the reference's check for the message.  Negative: no symmetry check.
 * adjoint: given l and the adjoints l_adj of its lower triangle, writes the adjoints of the lower triangle of a into
 *   a_adj (n x n, zeros above the diagonal: only the lower triangle of a is an operand of the reference's node). */
STAN_CUDA_API int stan_cuda_cholesky_decompose(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl,
                                               double symmetry_tolerance);
/* Squared-exponential covariance (the gp_exp_quad_cov hooks of
 * stan/math/prim/mat/fun/gp_exp_quad_cov.hpp:344-655, OpenCL kernels stan/math/opencl/kernels/gp_exp_quad_cov.hpp):
 *   cov(i, j) = sigma_sq * exp(factor * sum_d (x1[d][i] - x2[d][j])^2),   cov n1 x n2 column-major, host memory.
 * Points are stored dimension-major: coordinate d of point i at x[d * n + i] (a std::vector<double> of scalars is
 * dim = 1).  x2 == NULL: the one-sample form, cov is n1 x n1 with sigma_sq exactly on its diagonal (:53, :88).
 * factor = -0.5 / l^2 for a scalar length scale; -0.5 when the caller has divided the coordinates by a vector of
 * length scales (divide_columns, :233, :338).  The caller checks the operands (NaN, positivity) as the reference does. */
STAN_CUDA_API int stan_cuda_gp_exp_quad_cov(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2,
                                            double sigma_sq, double factor, double* cov, int64_t ldc);
/* Triangular solve x (n x m) = op(A)^-1 b (the mdivide_left_tri hooks of
 * stan/math/prim/mat/fun/mdivide_left_tri.hpp:58-100 and stan/math/rev/mat/fun/mdivide_left_tri.hpp:69-82, 91-105,
 * 145-159, 170-183, for which the OpenCL backend has tri_inverse + multiply, stan/math/opencl/tri_inverse.hpp).
 * A is n x n, triangular: lower != 0 reads its lower triangle, else its upper; op(A) = A' when trans != 0 (the
 * reverse sweep's A^-T adj).  Host matrices, column-major; x may be b. */
STAN_CUDA_API int stan_cuda_tri_solve(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b,
                                      int64_t ldb, double* x, int64_t ldx);
STAN_CUDA_API int stan_cuda_cholesky_adjoint(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj,
                                             double* a_adj, int64_t lda);
/* Page-locked staging memory of the calling thread (at least `bytes`, valid until the thread's next
 * call with a larger size).  Operands gathered into it upload by DMA without a bounce buffer. */
STAN_CUDA_API void* stan_cuda_host_staging(size_t bytes);

/* ---- data cache --------------------------------------------------------------------------- */
STAN_CUDA_API int stan_cuda_cache_invalidate(const void* host_ptr); /* drop entries keyed on this x (or y) pointer */
STAN_CUDA_API void stan_cuda_cache_clear(void);
STAN_CUDA_API int64_t stan_cuda_cache_bytes(void);                  /* device bytes held by the cache */
/* Byte budget of the cache (all devices together): when a new data set does not fit, least-recently-used
 * entries are dropped first (an evaluation that is still using one keeps it alive until it returns).
 * 0 = automatic: 80 % of the smallest device's memory, or environment STAN_CUDA_CACHE_BYTES.
 * (matrix_cl has no counterpart: its buffers live as long as the C++ object, stan/math/opencl/matrix_cl.hpp:32-454.) */
STAN_CUDA_API void stan_cuda_cache_set_limit(int64_t bytes);
/* Cache identity.  Outside a shard scope an entry is found by (x pointer, y pointer, shape, layout) and is
 * trusted only if a content fingerprint of 64 sampled elements of x and of y still matches; inside a scope by
 * (call_id, job, ordinal of the GLM call inside the job, shape, family) plus a 16-sample tripwire.  A host
 * array that is REWRITTEN IN PLACE at positions the samples miss is not noticed: data handed to these entry
 * points is immutable by contract (a Stan program's data block; x_r / x_i of map_rect,
 * stan/math/prim/mat/functor/map_rect.hpp:97-100) -- call stan_cuda_cache_invalidate after mutating it.
 * verify = 1 (or environment STAN_CUDA_CACHE_VERIFY=1) replaces the samples by a hash of EVERY element of x
 * and y on every call: O(n k) host work per evaluation, for debugging and for the test-suite. */
STAN_CUDA_API void stan_cuda_cache_set_verify(int verify);

/* ---- map_rect shard scope (thread-local) ---------------------------------------------------- */
/* Inside a scope, evaluations of the calling thread run on `device` and their data is keyed by
 * (call_id, job) instead of by host pointer: the functor rebuilds X from x_r on every call, so
 * its address is not an identity; the immutability contract is map_rect's own
 * (stan/math/prim/mat/functor/map_rect.hpp:97-100).  Scopes nest (up to 8 deep): exit restores the scope that was
 * current at the matching enter. */
STAN_CUDA_API void stan_cuda_shard_scope_enter(int call_id, int job, int device);
STAN_CUDA_API void stan_cuda_shard_scope_exit(void);

/* ---- multi-process communicator (one process per GPU) --------------------------------------- */
/* Rank 0 creates the id; the host plumbing (torch.distributed, MPI, a file) carries the 128 bytes
 * to every rank; every rank then calls comm_init_rank.  After that, sharded evaluations
 * all-reduce [lp-parts, dbeta] across ranks: inside the fused kernel over peer memory when
 * available (below), else with ncclAllReduce on the evaluation stream. */
STAN_CUDA_API int stan_cuda_comm_unique_id(void* id128);
STAN_CUDA_API int stan_cuda_comm_init_rank(int world_size, int rank, const void* id128);
STAN_CUDA_API int stan_cuda_comm_world_size(void);
/* 1 when comm_init_rank could map every rank's exchange block into every other rank (CUDA IPC over
 * NVLink/NVSwitch peer access): the all-reduce then runs inside the fused kernel -- the last CTA pushes
 * its K+4 doubles into every peer's block, waits for the peers' flags and adds the blocks in rank
 * order -- and no ncclAllReduce is launched.  0: ncclAllReduce behind the kernel (same results up to
 * summation order).  Environment STAN_CUDA_FUSED_ALLREDUCE=0 forces 0. */
STAN_CUDA_API int stan_cuda_comm_fused_available(void);
STAN_CUDA_API const char* stan_cuda_comm_describe(void);
STAN_CUDA_API void stan_cuda_comm_destroy(void);

/* ---- device-level API (streams and buffers owned by the caller, e.g. torch) ---------------- */
typedef struct stan_cuda_data* stan_cuda_data_t;

/* Upload + re-layout of (x, y) of one GLM onto the current device: explicit cache fill. */
STAN_CUDA_API int stan_cuda_data_create(int family, int64_t n, int64_t k, const double* x, int64_t ldx,
                          int x_row_major, const void* y, int y_is_double, stan_cuda_data_t* out);
STAN_CUDA_API void stan_cuda_data_destroy(stan_cuda_data_t d);
/* Number of doubles of the parameter block / result block of a launch. */
STAN_CUDA_API int64_t stan_cuda_data_params_len(stan_cuda_data_t d); /* = k_pad + 2: [beta(k_pad), alpha, sigma] */
STAN_CUDA_API int64_t stan_cuda_data_result_len(stan_cuda_data_t d); /* = k_pad + 4: [lp, sum_d, sum_s2, status, dbeta(k_pad)] */
STAN_CUDA_API int64_t stan_cuda_data_kpad(stan_cuda_data_t d);
STAN_CUDA_API int64_t stan_cuda_data_algorithmic_bytes(stan_cuda_data_t d); /* n*k*8 + n*sizeof(y) */
/* Smallest / largest integer outcome seen at fill (0, 0 for real outcomes). */
STAN_CUDA_API int stan_cuda_data_y_range(stan_cuda_data_t d, int64_t* y_min, int64_t* y_max);
/* Enqueue one fused evaluation on `stream` (a cudaStream_t).  d_params / d_result are device
 * pointers (layouts above); d_alpha_vec, d_sigma_vec, d_row_out may be NULL (scalar alpha/sigma,
 * no per-row outputs).  No host synchronisation, no copies. */
STAN_CUDA_API int stan_cuda_glm_launch(stan_cuda_data_t d, uint32_t flags, const double* d_params,
                         const double* d_alpha_vec, const double* d_sigma_vec, double* d_result,
                         double* d_row_dalpha, double* d_row_dsigma, void* stream);
/* ordered_logistic: d_params = [beta(k_pad), 0, 0, cuts(ncuts)], d_result = [lp, sum_d, 0, status,
 * dbeta(k_pad), dcuts(ncuts)]; the caller has checked 1 <= y <= ncuts + 1 (stan_cuda_data_y_range). */
STAN_CUDA_API int stan_cuda_glm_launch_ordered(stan_cuda_data_t d, uint32_t flags, const double* d_params, int ncuts,
                                 double* d_result, void* stream);
/* categorical_logit: d_params = [beta as n_classes rows of k_pad (class-major, zero padded), alpha(n_classes)],
 * d_result = [lp, 0, 0, status, dalpha(n_classes), dbeta as n_classes rows of k_pad]. */
STAN_CUDA_API int stan_cuda_glm_launch_categorical(stan_cuda_data_t d, uint32_t flags, const double* d_params,
                                     int n_classes, double* d_result, void* stream);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same evaluation and
 * the same argument meaning as stan_cuda_glm_launch for bernoulli_logit / poisson_log / normal_id data with a scalar
 * intercept / scale (no per-row operands, no collective), but the call allocates NOTHING: the reduction scratch of
 * the kernel (per-CTA partials, group sums, ticket words) lives in a file-scope __device__ array of the library,
 * 320 x 520 doubles = 1.27 MiB (+ 256 B of tickets) that every GPU which loads the library holds until exit.
 * Limit: (CTAs + reduction groups) x (k_pad + 4) + 2 <= 166400 doubles, i.e. every k <= 512 (one warp tile; wider
 * matrices run as several launches and are not served here) on a GPU of up to 160 SMs; beyond it the call returns STAN_CUDA_ERR_ARG with the two numbers in stan_cuda_last_error().
 * Every call on a device shares that one array: calls that OVERLAP on a GPU (two streams, two host threads) race
 * on it -- the caller serialises them.  The data handle, d_params and d_result are the caller's (created before
 * the no-allocation region); the first call per device resolves the array's device address (cudaGetSymbolAddress),
 * so make one call before starting a graph capture. */
STAN_CUDA_API int stan_cuda_glm_launch_noalloc(stan_cuda_data_t d, uint32_t flags, const double* d_params,
                                               double* d_result, void* stream);
/* All-reduce (sum) `count` doubles in place across the ranks of the communicator, on `stream`. */
STAN_CUDA_API int stan_cuda_allreduce(double* d_buf, int64_t count, void* stream);
/* Number of this library's kernels launched per stan_cuda_glm_launch call for this data. */
STAN_CUDA_API int stan_cuda_data_launches_per_eval(stan_cuda_data_t d);

#ifdef __cplusplus
}
#endif
#endif /* STAN_MATH_CUDA_C_API_H */
