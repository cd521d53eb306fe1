/* tests/mock_device/mock_stan_cuda.c -- TEST DOUBLE of libstanmath_cuda.so.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The CPU suite (`pytest -m "not gpu"`) has no GPU, and the product library has no CPU path: every
 * stan_cuda_*_eval fails loudly without a device.  To exercise the C++ HOST LAYER on CPU all the same -- the
 * include-path overlay and its routing, argument marshalling (stan/math/cuda/marshal.hpp), the mapping of ABI
 * return codes onto the reference's exceptions, the assembly of the tape node from the returned partials,
 * map_rect's shard scopes -- tests/test_cxx_mock_cpu.py runs the C++ test binaries (the reference's own unit
 * tests compiled against the overlay, test/unit/math/cuda) with THIS library in front of them on
 * LD_LIBRARY_PATH.  It implements the entry points those binaries import by forwarding the argument structs to
 * the plain-C oracle (oracle/glm_oracle.c, whose functions take the very same structs) and adds the return-code
 * contract of include/stan_cuda.h that lives in the product's host code rather than in the oracle
 * (STAN_CUDA_BAD_OUTCOME from the outcome range, STAN_CUDA_BAD_OPERAND for normal_lpdf).
 *
 * It is built into build/mock/ by the test that uses it, is never installed, never on the product's library
 * path, never loaded by a `-m gpu` test, by smoke() or by bench.py.  What it proves is the host layer, not the
 * kernels: the GPU parity tests run the same binaries against the real library.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/stan_cuda.h"

int oracle_glm_eval(const stan_cuda_glm_args* a);
int oracle_normal_lpdf_eval(const stan_cuda_normal_args* a);
int oracle_student_t_lpdf_eval(const stan_cuda_student_t_args* a);
int oracle_multiply(const stan_cuda_multiply_args* g);
int oracle_tri_solve(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b, int64_t ldb, double* x, int64_t ldx);
int oracle_gp_exp_quad_cov(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2, double sigma_sq, double factor, double* cov, int64_t ldc);
int oracle_cholesky_decompose(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl);
int oracle_cholesky_adjoint(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj, double* a_adj, int64_t lda);
int oracle_density_eval(const stan_cuda_density_args* a);

static __thread char tl_error[256];
static __thread int tl_scope_active, tl_scope_call, tl_scope_job, tl_scope_device;
static unsigned long long g_evals, g_scoped_evals, g_invalidations;

/* STAN_CUDA_MOCK_TIMING=1: every 10th evaluation prints the cumulative wall time spent INSIDE the three eval entry
 * points, so that a bench binary run against the mock shows the host layer's own cost per evaluation (its total
 * minus this).  (Not at exit: bench/stan_bench.cpp leaves through _Exit.) */
static unsigned long long g_eval_ns;
static int g_timing = -1;
static double now_ns(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return 1e9 * (double)t.tv_sec + (double)t.tv_nsec;
}
static void timed_done(double t0) {
  const unsigned long long ns = __atomic_add_fetch(&g_eval_ns, (unsigned long long)(now_ns() - t0), __ATOMIC_RELAXED);
  if (g_timing < 0) {
    const char* v = getenv("STAN_CUDA_MOCK_TIMING");
    g_timing = v && *v == '1';
  }
  if (g_timing && g_evals % 10 == 0)
    fprintf(stderr, "mock device: %llu evals, %.3f ms inside the eval entry points\n", g_evals, (double)ns * 1e-6);
}
#define TIMED(call)           \
  const double t0_ = now_ns(); \
  const int rc = (call);      \
  timed_done(t0_)

static int fail(int rc, const char* what) {
  snprintf(tl_error, sizeof tl_error, "mock device: %s", what);
  return rc;
}

/* mock-only introspection (found with dlsym by test/unit/math/cuda/map_rect_dispatch_mock_test.cpp; the product
 * library has no such symbol): the (call_id, job, device) of every evaluation that ran inside a shard scope */
#define MOCK_LOG_CAP 4096
static int g_log[MOCK_LOG_CAP][3];
static unsigned g_log_n;
static void log_scope(void) {
  if (!tl_scope_active) return;
  const unsigned i = __atomic_fetch_add(&g_log_n, 1u, __ATOMIC_RELAXED);
  if (i < MOCK_LOG_CAP) {
    g_log[i][0] = tl_scope_call;
    g_log[i][1] = tl_scope_job;
    g_log[i][2] = tl_scope_device;
  }
}
int stan_cuda_mock_scope_log(int* out, int cap) { /* out[3 * i + {0,1,2}]; returns the number of entries; resets the log */
  const unsigned n = g_log_n < MOCK_LOG_CAP ? g_log_n : MOCK_LOG_CAP;
  for (unsigned i = 0; i < n && (int)i < cap; ++i) memcpy(out + 3 * i, g_log[i], sizeof g_log[i]);
  g_log_n = 0;
  return (int)n;
}
int stan_cuda_mock_scope_active(void) { return tl_scope_active; }

int stan_cuda_init(int n_devices) { (void)n_devices; return STAN_CUDA_OK; }
int stan_cuda_device_count(void) {
  const char* v = getenv("STAN_CUDA_MOCK_DEVICES");   /* lets a test see map_rect's job -> device dispatch */
  return v && atoi(v) > 0 ? atoi(v) : 1;
}
const char* stan_cuda_describe(void) { return "mock device (tests/mock_device): C ABI forwarded to the plain-C oracle"; }
const char* stan_cuda_last_error(void) { return tl_error; }

int stan_cuda_stats(stan_cuda_stats_t* out) {
  if (!out) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_stats: null");
  memset(out, 0, sizeof *out);
  out->evals = g_evals;
  out->cache_hits = g_scoped_evals;        /* the mock's meaning: evaluations that ran inside a shard scope */
  out->cache_misses = g_invalidations;     /* the mock's meaning: stan_cuda_cache_invalidate calls */
  return STAN_CUDA_OK;
}

/* include/stan_cuda.h, code 2: the product answers the eager outcome check from the range computed at cache fill */
static int outcome_out_of_range(const stan_cuda_glm_args* a) {
  if (a->family == STAN_CUDA_NORMAL_ID || a->y_is_double) return 0;
  const int32_t* y = (const int32_t*)a->y;
  for (int64_t i = 0; i < a->n; ++i) {
    const int64_t v = y[i];
    switch (a->family) {
      case STAN_CUDA_BERNOULLI_LOGIT: if (v < 0 || v > 1) return 1; break;
      case STAN_CUDA_ORDERED_LOGISTIC: if (v < 1 || v > a->cuts_len + 1) return 1; break;
      case STAN_CUDA_CATEGORICAL_LOGIT: if (v < 1 || v > a->n_classes) return 1; break;
      default: if (v < 0) return 1; break;
    }
  }
  return 0;
}

/* STAN_CUDA_MOCK_NO_COMPUTE=1 (timing of the host layer alone, bench binaries only): the GLM entry point returns at
 * once with zeroed outputs, so that what a bench binary measures is the C++ layer in front of the ABI */
static int no_compute(void) {
  static int on = -1;
  if (on < 0) {
    const char* v = getenv("STAN_CUDA_MOCK_NO_COMPUTE");
    on = v && *v == '1';
  }
  return on;
}

int stan_cuda_glm_eval(const stan_cuda_glm_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null args");
  if (no_compute()) {
    if (a->lp) *a->lp = -1.0;
    if (a->dbeta) memset(a->dbeta, 0, (size_t)(a->k * (a->n_classes > 0 ? a->n_classes : 1)) * sizeof(double));
    if (a->dalpha && a->alpha_len == 1) a->dalpha[0] = 0.0;
    if (a->dsigma && a->sigma_len == 1) a->dsigma[0] = 0.0;
    if (a->aux) memset(a->aux, 0, 4 * sizeof(double));
    return STAN_CUDA_OK;
  }
  if (a->n > 0 && (!a->y || (!a->x && a->k > 0))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null x or y");
  if (outcome_out_of_range(a)) return STAN_CUDA_BAD_OUTCOME;
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  if (tl_scope_active) __atomic_fetch_add(&g_scoped_evals, 1, __ATOMIC_RELAXED);
  log_scope();
  TIMED(oracle_glm_eval(a));
  return rc < 0 ? fail(rc, "stan_cuda_glm_eval: bad argument") : rc;
}

int stan_cuda_normal_lpdf_eval(const stan_cuda_normal_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_normal_lpdf_eval: null args");
  /* include/stan_cuda.h, code 3: y NaN, mu not finite, sigma not > 0 (normal_lpdf.hpp:51-53) */
  for (int64_t i = 0; i < a->y_len; ++i) if (isnan(a->y[i])) return STAN_CUDA_BAD_OPERAND;
  for (int64_t i = 0; i < a->mu_len; ++i) if (!isfinite(a->mu[i])) return STAN_CUDA_BAD_OPERAND;
  for (int64_t i = 0; i < a->sigma_len; ++i) if (!(a->sigma[i] > 0.0)) return STAN_CUDA_BAD_OPERAND;
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  TIMED(oracle_normal_lpdf_eval(a));
  return rc < 0 ? fail(rc, "stan_cuda_normal_lpdf_eval: bad argument") : rc;
}

int oracle_locscale_lpdf_eval(int density, const stan_cuda_normal_args* a);
int stan_cuda_locscale_lpdf_eval(int32_t density, const stan_cuda_normal_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_locscale_lpdf_eval: null args");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  TIMED(oracle_locscale_lpdf_eval(density, a)); /* code 3 / LOG_ZERO as the header describes */
  return rc < 0 ? fail(rc, "stan_cuda_locscale_lpdf_eval: bad argument") : rc;
}

int stan_cuda_student_t_lpdf_eval(const stan_cuda_student_t_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_student_t_lpdf_eval: null args");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  TIMED(oracle_student_t_lpdf_eval(a)); /* code 3 for the operands student_t_lpdf.hpp:67-70 rejects */
  return rc < 0 ? fail(rc, "stan_cuda_student_t_lpdf_eval: bad argument") : rc;
}

int stan_cuda_density_eval(const stan_cuda_density_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: null args");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  TIMED(oracle_density_eval(a));
  return rc < 0 ? fail(rc, "stan_cuda_density_eval: bad argument") : rc;
}

/* dense matrices: a handle is a host array here (column-major, ld = rows) */
struct stan_cuda_matrix {
  int64_t rows, cols;
  double* d;
};

int stan_cuda_matrix_create(int64_t rows, int64_t cols, stan_cuda_matrix** out) {
  if (!out || rows < 0 || cols < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_create: bad argument");
  stan_cuda_matrix* m = (stan_cuda_matrix*)malloc(sizeof(stan_cuda_matrix));
  if (!m) return fail(STAN_CUDA_ERR_CUDA, "stan_cuda_matrix_create: out of memory");
  m->rows = rows;
  m->cols = cols;
  m->d = (double*)calloc((size_t)(rows * cols > 0 ? rows * cols : 1), sizeof(double));
  if (!m->d) {
    free(m);
    return fail(STAN_CUDA_ERR_CUDA, "stan_cuda_matrix_create: out of memory");
  }
  *out = m;
  return STAN_CUDA_OK;
}

void stan_cuda_matrix_free(stan_cuda_matrix* m) {
  if (m) {
    free(m->d);
    free(m);
  }
}

int64_t stan_cuda_matrix_rows(const stan_cuda_matrix* m) { return m ? m->rows : 0; }
int64_t stan_cuda_matrix_cols(const stan_cuda_matrix* m) { return m ? m->cols : 0; }

int stan_cuda_matrix_upload(stan_cuda_matrix* m, const double* host, int64_t ld) {
  if (!m || (!host && m->rows * m->cols > 0) || ld < m->rows) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_upload: bad argument");
  for (int64_t j = 0; j < m->cols; ++j) memcpy(m->d + j * m->rows, host + j * ld, (size_t)m->rows * sizeof(double));
  return STAN_CUDA_OK;
}

int stan_cuda_matrix_download(const stan_cuda_matrix* m, double* host, int64_t ld) {
  if (!m || (!host && m->rows * m->cols > 0) || ld < m->rows) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_download: bad argument");
  for (int64_t j = 0; j < m->cols; ++j) memcpy(host + j * ld, m->d + j * m->rows, (size_t)m->rows * sizeof(double));
  return STAN_CUDA_OK;
}

int stan_cuda_matrix_multiply(const stan_cuda_matrix* a, int trans_a, const stan_cuda_matrix* b, int trans_b, stan_cuda_matrix* c) {
  if (!a || !b || !c || c == a || c == b) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: bad argument");
  stan_cuda_multiply_args g;
  memset(&g, 0, sizeof(g));
  g.m = trans_a ? a->cols : a->rows;
  g.k = trans_a ? a->rows : a->cols;
  g.n = trans_b ? b->rows : b->cols;
  if ((trans_b ? b->cols : b->rows) != g.k) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: inner dimensions differ");
  if (c->rows != g.m || c->cols != g.n) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: result has the wrong shape");
  g.a = a->d; g.lda = a->rows > 0 ? a->rows : 1; g.trans_a = trans_a;
  g.b = b->d; g.ldb = b->rows > 0 ? b->rows : 1; g.trans_b = trans_b;
  g.c = c->d; g.ldc = c->rows > 0 ? c->rows : 1;
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  return oracle_multiply(&g);
}

int stan_cuda_multiply(const stan_cuda_multiply_args* g) {
  if (!g) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_multiply: null args");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  TIMED(oracle_multiply(g));
  return rc < 0 ? fail(rc, "stan_cuda_multiply: bad argument") : rc;
}

int stan_cuda_tri_solve(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b, int64_t ldb,
                        double* x, int64_t ldx) {
  if (n < 0 || m < 0 || (n > 0 && !a) || (n > 0 && m > 0 && (!b || !x))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_tri_solve: bad argument");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  if (x == b) { /* the oracle copies b into x first: aliasing is fine */ }
  return oracle_tri_solve(lower, trans, n, m, a, lda, b, ldb, x, ldx);
}

int stan_cuda_gp_exp_quad_cov(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2, double sigma_sq, double factor,
                              double* cov, int64_t ldc) {
  if (n1 < 0 || n2 < 0 || dim < 1 || (n1 > 0 && !x1)) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_gp_exp_quad_cov: bad argument");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  return oracle_gp_exp_quad_cov(n1, n2, dim, x1, x2, sigma_sq, factor, cov, ldc);
}

int stan_cuda_cholesky_decompose(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl, double symmetry_tolerance) {
  if (n < 0 || (n > 0 && (!a || !l || lda < n || ldl < n))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_cholesky_decompose: bad argument");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  if (symmetry_tolerance >= 0.0) /* check_symmetric.hpp:39-46 */
    for (int64_t m = 0; m < n; ++m)
      for (int64_t k = m + 1; k < n; ++k)
        if (!(fabs(a[k * lda + m] - a[m * lda + k]) <= symmetry_tolerance)) return STAN_CUDA_NOT_SYMMETRIC;
  return oracle_cholesky_decompose(n, a, lda, l, ldl);
}

int stan_cuda_cholesky_adjoint(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj, double* a_adj, int64_t lda) {
  if (n < 0 || (n > 0 && (!l || !l_adj || !a_adj))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_cholesky_adjoint: bad argument");
  __atomic_fetch_add(&g_evals, 1, __ATOMIC_RELAXED);
  return oracle_cholesky_adjoint(n, l, ldl, l_adj, ldadj, a_adj, lda);
}

/* the product returns pinned memory of the calling thread; pageable memory serves a host-only run */
void* stan_cuda_host_staging(size_t bytes) {
  static __thread void* p;
  static __thread size_t cap;
  if (bytes > cap) {
    free(p);
    cap = bytes + bytes / 4 + 4096;
    p = malloc(cap);
    if (!p) cap = 0;
  }
  return p;
}

int stan_cuda_cache_invalidate(const void* host_ptr) {
  (void)host_ptr;
  __atomic_fetch_add(&g_invalidations, 1, __ATOMIC_RELAXED);
  return 0;
}
void stan_cuda_cache_clear(void) {}
int64_t stan_cuda_cache_bytes(void) { return 0; }
void stan_cuda_cache_set_verify(int verify) { (void)verify; }

/* scopes nest, as in the product (api.cu): exit restores the scope current at the matching enter */
static __thread int tl_saved[8][4];
static __thread int tl_depth;
void stan_cuda_shard_scope_enter(int call_id, int job, int device) {
  if (tl_depth < 8) {
    tl_saved[tl_depth][0] = tl_scope_active;
    tl_saved[tl_depth][1] = tl_scope_call;
    tl_saved[tl_depth][2] = tl_scope_job;
    tl_saved[tl_depth][3] = tl_scope_device;
  }
  ++tl_depth;
  tl_scope_active = 1;
  tl_scope_call = call_id;
  tl_scope_job = job;
  tl_scope_device = device;
}
void stan_cuda_shard_scope_exit(void) {
  if (tl_depth > 0) --tl_depth;
  if (tl_depth < 8) {
    tl_scope_active = tl_saved[tl_depth][0];
    tl_scope_call = tl_saved[tl_depth][1];
    tl_scope_job = tl_saved[tl_depth][2];
    tl_scope_device = tl_saved[tl_depth][3];
  }
  if (tl_depth == 0) tl_scope_active = 0;
}
