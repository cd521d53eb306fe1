// common.cuh -- shared declarations of libstanmath_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/stan_cuda.h"

#include <nvtx3/nvToolsExt.h>  // header-only (NVTX v3): ranges are no-ops unless a profiler is attached

namespace stancuda {

// NVTX range for the duration of a scope (upload / evaluation / exchange ranges for nsys / ncu timelines)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

// ---- error plumbing: no exception leaves the library, text goes to a thread-local string ----
void set_last_error(const std::string& s);
int fail(int code, const std::string& what);
int fail_cuda(cudaError_t e, const char* expr, const char* file, int line);

#define SC_CUDA(expr)                                                              \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) return ::stancuda::fail_cuda(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define SC_TRY(expr)            \
  do {                          \
    int _rc = (expr);           \
    if (_rc < 0) return _rc;    \
  } while (0)

// ---- tiling of one GLM data block (see DESIGN.md "Data layout in HBM") -----------------------
// X is stored tile by tile; a tile holds R consecutive rows and all kpad columns, column-major
// inside the tile:  x_dev[(tile * kpad + col) * R + r].  One tile is one contiguous span, so a
// single 1-D TMA bulk copy (cp.async.bulk) stages it.  Thread (cg, r) of the fused kernel owns
// row r and the kc columns [cg*kc, (cg+1)*kc);  C * R == threads,  kpad == C * kc,  kc <= 16.
constexpr int kMaxKc = 16;

// Two tile layouts, one per kernel variant (STAN_CUDA_KERNEL=cta|warp, default warp):
//   kLayoutCta  (glm_fused.cu) x_dev[(tile * kpad + col) * R + r]
//   kLayoutWarp (glm_warp.cu)  8 consumer warps; the C column groups of a row sit in one warp:
//               rows per warp rpw = 32 / C, R = 8 * rpw, element (r, col) of a tile with
//               w = r / rpw, rl = r % rpw, cg = col / kc, j = col % kc lives at
//               x_dev[tile * kpad * R + (j * 8 + w) * 32 + rl * C + cg]
//   kLayoutMma  (glm_categorical_mma.cu) fragments of 8 rows x 4 columns in the lane order of the FP64
//               mma.m8n8k4 A operand; kpad = 8 * ceil(k / 8), R = 64 (kpad <= 64) or 32:
//               x_dev[tile * kpad * R + ((r / 8) * (kpad / 4) + col / 4) * 32 + (r % 8) * 4 + col % 4]
constexpr int kLayoutCta = 0;
constexpr int kLayoutWarp = 1;
constexpr int kLayoutMma = 2;
constexpr int kLayoutPlain = 3;  // glm_categorical_wide.cu: plain column-major (one "tile" of R = padded row count)

struct Tiling {
  int layout;   // kLayoutCta | kLayoutWarp
  int threads;  // consumer threads per CTA (128 | 256 | 512)
  int C;        // column groups (power of two)
  int kc;       // columns per thread, 1..16
  int R;        // rows per tile = threads / C (power of two, >= 8)
  int kpad;     // C * kc >= k
  int64_t tiles;
  size_t tile_bytes() const { return size_t(kpad) * R * sizeof(double); }
};

// Picks C, kc, R for k columns and n rows; returns false if k is outside the fused kernel's range.
bool choose_tiling(int64_t n, int64_t k, int threads_override, Tiling* t, int force_layout = -1);

struct Workspace {   // per (host thread, device) or per data handle: one launch in flight at a time
  int device = -1;
  double* scratch = nullptr;  // [max_grid][kpad_cap + 4] per-CTA partials
  unsigned int* ticket = nullptr;
  size_t scratch_doubles = 0;
  bool fixed = false;  // a view of memory that must never be reallocated (stan_cuda_glm_launch_noalloc): reserve fails instead
};
int workspace_reserve(Workspace* w, int device, size_t doubles);
void workspace_free(Workspace* w);

}  // namespace stancuda

// One GLM data block resident on one device.
struct stan_cuda_data {
  int device = 0;
  int family = 0;
  int64_t n = 0, k = 0;
  int y_is_double = 0;
  stancuda::Tiling t{};
  double* x = nullptr;   // tiled layout, zero padded
  void* y = nullptr;     // int32 or double, padded to tiles * R
  double lgamma_sum = 0; // sum lgamma(y+1) (poisson), computed once at fill
  int64_t y_min = 0, y_max = 0;  // range of the integer outcomes, computed once at fill
  unsigned int* hist = nullptr;  // neg_binomial_2_log: histogram of y (counts of 0 .. y_max), built at fill
  int hist_len = 0;
  size_t bytes = 0;
  stancuda::Workspace ws;  // used by the device-level API (stan_cuda_glm_launch)
  // k > 512: the columns are split into blocks of kColBlock, each a complete data block of its own (x tiled for
  // the warp kernel; the last one carries the real family, the others kFamilyLinearBlock).  This head block then
  // holds no x of its own; its t is the last block's tiling with kpad = the sum over the blocks.
  std::vector<stan_cuda_data*> colblocks;
};

namespace stancuda {

struct PeerXchg;

struct LaunchIO {
  uint32_t flags;
  const double* d_params;     // [kpad] beta, [kpad] alpha, [kpad+1] sigma
  const double* d_alpha_vec;  // n or null
  const double* d_sigma_vec;  // n or null
  double* d_result;           // [kpad + 4]
  double* d_row_dalpha;       // n or null
  double* d_row_dsigma;       // n or null
  const struct PeerXchg* xchg = nullptr;  // non-null: all-reduce fused into the kernel (comm.cu)
  unsigned long long xchg_seq = 0;
  // low-latency host path (api.cu): parameter block passed by value in the kernel arguments
  // (no H2D copy) and a completion flag in mapped pinned host memory, written by the kernel after
  // the result block (which then also lives in mapped host memory: no D2H copy, no stream sync)
  const double* h_params_inline = nullptr;   // host pointer, params_len doubles, <= kInlineParams
  unsigned long long* host_flag = nullptr;   // device-visible address of the flag
  unsigned long long host_seq = 0;           // value to store
  // ordered_logistic: cut-points follow the parameter block (d_params[kpad + 2 + j]), d/dcuts follow
  // the result block (d_result[kpad + 4 + j])
  int ncuts = 0;
  // categorical_logit: d_params = [beta: ncls rows of kpad (class-major), alpha(ncls)],
  // d_result = [lp, 0, 0, status, dalpha(ncls), dbeta: ncls rows of kpad]
  int ncls = 0;
  bool shared_gpu = false;  // other launches are expected to run concurrently on this device (map_rect shard scope)
  // per-call outcomes (stan_cuda_glm_args.y_per_call): device vector padded to tiles * R elements that replaces the
  // cached y of the data block for this launch, with its own sum lgamma(y + 1); no histogram of y is available
  const void* d_y = nullptr;
  double y_lgamma_sum = 0.0;
  // var x: d/dx = d * beta^T streamed out tile by tile (device matrix, column-major rows x k, leading dimension ld_dx)
  double* d_dx = nullptr;
  int64_t ld_dx = 0;
  // run the block as this family instead of its own (the many-cut-points path of ordered_logistic drives every block of
  // the matrix as a kFamilyLinearBlock: glm_ordered_wide.cu)
  int family_override = -1;
};
// Internal pseudo-family of the column-block path (k > 512, glm_warp.cu launch_glm_colblocks): a block of columns
// contributes eta_b = X_b beta_b to a per-row carry vector, or dbeta_b = X_b^T d from a per-row d vector.  Never
// seen at the C ABI.
constexpr int kFamilyLinearBlock = 6;
constexpr int kColBlock = 512;               // columns per block: the widest tile of the warp kernel (C = 32, kc = 16)
constexpr int kMaxClasses = 16;              // categorical_logit: classes on the device path
constexpr int kMaxClassWeights = 4096;       // categorical_logit: classes * k_pad (weights live in shared memory)
constexpr int kMaxCuts = 1024;               // ordered_logistic: up to here the per-warp bins and the class table live in shared
                                             // memory of the fused kernel; beyond: glm_ordered_wide.cu
constexpr int kOrdPrivateDoubles = 2048;     // ordered_logistic: ncuts * R up to this -> one private bin row per row slot
constexpr int kInlineParams = 264;           // covers kpad <= 256 (+ alpha, sigma)

// Enqueue the fused kernel for one data block.  ws must belong to the block's device.
int launch_glm(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
int launch_glm_warp(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
// var x: dx[j * rows + i] = d[i] * params[j]  (d/dx = d * beta^T; data.cu)
int launch_dx_outer(const double* d_row, const double* d_params, int64_t rows, int64_t k, double* dx, cudaStream_t stream);
int launch_glm_colblocks(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
// ordered_logistic with more than kMaxCuts cut-points (glm_ordered_wide.cu)
int launch_glm_ordered_wide(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
int launch_glm_categorical(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
int launch_glm_categorical_mma(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
int launch_glm_categorical_wide(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream);
// categorical_logit shapes outside the two fused kernels (weights in shared memory: <= 16 classes, classes * k_pad <= 4096,
// k <= 512) run on the general three-pass path over plain column-major data (glm_categorical_wide.cu)
inline bool categorical_is_wide(int64_t k, int64_t ncls) {
  return ncls > 16 || k > 512 || ncls * ((k + 15) / 16 * 16) > 4096;
}
// number of classes of the categorical_logit call for which the next data block of this thread is created (0: unknown)
void set_categorical_classes_hint(int64_t ncls);
// width of the result block (and of every partial) of one launch
inline size_t result_cols(const stan_cuda_data* d, const LaunchIO& io) {
  if (d->family == STAN_CUDA_CATEGORICAL_LOGIT) return size_t(4) + size_t(io.ncls) * (d->t.kpad + 1);
  return size_t(d->t.kpad) + 4 + (d->family == STAN_CUDA_ORDERED_LOGISTIC ? io.ncuts : 0);
}
int sm_count(int device);

// tuning knobs (env STAN_CUDA_THREADS / STAN_CUDA_STAGES / STAN_CUDA_CTAS_PER_SM; 0 = auto)
struct Tuning { int threads = 0, stages = 0, ctas_per_sm = 0, layout = kLayoutWarp, blocked = 0, cat_mma = 1; };
const Tuning& tuning();

// relayout kernels (data.cu)
int upload_glm_data(stan_cuda_data* d, const double* x, int64_t ldx, int x_row_major, const void* y,
                    int64_t row0, cudaStream_t stream, uint64_t* bytes_uploaded);

// normal_lpdf kernel (normal_kernel.cu)
struct NormalIO {
  uint32_t flags;
  int family;                 // STAN_CUDA_LOCSCALE_*
  int64_t n;
  const double* y; int y_vec;
  const double* mu; int mu_vec;
  const double* sigma; int sigma_vec;
  double y0, mu0, sigma0;     // scalar operands by value
  double* dmu_row; double* dsigma_row; double* dy_row;  // per-row outputs (vector operands) or null
  double* result;             // [6]: lp, sum dmu, sum dsigma, witness, status bits, sum dy
};
int launch_normal_lpdf(const NormalIO& io, Workspace* ws, int device, cudaStream_t stream);

// dense FP64 product on raw device pointers (gemm.cu): C = alpha * op(A) * op(B) + beta * C, column-major
int dgemm(int device, bool trans_a, bool trans_b, bool lower_only, int64_t M, int64_t N, int64_t K, double alpha, const double* A,
          int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc, cudaStream_t stream);

// student_t_lpdf kernel (student_t.cu)
struct StudentTIO {
  uint32_t flags;
  int64_t n;
  const double* y; const double* nu; const double* mu; const double* sigma;  // device vectors, or null for a scalar operand
  double y0, nu0, mu0, sigma0;                                              // scalar operands by value
  // scalar nu: the per-nu terms of :94-117, evaluated once on the host (student_t_scalar_dof_terms)
  double dof_term0;   // lgamma(nu/2 + 1/2) - lgamma(nu/2) - log(nu)/2
  double dnu_const0;  // digamma(nu/2 + 1/2)/2 - digamma(nu/2)/2 - 1/(2 nu)
  double* dy_row; double* dnu_row; double* dmu_row; double* dsigma_row;     // per-element outputs (vector operands) or null
  double* result;  // [8]: lp, sum dy, sum dnu, sum dmu, sum dsigma, witness, bad-operand bits, 0
};
void student_t_scalar_dof_terms(StudentTIO* io);
int launch_student_t_lpdf(const StudentTIO& io, Workspace* ws, int device, cudaStream_t stream);

// vectorised elementwise densities (density.cu)
struct DensityIO {
  int density;
  uint32_t flags;
  int64_t n;
  const int* y; int y_vec; int y0;
  const double* theta; int theta_vec; double theta0;
  // poisson_log: sum lgamma(n + 1) over all elements, computed once when the outcomes were cached (data only).  The
  // elements poisson_log_lpmf.hpp:77 skips have n == 0, i.e. lgamma(1) = 0: the total needs no correction.
  int lgamma_cached;
  double lgamma_total;
  double* dtheta_row;         // per-element partials (vector operand) or null
  double* result;             // [5]: lp, sum dtheta, 0, witness, bits (1: outcome outside the support, 2: operand NaN, 4: LOG_ZERO)
};
int launch_density(const DensityIO& io, Workspace* ws, int device, cudaStream_t stream);
int int_vector_stats(const int* d_y, int64_t n, int want_lgamma, cudaStream_t stream, double* lgamma_sum, int64_t* y_min,
                     int64_t* y_max);

// counters
struct Counters {
  std::atomic<uint64_t> cache_hits{0}, cache_misses{0}, bytes_uploaded{0}, evals{0}, launches{0},
      allreduces{0}, fused_allreduces{0}, cache_evictions{0};
  std::atomic<double> last_fill_ms{0};
};
Counters& counters();

}  // namespace stancuda
