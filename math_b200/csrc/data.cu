// data.cu -- cache fill: host (x, y) -> tiled device layout.  Replaces matrix_cl + to_matrix_cl
// (stan/math/opencl/matrix_cl.hpp:32-454, stan/math/opencl/copy.hpp:37-93) for the GLM data.
//
// The device layout is chosen for the fused kernel, not for the host: tile by tile, each tile
// R rows x kpad columns, column-major inside the tile (see common.cuh / DESIGN.md).  The host
// matrix (Eigen column-major, or row-major) is streamed through a bounded staging buffer and
// re-tiled by a kernel, so the fill never needs a second full-size copy of X on the device.
#include <limits.h>
#include <math.h>

#include <chrono>

#include "context.cuh"

namespace stancuda {

// Position of element (r, col) inside a tile, for either layout (common.cuh).
struct TileMap {
  int layout, R, C, kc, kpad;
};
__device__ __forceinline__ int64_t tile_elem_offset(const TileMap& m, int r, int col) {
  if (m.layout == kLayoutCta || m.layout == kLayoutPlain) return int64_t(col) * m.R + r;
  if (m.layout == kLayoutMma) return (int64_t(r >> 3) * (m.kpad >> 2) + (col >> 2)) * 32 + (r & 7) * 4 + (col & 3);
  const int rpw = 32 / m.C;
  const int w = r / rpw, rl = r - w * rpw;
  const int cg = col / m.kc, j = col - cg * m.kc;
  return int64_t(j * 8 + w) * 32 + rl * m.C + cg;
}

// staging chunk -> tiled X.  Column-major chunks hold rows [0, n) x columns [c0, c0 + nc) of the
// block: src[jj * src_ld + i]; row-major chunks hold rows [i0, i0 + ni) x all k columns:
// src[(i - i0) * src_ld + j].
__global__ void retile_colmajor_kernel(const double* __restrict__ src, int64_t src_ld, int64_t n, int c0, int nc,
                                       double* __restrict__ dst, TileMap m, int64_t tiles) {
  const int64_t per_col = tiles * m.R;
  const int64_t total = per_col * nc;
  const int64_t tile_elems = int64_t(m.kpad) * m.R;
  for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < total;
       idx += int64_t(gridDim.x) * blockDim.x) {
    const int64_t jj = idx / per_col;
    const int64_t i = idx - jj * per_col;  // row index (padded range)
    const int64_t tile = i / m.R;
    const int r = int(i - tile * m.R);
    const double v = i < n ? src[jj * src_ld + i] : 0.0;
    dst[tile * tile_elems + tile_elem_offset(m, r, int(c0 + jj))] = v;
  }
}

__global__ void retile_rowmajor_kernel(const double* __restrict__ src, int64_t src_ld, int64_t i0, int64_t ni,
                                       int k, double* __restrict__ dst, TileMap m) {
  // one thread per destination element of the row range; r fastest
  const int64_t total = ni * k;
  const int64_t tile_elems = int64_t(m.kpad) * m.R;
  for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < total;
       idx += int64_t(gridDim.x) * blockDim.x) {
    const int64_t j = idx / ni;
    const int64_t ii = idx - j * ni;
    const int64_t i = i0 + ii;
    const int64_t tile = i / m.R;
    const int r = int(i - tile * m.R);
    dst[tile * tile_elems + tile_elem_offset(m, r, int(j))] = src[ii * src_ld + j];
  }
}

// Statistics of the integer outcomes, computed once per fill (the data is immutable):
//   min / max   -> the eager domain checks check_bounded(y, 0, 1) / check_nonnegative(y)
//                  (bernoulli_logit_glm_lpmf.hpp:70, poisson_log_glm_lpmf.hpp:70) without a per-call
//                  O(N) host loop;
//   sum lgamma(y + 1) (poisson_log_glm_lpmf.hpp:102-108), depends on data only.
// Per-block partials in block order, combined on the host in order.
// Histogram of the integer outcomes (neg_binomial_2_log; integer atomics: exact, order-free).
__global__ void y_hist_kernel(const int* __restrict__ y, int64_t n, unsigned int* __restrict__ hist) {
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
    atomicAdd(&hist[y[i]], 1u);
}

__global__ void y_stats_kernel(const int* __restrict__ y, int64_t n, int want_lgamma, double* __restrict__ partial,
                               int* __restrict__ minmax) {
  __shared__ double sh[256];
  __shared__ int smin[256], smax[256];
  double s = 0.0;
  int lo = INT_MAX, hi = INT_MIN;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
    const int v = y[i];
    lo = v < lo ? v : lo;
    hi = v > hi ? v : hi;
    if (want_lgamma) s += v >= 0 ? lgamma(double(v) + 1.0) : 0.0;
  }
  sh[threadIdx.x] = s;
  smin[threadIdx.x] = lo;
  smax[threadIdx.x] = hi;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (int(threadIdx.x) < off) {
      sh[threadIdx.x] += sh[threadIdx.x + off];
      smin[threadIdx.x] = min(smin[threadIdx.x], smin[threadIdx.x + off]);
      smax[threadIdx.x] = max(smax[threadIdx.x], smax[threadIdx.x + off]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partial[blockIdx.x] = sh[0];
    minmax[2 * blockIdx.x] = smin[0];
    minmax[2 * blockIdx.x + 1] = smax[0];
  }
}

// min / max / sum lgamma(v + 1) of a device vector of ints (current device, synchronises the stream)
int int_vector_stats(const int* d_y, int64_t n, int want_lgamma, cudaStream_t stream, double* lgamma_sum, int64_t* y_min,
                     int64_t* y_max) {
  const int nb = 256;
  double* partial = nullptr;  // [nb] doubles followed by [2 * nb] ints
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&partial), nb * (sizeof(double) + 2 * sizeof(int))));
  int* mm = reinterpret_cast<int*>(partial + nb);
  y_stats_kernel<<<nb, 256, 0, stream>>>(d_y, n, want_lgamma, partial, mm);
  double host[256];
  int hmm[512];
  cudaError_t e = cudaMemcpyAsync(host, partial, sizeof(host), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(hmm, mm, sizeof(hmm), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  cudaFree(partial);
  SC_CUDA(e);
  double s = 0.0;
  int lo = INT_MAX, hi = INT_MIN;
  for (int i = 0; i < nb; ++i) {
    s += host[i];
    lo = hmm[2 * i] < lo ? hmm[2 * i] : lo;
    hi = hmm[2 * i + 1] > hi ? hmm[2 * i + 1] : hi;
  }
  if (lgamma_sum) *lgamma_sum = s;
  if (y_min) *y_min = lo;
  if (y_max) *y_max = hi;
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

// var x (bernoulli_logit_glm_lpmf.hpp:125-128, poisson_log_glm_lpmf.hpp:119-122, normal_id_glm_lpdf.hpp:121-124):
// d/dx = (beta d^T)^T, an n x k outer product -- a pure HBM write stream of n * k * 8 bytes (the n + k inputs stay in
// L1/L2).  One thread writes two consecutive rows of one column (16-byte stores, coalesced along the rows).
__global__ void __launch_bounds__(256) dx_outer_kernel(const double* __restrict__ d_row, const double* __restrict__ beta,
                                                       int64_t rows, int64_t k, double* __restrict__ dx) {
  const int64_t pairs = (rows + 1) / 2;
  const int64_t total = pairs * k;
  for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < total; idx += int64_t(gridDim.x) * blockDim.x) {
    const int64_t j = idx / pairs, i = (idx - j * pairs) * 2;
    const double b = beta[j];
    double* dst = dx + j * rows + i;
    if (i + 1 < rows && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
      *reinterpret_cast<double2*>(dst) = make_double2(d_row[i] * b, d_row[i + 1] * b);
    } else {
      dst[0] = d_row[i] * b;
      if (i + 1 < rows) dst[1] = d_row[i + 1] * b;
    }
  }
}

int launch_dx_outer(const double* d_row, const double* d_params, int64_t rows, int64_t k, double* dx, cudaStream_t stream) {
  if (rows <= 0 || k <= 0) return STAN_CUDA_OK;
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  const int64_t total = (rows + 1) / 2 * k;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = int64_t(sm_count(dev)) * 8;
  if (blocks > cap) blocks = cap;
  dx_outer_kernel<<<unsigned(blocks), 256, 0, stream>>>(d_row, d_params, rows, k, dx);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

static const size_t kStageBytes = size_t(256) << 20;  // staging buffer for the fill

// Upload rows [row0, row0 + d->n) of the host matrix into d (already sized; d->t chosen).
int upload_glm_data(stan_cuda_data* d, const double* x, int64_t ldx, int x_row_major, const void* y,
                    int64_t row0, cudaStream_t stream, uint64_t* bytes_uploaded) {
  NvtxRange range("stan_cuda upload_glm_data");
  const Tiling& t = d->t;
  const int64_t n = d->n, k = d->k;
  SC_CUDA(cudaSetDevice(d->device));
  const size_t x_bytes = size_t(t.tiles) * t.tile_bytes();
  const size_t y_elem = d->y_is_double ? sizeof(double) : sizeof(int);
  const size_t y_bytes = size_t(t.tiles) * t.R * y_elem;
  // (a refill of an existing block -- per-call x, refill_data_block -- keeps its allocations: same shape, same sizes)
  if (!d->x) SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d->x), x_bytes));
  if (!d->y) SC_CUDA(cudaMalloc(&d->y, y_bytes));
  if (d->hist) {
    SC_CUDA(cudaFree(d->hist));
    d->hist = nullptr;
    d->hist_len = 0;
  }
  d->bytes = x_bytes + y_bytes;
  // zero fill covers the padding columns [k, kpad) (padding rows are written by the retile kernel)
  if (t.kpad != k || x_row_major) SC_CUDA(cudaMemsetAsync(d->x, 0, x_bytes, stream));
  SC_CUDA(cudaMemsetAsync(d->y, 0, y_bytes, stream));
  SC_CUDA(cudaMemcpyAsync(d->y, static_cast<const char*>(y) + size_t(row0) * y_elem, size_t(n) * y_elem,
                          cudaMemcpyHostToDevice, stream));
  uint64_t up = uint64_t(n) * y_elem;

  double* stage[2] = {nullptr, nullptr};
  cudaEvent_t freed[2] = {nullptr, nullptr};
  int rc = STAN_CUDA_OK;
  auto cleanup = [&]() {
    for (int i = 0; i < 2; ++i) {
      if (stage[i]) cudaFree(stage[i]);
      if (freed[i]) cudaEventDestroy(freed[i]);
    }
  };
#define SC_CUDA_CLEAN(expr)                                          \
  do {                                                               \
    cudaError_t _e = (expr);                                         \
    if (_e != cudaSuccess) {                                         \
      rc = fail_cuda(_e, #expr, __FILE__, __LINE__);                 \
      cleanup();                                                     \
      return rc;                                                     \
    }                                                                \
  } while (0)

  const int blocks = sm_count(d->device) * 8;
  const TileMap tmap{t.layout, t.R, t.C, t.kc, t.kpad};
  if (!x_row_major) {
    // chunk = a range of whole columns (rows row0 .. row0+n of each)
    int64_t cols_per_chunk = int64_t(kStageBytes / (size_t(n) * sizeof(double)));
    if (cols_per_chunk < 1) cols_per_chunk = 1;
    if (cols_per_chunk > k) cols_per_chunk = k;
    const size_t stage_doubles = size_t(cols_per_chunk) * n;
    for (int i = 0; i < 2; ++i) {
      SC_CUDA_CLEAN(cudaMalloc(reinterpret_cast<void**>(&stage[i]), stage_doubles * sizeof(double)));
      SC_CUDA_CLEAN(cudaEventCreateWithFlags(&freed[i], cudaEventDisableTiming));
    }
    int which = 0;
    for (int64_t c0 = 0; c0 < k; c0 += cols_per_chunk, which ^= 1) {
      const int64_t nc = (k - c0 < cols_per_chunk) ? (k - c0) : cols_per_chunk;
      SC_CUDA_CLEAN(cudaEventSynchronize(freed[which]));
      if (ldx == n && row0 == 0) {
        SC_CUDA_CLEAN(cudaMemcpyAsync(stage[which], x + size_t(c0) * ldx, size_t(nc) * n * sizeof(double),
                                      cudaMemcpyHostToDevice, stream));
      } else {
        SC_CUDA_CLEAN(cudaMemcpy2DAsync(stage[which], size_t(n) * sizeof(double), x + size_t(c0) * ldx + row0,
                                        size_t(ldx) * sizeof(double), size_t(n) * sizeof(double), size_t(nc),
                                        cudaMemcpyHostToDevice, stream));
      }
      retile_colmajor_kernel<<<blocks, 256, 0, stream>>>(stage[which], n, n, int(c0), int(nc), d->x, tmap, t.tiles);
      SC_CUDA_CLEAN(cudaGetLastError());
      SC_CUDA_CLEAN(cudaEventRecord(freed[which], stream));
      up += uint64_t(nc) * n * sizeof(double);
    }
  } else {
    int64_t rows_per_chunk = int64_t(kStageBytes / (size_t(ldx) * sizeof(double)));
    if (rows_per_chunk < 1) rows_per_chunk = 1;
    if (rows_per_chunk > n) rows_per_chunk = n;
    const size_t stage_doubles = size_t(rows_per_chunk) * ldx;
    for (int i = 0; i < 2; ++i) {
      SC_CUDA_CLEAN(cudaMalloc(reinterpret_cast<void**>(&stage[i]), stage_doubles * sizeof(double)));
      SC_CUDA_CLEAN(cudaEventCreateWithFlags(&freed[i], cudaEventDisableTiming));
    }
    int which = 0;
    for (int64_t i0 = 0; i0 < n; i0 += rows_per_chunk, which ^= 1) {
      const int64_t ni = (n - i0 < rows_per_chunk) ? (n - i0) : rows_per_chunk;
      SC_CUDA_CLEAN(cudaEventSynchronize(freed[which]));
      // last row of the chunk may be shorter than ldx in the host allocation: copy (ni-1)*ldx + k
      const size_t cnt = size_t(ni - 1) * ldx + k;
      SC_CUDA_CLEAN(cudaMemcpyAsync(stage[which], x + size_t(row0 + i0) * ldx, cnt * sizeof(double),
                                    cudaMemcpyHostToDevice, stream));
      retile_rowmajor_kernel<<<blocks, 256, 0, stream>>>(stage[which], ldx, i0, ni, int(k), d->x, tmap);
      SC_CUDA_CLEAN(cudaGetLastError());
      SC_CUDA_CLEAN(cudaEventRecord(freed[which], stream));
      up += uint64_t(cnt) * sizeof(double);
    }
  }
  counters().launches.fetch_add(1, std::memory_order_relaxed);

  d->lgamma_sum = 0.0;
  d->y_min = d->y_max = 0;
  if (!d->y_is_double) {
    const int want_lgamma = d->family == STAN_CUDA_POISSON_LOG || d->family == STAN_CUDA_NEG_BINOMIAL_2_LOG;
    double s = 0.0;
    int64_t lo = 0, hi = 0;
    {
      const int rc_stats = int_vector_stats(static_cast<const int*>(d->y), n, want_lgamma, stream, &s, &lo, &hi);
      if (rc_stats < 0) {
        cleanup();
        return rc_stats;
      }
    }
    d->lgamma_sum = s;
    d->y_min = lo;
    d->y_max = hi;
    // neg_binomial_2_log: histogram of y, so that sum lgamma(y + phi) and sum digamma(y + phi) cost
    // O(y_max) per evaluation instead of O(N) special-function calls in the hot loop
    if (d->family == STAN_CUDA_NEG_BINOMIAL_2_LOG && lo >= 0 && hi < (1 << 20)) {
      d->hist_len = hi + 1;
      SC_CUDA_CLEAN(cudaMalloc(reinterpret_cast<void**>(&d->hist), size_t(d->hist_len) * sizeof(unsigned int)));
      SC_CUDA_CLEAN(cudaMemsetAsync(d->hist, 0, size_t(d->hist_len) * sizeof(unsigned int), stream));
      y_hist_kernel<<<sm_count(d->device) * 4, 256, 0, stream>>>(static_cast<const int*>(d->y), n, d->hist);
      SC_CUDA_CLEAN(cudaGetLastError());
      d->bytes += size_t(d->hist_len) * sizeof(unsigned int);
    }
  }
  SC_CUDA_CLEAN(cudaStreamSynchronize(stream));
#undef SC_CUDA_CLEAN
  cleanup();
  if (bytes_uploaded) *bytes_uploaded = up;
  return STAN_CUDA_OK;
}

static thread_local int64_t tl_categorical_classes = 0;
void set_categorical_classes_hint(int64_t ncls) { tl_categorical_classes = ncls; }

static thread_local bool tl_block_of_wider_matrix = false;  // set while the column blocks of a k > 512 matrix are built

int make_data_block(int device, int family, int64_t n_rows, int64_t k, const double* x, int64_t ldx,
                    int x_row_major, const void* y, int y_is_double, int64_t row0, cudaStream_t stream,
                    stan_cuda_data** out) {
  if (n_rows <= 0 || k < 0 || (!x && k > 0) || !y) return fail(STAN_CUDA_ERR_ARG, "data block: bad shape or null pointer");
  if (family < 0 || (family > STAN_CUDA_CATEGORICAL_LOGIT && family != kFamilyLinearBlock))
    return fail(STAN_CUDA_ERR_ARG, "data block: unknown family");
  if (k > kColBlock && family != STAN_CUDA_CATEGORICAL_LOGIT && family != kFamilyLinearBlock) {
    // wider than one warp tile: column blocks (glm_warp.cu, launch_glm_colblocks).  Only the last block is padded,
    // so the concatenated parameter / gradient layout is [beta(k), zeros] like that of a single block.
    const int64_t B = (k + kColBlock - 1) / kColBlock;
    if (B > 4096) return fail(STAN_CUDA_ERR_ARG, "data block: more than 2097152 columns");
    stan_cuda_data* head = new stan_cuda_data();
    head->device = device;
    head->family = family;
    head->n = n_rows;
    head->k = k;
    head->y_is_double = y_is_double;
    for (int64_t b = 0; b < B; ++b) {
      const int64_t c0 = b * kColBlock;
      const int64_t kb = (k - c0 < kColBlock) ? (k - c0) : int64_t(kColBlock);
      const double* xb = x_row_major ? x + c0 : x + c0 * ldx;
      stan_cuda_data* sub = nullptr;
      tl_block_of_wider_matrix = true;  // every block on the warp kernel's layout, whatever variant is selected
      const int rc = make_data_block(device, b == B - 1 ? family : kFamilyLinearBlock, n_rows, kb, xb, ldx, x_row_major, y,
                                     y_is_double, row0, stream, &sub);
      tl_block_of_wider_matrix = false;
      if (rc < 0) {
        free_data_block(head);
        return rc;
      }
      head->colblocks.push_back(sub);
      head->bytes += sub->bytes;
    }
    const stan_cuda_data* last = head->colblocks.back();
    head->t = last->t;
    head->t.kpad = int((B - 1) * kColBlock) + last->t.kpad;
    head->lgamma_sum = last->lgamma_sum;
    head->y_min = last->y_min;
    head->y_max = last->y_max;
    *out = head;
    return STAN_CUDA_OK;
  }
  stan_cuda_data* d = new stan_cuda_data();
  d->device = device;
  d->family = family;
  d->n = n_rows;
  d->k = k;
  d->y_is_double = y_is_double;
  // neg_binomial_2_log and ordered_logistic exist on the warp kernel only; categorical_logit has its own two
  // kernels: FP64 tensor-core (fragment layout, k <= 256) or vector (warp layout)
  int force = (family >= STAN_CUDA_NEG_BINOMIAL_2_LOG || tl_block_of_wider_matrix) ? kLayoutWarp : -1;
  if (family == STAN_CUDA_CATEGORICAL_LOGIT && tuning().cat_mma && k <= 256) force = kLayoutMma;
  if (family == STAN_CUDA_CATEGORICAL_LOGIT && tl_categorical_classes > 0 && categorical_is_wide(k, tl_categorical_classes))
    force = kLayoutPlain;
  if (force == kLayoutPlain && n_rows >= (int64_t(1) << 31) - 2) {
    delete d;
    return fail(STAN_CUDA_ERR_ARG, "data block: the general categorical_logit path holds fewer than 2^31 rows per device");
  }
  if (!choose_tiling(n_rows, k, 0, &d->t, force)) {
    delete d;
    return fail(STAN_CUDA_ERR_ARG, "data block: categorical_logit with k > 512 needs the number of classes (stan_cuda_glm_eval); "
                                   "the device-level stan_cuda_data_create serves k <= 512");
  }
  auto t0 = std::chrono::steady_clock::now();
  uint64_t up = 0;
  int rc = upload_glm_data(d, x, ldx, x_row_major, y, row0, stream, &up);
  if (rc < 0) {
    if (d->x) cudaFree(d->x);
    if (d->y) cudaFree(d->y);
    if (d->hist) cudaFree(d->hist);
    delete d;
    return rc;
  }
  auto t1 = std::chrono::steady_clock::now();
  counters().bytes_uploaded.fetch_add(up);
  counters().last_fill_ms.store(std::chrono::duration<double, std::milli>(t1 - t0).count());
  *out = d;
  return STAN_CUDA_OK;
}

// Per-call x (stan_cuda_glm_args.x_per_call): the same shape again with new contents, into the block's existing
// device buffers (no cudaMalloc / cudaFree per evaluation).
int refill_data_block(stan_cuda_data* d, const double* x, int64_t ldx, int x_row_major, const void* y, int64_t row0,
                      cudaStream_t stream) {
  uint64_t up = 0;
  if (!d->colblocks.empty()) {
    for (size_t b = 0; b < d->colblocks.size(); ++b) {
      const int64_t c0 = int64_t(b) * kColBlock;
      SC_TRY(refill_data_block(d->colblocks[b], x_row_major ? x + c0 : x + c0 * ldx, ldx, x_row_major, y, row0, stream));
    }
    const stan_cuda_data* last = d->colblocks.back();
    d->lgamma_sum = last->lgamma_sum;
    d->y_min = last->y_min;
    d->y_max = last->y_max;
    return STAN_CUDA_OK;
  }
  SC_TRY(upload_glm_data(d, x, ldx, x_row_major, y, row0, stream, &up));
  counters().bytes_uploaded.fetch_add(up);
  return STAN_CUDA_OK;
}

void free_data_block(stan_cuda_data* d) {
  if (!d) return;
  for (stan_cuda_data* sub : d->colblocks) free_data_block(sub);
  d->colblocks.clear();
  cudaSetDevice(d->device);
  if (d->x) cudaFree(d->x);
  if (d->y) cudaFree(d->y);
  if (d->hist) cudaFree(d->hist);
  workspace_free(&d->ws);
  delete d;
}

}  // namespace stancuda

using namespace stancuda;

extern "C" {

int stan_cuda_data_create(int family, int64_t n, int64_t k, const double* x, int64_t ldx, int x_row_major,
                          const void* y, int y_is_double, stan_cuda_data_t* out) {
  if (!out) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_data_create: null out");
  SC_TRY(ensure_init());
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  // the caller's current device must be one of ours; default to the first
  bool ok = false;
  for (int d : context().devices) ok = ok || d == dev;
  if (!ok) dev = context().devices[0];
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  set_categorical_classes_hint(0);  // the number of classes is unknown here: the fused kernels' layouts
  return make_data_block(dev, family, n, k, x, ldx, x_row_major, y, y_is_double, 0, td->stream, out);
}

void stan_cuda_data_destroy(stan_cuda_data_t d) { free_data_block(d); }

int64_t stan_cuda_data_params_len(stan_cuda_data_t d) { return d ? d->t.kpad + 2 : 0; }
int64_t stan_cuda_data_result_len(stan_cuda_data_t d) { return d ? d->t.kpad + 4 : 0; }
int64_t stan_cuda_data_kpad(stan_cuda_data_t d) { return d ? d->t.kpad : 0; }
int64_t stan_cuda_data_algorithmic_bytes(stan_cuda_data_t d) {
  if (!d) return 0;
  return d->n * d->k * int64_t(sizeof(double)) + d->n * int64_t(d->y_is_double ? sizeof(double) : sizeof(int));
}
int stan_cuda_data_launches_per_eval(stan_cuda_data_t d) { return d ? 1 : 0; }
int stan_cuda_data_y_range(stan_cuda_data_t d, int64_t* y_min, int64_t* y_max) {
  if (!d) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_data_y_range: null handle");
  if (y_min) *y_min = d->y_min;
  if (y_max) *y_max = d->y_max;
  return STAN_CUDA_OK;
}

static int glm_launch_common(stan_cuda_data_t d, uint32_t flags, LaunchIO io, void* stream) {
  SC_CUDA(cudaSetDevice(d->device));
  if (flags & STAN_CUDA_FUSED_ALLREDUCE) {
    PeerXchg* x = comm_rank_xchg();
    if (!x) return fail(STAN_CUDA_ERR_STATE, "stan_cuda_glm_launch: STAN_CUDA_FUSED_ALLREDUCE without a peer exchange "
                                             "(stan_cuda_comm_fused_available() is 0)");
    io.xchg = x;
    io.xchg_seq = ++x->seq;
    counters().fused_allreduces.fetch_add(1, std::memory_order_relaxed);
  }
  return launch_glm(d, io, &d->ws, static_cast<cudaStream_t>(stream));
}

int stan_cuda_glm_launch(stan_cuda_data_t d, uint32_t flags, const double* d_params, const double* d_alpha_vec,
                         const double* d_sigma_vec, double* d_result, double* d_row_dalpha,
                         double* d_row_dsigma, void* stream) {
  if (!d || !d_params || !d_result) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch: null argument");
  if (d->family == STAN_CUDA_ORDERED_LOGISTIC || d->family == STAN_CUDA_CATEGORICAL_LOGIT)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch: ordered_logistic / categorical_logit data goes through "
                                   "stan_cuda_glm_launch_ordered / _categorical");
  LaunchIO io{flags, d_params, d_alpha_vec, d_sigma_vec, d_result, d_row_dalpha, d_row_dsigma};
  return glm_launch_common(d, flags, io, stream);
}

int stan_cuda_glm_launch_categorical(stan_cuda_data_t d, uint32_t flags, const double* d_params, int n_classes,
                                     double* d_result, void* stream) {
  if (!d || !d_params || !d_result) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_categorical: null argument");
  if (d->family != STAN_CUDA_CATEGORICAL_LOGIT)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_categorical: data block of another family");
  LaunchIO io{flags, d_params, nullptr, nullptr, d_result, nullptr, nullptr};
  io.ncls = n_classes;
  return glm_launch_common(d, flags, io, stream);
}

int stan_cuda_glm_launch_ordered(stan_cuda_data_t d, uint32_t flags, const double* d_params, int ncuts,
                                 double* d_result, void* stream) {
  if (!d || !d_params || !d_result) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_ordered: null argument");
  if (d->family != STAN_CUDA_ORDERED_LOGISTIC)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_ordered: data block of another family");
  LaunchIO io{flags, d_params, nullptr, nullptr, d_result, nullptr, nullptr};
  io.ncuts = ncuts;
  return glm_launch_common(d, flags, io, stream);
}

}  // extern "C"
