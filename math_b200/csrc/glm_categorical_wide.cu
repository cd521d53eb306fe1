// glm_categorical_wide.cu -- categorical_logit_glm_lpmf for the shapes the two fused kernels do not hold:
// more than 16 classes, classes * k_pad > 4096, or more than 512 columns.
//
// The reference accepts any shape (stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:56-64 checks consistency
// only).  The fused kernels (glm_categorical_mma.cu, glm_categorical.cu) keep the whole K x C weight matrix in shared
// memory and one row's C linear predictors in a handful of lanes; beyond that the evaluation runs as the reference
// itself does it (:74-190), as three passes with an N x C temporary in device memory:
//   1. lin = x * beta + alpha'                                   cat_wide_lin_kernel      (:74-76)
//   2. per row: max, exp, sum -> log-density term, d = onehot(y) - softmax(lin) in place of lin
//                                                                cat_wide_softmax_kernel  (:82-104, :145-160)
//   3. dalpha = colsum(d), dbeta = x' * d                        cat_wide_grad_kernel     (:150-181)
// x is stored column-major on the device (kLayoutPlain: one "tile" holding all rows), every sum runs in a fixed
// order (per-thread strided sums, shuffle trees, per-CTA partials added in CTA order): bit-reproducible like the
// fused kernels.  Traffic: x is read ceil(C / 8) + ceil(C / 8) times (L2 serves most of it at these sizes) -- this is
// the general path, not the roofline path.
#include <algorithm>

#include "glm_device.cuh"

namespace stancuda {

constexpr int kWideThreads = 256;
constexpr int kWideClassChunk = 8;   // classes per thread in passes 1 and 3
constexpr int kWideColChunk = 8;     // columns per CTA in pass 3
constexpr int kWideKStage = 64;      // rows of beta staged in shared memory at a time (pass 1)

// pass 1: thread = one row, kWideClassChunk classes; grid = (row blocks, class chunks)
__global__ void __launch_bounds__(kWideThreads) cat_wide_lin_kernel(const double* __restrict__ x, int64_t ldx, int64_t n,
                                                                    int k, int ncls, const double* __restrict__ params,
                                                                    double* __restrict__ lin) {
  __shared__ double bs[kWideKStage][kWideClassChunk];
  const int64_t row = blockIdx.x * int64_t(kWideThreads) + threadIdx.x;
  const int c0 = blockIdx.y * kWideClassChunk;
  double acc[kWideClassChunk];
#pragma unroll
  for (int c = 0; c < kWideClassChunk; ++c) acc[c] = 0.0;
  for (int k0 = 0; k0 < k; k0 += kWideKStage) {
    const int kn = min(kWideKStage, k - k0);
    __syncthreads();
    for (int q = threadIdx.x; q < kn * kWideClassChunk; q += kWideThreads) {
      const int kk = q / kWideClassChunk, c = q - kk * kWideClassChunk;
      bs[kk][c] = c0 + c < ncls ? params[size_t(c0 + c) * k + k0 + kk] : 0.0;  // beta class-major: [class][k]
    }
    __syncthreads();
    if (row < n) {
      for (int kk = 0; kk < kn; ++kk) {
        const double xv = x[int64_t(k0 + kk) * ldx + row];
#pragma unroll
        for (int c = 0; c < kWideClassChunk; ++c) acc[c] = fma(xv, bs[kk][c], acc[c]);
      }
    }
  }
  if (row < n) {
#pragma unroll
    for (int c = 0; c < kWideClassChunk; ++c)
      if (c0 + c < ncls) lin[int64_t(c0 + c) * n + row] = acc[c] + params[size_t(ncls) * k + c0 + c];
  }
}

// pass 2: thread = one row over all classes (lin is column-major n x ncls: coalesced); per-CTA partial of the
// log-density into lp_part[blockIdx.x]
__global__ void __launch_bounds__(kWideThreads) cat_wide_softmax_kernel(double* __restrict__ lin, const int* __restrict__ y,
                                                                        int64_t n, int ncls, double* __restrict__ lp_part) {
  __shared__ double sh[kWideThreads / 32];
  const int64_t row = blockIdx.x * int64_t(kWideThreads) + threadIdx.x;
  double lp = 0.0;
  if (row < n) {
    double mx = lin[row];
    for (int c = 1; c < ncls; ++c) mx = fmax(mx, lin[int64_t(c) * n + row]);          // :83
    double sum = 0.0;
    for (int c = 0; c < ncls; ++c) sum += exp(lin[int64_t(c) * n + row] - mx);        // :84-85
    const int yc = y[row] - 1;                                                        // 1 <= y <= ncls: checked at fill
    lp = log(1.0 / sum) - mx + lin[int64_t(yc) * n + row];                            // :92-104
    const double inv = 1.0 / sum;
    for (int c = 0; c < ncls; ++c) {
      const double sm = exp(lin[int64_t(c) * n + row] - mx) * inv;
      lin[int64_t(c) * n + row] = (c == yc ? 1.0 : 0.0) - sm;                         // :145-160
    }
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) lp += __shfl_xor_sync(0xffffffffu, lp, off);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = lp;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < kWideThreads / 32; ++w) s += sh[w];
    lp_part[blockIdx.x] = s;
  }
}

// pass 3: CTA = kWideColChunk columns of x (or the intercept when blockIdx.x == column chunks) x kWideClassChunk
// classes, summed over all rows: out[class][col]
__global__ void __launch_bounds__(kWideThreads) cat_wide_grad_kernel(const double* __restrict__ x, int64_t ldx, int64_t n, int k,
                                                                     int ncls, const double* __restrict__ d,
                                                                     double* __restrict__ result) {
  __shared__ double sh[kWideThreads / 32][kWideColChunk * kWideClassChunk];
  const int col_chunks = (k + kWideColChunk - 1) / kWideColChunk;
  const bool intercept = int(blockIdx.x) == col_chunks;
  const int j0 = intercept ? 0 : blockIdx.x * kWideColChunk;
  const int c0 = blockIdx.y * kWideClassChunk;
  double acc[kWideColChunk][kWideClassChunk];
#pragma unroll
  for (int j = 0; j < kWideColChunk; ++j)
#pragma unroll
    for (int c = 0; c < kWideClassChunk; ++c) acc[j][c] = 0.0;
  for (int64_t row = threadIdx.x; row < n; row += kWideThreads) {
    double dv[kWideClassChunk];
#pragma unroll
    for (int c = 0; c < kWideClassChunk; ++c) dv[c] = c0 + c < ncls ? d[int64_t(c0 + c) * n + row] : 0.0;
    if (intercept) {
#pragma unroll
      for (int c = 0; c < kWideClassChunk; ++c) acc[0][c] += dv[c];
    } else {
#pragma unroll
      for (int j = 0; j < kWideColChunk; ++j) {
        const double xv = j0 + j < k ? x[int64_t(j0 + j) * ldx + row] : 0.0;
#pragma unroll
        for (int c = 0; c < kWideClassChunk; ++c) acc[j][c] = fma(xv, dv[c], acc[j][c]);
      }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < kWideColChunk; ++j)
#pragma unroll
    for (int c = 0; c < kWideClassChunk; ++c) {
      double v = acc[j][c];
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0) sh[warp][j * kWideClassChunk + c] = v;
    }
  __syncthreads();
  if (threadIdx.x < kWideColChunk * kWideClassChunk) {
    const int j = threadIdx.x / kWideClassChunk, c = threadIdx.x - j * kWideClassChunk;
    double s = 0.0;
    for (int w = 0; w < kWideThreads / 32; ++w) s += sh[w][threadIdx.x];
    if (c0 + c < ncls) {
      if (intercept) {
        if (j == 0) result[4 + c0 + c] = s;                                          // dalpha
      } else if (j0 + j < k) {
        result[4 + ncls + size_t(c0 + c) * k + j0 + j] = s;                          // dbeta, class-major rows of k
      }
    }
  }
}

// final scalars: lp = sum of the per-CTA partials in CTA order; status word = lp (categorical_logit_glm_lpmf.hpp:109)
__global__ void cat_wide_finish_kernel(const double* __restrict__ lp_part, int nparts, uint32_t flags, double* __restrict__ result) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double s = 0.0;
    for (int b = 0; b < nparts; ++b) s += lp_part[b];
    result[0] = (flags & STAN_CUDA_LP_MAIN) ? s : 0.0;
    result[1] = 0.0;
    result[2] = 0.0;
    result[3] = s;
  }
}

int launch_glm_categorical_wide(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  const Tiling& t = d->t;
  if (t.layout != kLayoutPlain) return fail(STAN_CUDA_ERR_ARG, "launch_glm: the general categorical_logit path needs column-major data");
  const int ncls = io.ncls;
  const int64_t n = d->n;
  const int k = int(d->k);
  if (ncls < 1) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit needs n_classes >= 1");
  if (!io.d_params) return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  if (io.xchg || io.host_flag) return fail(STAN_CUDA_ERR_ARG, "launch_glm: the general categorical_logit path has no in-kernel all-reduce");
  const int64_t row_blocks = (n + kWideThreads - 1) / kWideThreads;
  const int class_chunks = (ncls + kWideClassChunk - 1) / kWideClassChunk;
  if (class_chunks > 65535) return fail(STAN_CUDA_ERR_ARG, "launch_glm: more than 524280 classes");
  // workspace: lin / d (n x ncls) + per-CTA log-density partials
  const size_t lin_doubles = size_t(n) * size_t(ncls);
  SC_CUDA(cudaSetDevice(d->device));
  SC_TRY(workspace_reserve(ws, d->device, lin_doubles + size_t(row_blocks) + 8));
  double* lin = ws->scratch;
  double* lp_part = lin + lin_doubles;
  const int64_t ldx = int64_t(t.R);  // kLayoutPlain: one tile of R = padded row count, column-major
  cat_wide_lin_kernel<<<dim3(unsigned(row_blocks), unsigned(class_chunks)), kWideThreads, 0, stream>>>(d->x, ldx, n, k, ncls,
                                                                                                      io.d_params, lin);
  SC_CUDA(cudaGetLastError());
  cat_wide_softmax_kernel<<<unsigned(row_blocks), kWideThreads, 0, stream>>>(lin, static_cast<const int*>(io.d_y ? io.d_y : d->y), n,
                                                                            ncls, lp_part);
  SC_CUDA(cudaGetLastError());
  if (io.flags & (STAN_CUDA_NEED_DBETA | STAN_CUDA_NEED_DALPHA)) {
    const int col_chunks = (k + kWideColChunk - 1) / kWideColChunk;
    cat_wide_grad_kernel<<<dim3(unsigned(col_chunks + 1), unsigned(class_chunks)), kWideThreads, 0, stream>>>(d->x, ldx, n, k, ncls,
                                                                                                             lin, io.d_result);
    SC_CUDA(cudaGetLastError());
  }
  cat_wide_finish_kernel<<<1, 32, 0, stream>>>(lp_part, int(row_blocks), io.flags, io.d_result);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(4, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
