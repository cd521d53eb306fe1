// glm_ordered_wide.cu -- ordered_logistic_glm_lpmf with more cut-points than the fused kernel holds in shared memory
// (kMaxCuts = 1024).
//
// The reference accepts any number of classes (stan/math/prim/mat/prob/ordered_logistic_glm_lpmf.hpp:69-102 checks
// consistency and the ordering of the cut-points only).  The fused kernel keeps a per-class table and per-warp d/dcuts
// bins in shared memory; beyond that the evaluation runs the way the reference itself does it (:104-200), as three
// passes with two N-vectors in device memory:
//   1. location = x * beta                          the fused kernel as a kFamilyLinearBlock, block by block   (:104-105)
//   2. per row: log-density term, d = dL/dlocation, and the row's two d/dcuts terms into bins in GLOBAL memory that
//      belong to one warp each                      ordered_wide_rows_kernel                                    (:119-200)
//   3. dbeta = x' * d                               the fused kernel again (kFamilyLinearBlock with a given d)  (:175-181)
//   4. warp partials and bins added in warp order, result block assembled      ordered_wide_finish_kernel
// Every block of columns crosses HBM twice (2 x the bytes of the fused path) and every sum runs in a fixed order:
// bit-reproducible like the fused kernels.  This is the general path, not the roofline path.
#include <algorithm>

#include "glm_device.cuh"

namespace stancuda {

constexpr int kOrdWideThreads = 256;
constexpr int kOrdWideWarps = kOrdWideThreads / 32;
constexpr size_t kOrdWideBinDoubles = size_t(32) << 20;  // bins of all warps together (256 MiB): fewer warps beyond

// pass 2.  Warp w owns rows [w * rows_per_warp, (w + 1) * rows_per_warp) and bins[w][ncuts]; lanes take consecutive
// rows (coalesced), 32 rows per step, steps in row order.
__global__ void __launch_bounds__(kOrdWideThreads)
    ordered_wide_rows_kernel(const double* __restrict__ loc, const int* __restrict__ y, int64_t n, int ncuts,
                             const double* __restrict__ cuts, int64_t rows_per_warp, uint32_t flags,
                             double* __restrict__ drow, double* __restrict__ partial, double* __restrict__ bins_all) {
  const int lane = threadIdx.x & 31;
  const int64_t w = int64_t(blockIdx.x) * kOrdWideWarps + (threadIdx.x >> 5);
  const int64_t r0 = w * rows_per_warp;
  const int64_t r1 = r0 + rows_per_warp < n ? r0 + rows_per_warp : n;
  double* bins = bins_all + size_t(w) * ncuts;
  const bool want_dcuts = (flags & STAN_CUDA_NEED_DCUTS) != 0;
  const int ncls = ncuts + 1;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0;
  for (int64_t base = r0; base < r1; base += 32) {
    const int64_t row = base + lane;
    const bool valid = row < r1;
    const int c = valid ? y[row] : 1;  // 1 <= c <= ncuts + 1 was checked from the range of y cached at fill
    const double eta = valid ? loc[row] : 0.0;
    const bool first = c == 1, last = c == ncls, inner = !first && !last;
    const double cy1 = last ? inf : cuts[c - 1];
    const double cy2 = first ? -inf : cuts[c - 2];
    // the class terms the fused kernel tabulates (ordered_class_table), here per row
    const double ecd = inner ? exp(cy2 - cy1) : 0.0;
    const double r = 1.0 / (1.0 - ecd);
    const double l1m = inner ? log1m_exp_dev(cy2 - cy1) : 0.0;
    const RowTerms o = ordered_row_terms(eta, first, last, cy1, cy2, ecd, r, l1m);
    if (valid) {
      a0 += o.t0;
      a1 += o.d;
      a2 += eta;  // ordered_logistic_glm_lpmf.hpp:119: finiteness of sum(location)
      drow[row] = o.d;
    }
    if (want_dcuts) {  // :192-200
      warp_bin_add(bins, lane, valid && !last, c - 1, o.t3);
      warp_bin_add(bins, lane, valid && !first, c - 2, -o.t2);
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, off);
    a1 += __shfl_xor_sync(0xffffffffu, a1, off);
    a2 += __shfl_xor_sync(0xffffffffu, a2, off);
  }
  if (lane == 0) {
    partial[w * 4 + 0] = a0;
    partial[w * 4 + 1] = a1;
    partial[w * 4 + 2] = a2;
  }
}

// parameter block of column block b for the kFamilyLinearBlock launches: [beta_b(kpad_b), 0 (carry starts at zero:
// ordered_logistic has no intercept), 1]; full = [beta(kpad_total), ., ., cuts]
__global__ void ordered_wide_params_kernel(const double* __restrict__ full, double* __restrict__ blocks, int nblocks,
                                           int kpad_last, size_t stride) {
  const int b = blockIdx.x;
  const int kpad_b = b == nblocks - 1 ? kpad_last : kColBlock;
  double* dst = blocks + size_t(b) * stride;
  for (int j = threadIdx.x; j < kpad_b; j += blockDim.x) dst[j] = full[size_t(b) * kColBlock + j];
  if (threadIdx.x == 0) {
    dst[kpad_b] = 0.0;
    dst[kpad_b + 1] = 1.0;
  }
}

// pass 4.  One thread per column of the result block [lp, sum d, 0, witness][dbeta(kpad_total)][dcuts(ncuts)].
// blocks_res[b] = result block of column block b of pass 3: [4][dbeta_b] at stride res_stride (block b holds
// columns [b * kColBlock, ...)).
__global__ void __launch_bounds__(256)
    ordered_wide_finish_kernel(const double* __restrict__ partial, const double* __restrict__ bins_all, int64_t nwarps,
                               int ncuts, const double* __restrict__ blocks_res, size_t res_stride, int kpad_total,
                               uint32_t flags, double* __restrict__ result) {
  const int64_t col = int64_t(blockIdx.x) * 256 + threadIdx.x;
  const int64_t cols = int64_t(4) + kpad_total + ncuts;
  if (col >= cols) return;
  if (col < 4) {
    double s = 0.0;
    if (col != 2) {
      const int src = col == 3 ? 2 : int(col);  // witness = sum(location)
      for (int64_t w = 0; w < nwarps; ++w) s += partial[w * 4 + src];
    }
    if (col == 0 && !(flags & STAN_CUDA_LP_MAIN)) s = 0.0;
    result[col] = s;
  } else if (col < 4 + kpad_total) {
    const int64_t j = col - 4;
    const int64_t b = j / kColBlock;
    result[col] = (flags & STAN_CUDA_NEED_DBETA) ? blocks_res[size_t(b) * res_stride + 4 + (j - b * kColBlock)] : 0.0;
  } else {
    const int64_t j = col - 4 - kpad_total;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    if (flags & STAN_CUDA_NEED_DCUTS) {
      int64_t w = 0;
      for (; w + 4 <= nwarps; w += 4) {
        s0 += bins_all[size_t(w + 0) * ncuts + j];
        s1 += bins_all[size_t(w + 1) * ncuts + j];
        s2 += bins_all[size_t(w + 2) * ncuts + j];
        s3 += bins_all[size_t(w + 3) * ncuts + j];
      }
      for (; w < nwarps; ++w) s0 += bins_all[size_t(w) * ncuts + j];
    }
    result[col] = (s0 + s1) + (s2 + s3);
  }
}

int launch_glm_ordered_wide(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  if (d->family != STAN_CUDA_ORDERED_LOGISTIC) return fail(STAN_CUDA_ERR_STATE, "launch_glm: not an ordered_logistic block");
  if (io.xchg || io.host_flag || io.h_params_inline)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: ordered_logistic with more than 1024 cut-points runs as several launches: no "
                                   "in-kernel all-reduce / inline parameters (use stan_cuda_allreduce behind the launch)");
  if (!io.d_params) return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  if (io.d_alpha_vec || io.d_sigma_vec || io.d_row_dalpha || io.d_row_dsigma)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: ordered_logistic takes no per-row operands");
  const int ncuts = io.ncuts;
  const int64_t n = d->n;
  if (n <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_glm: empty data block");
  // the matrix as a list of column blocks (one block when k <= 512)
  const bool split = !d->colblocks.empty();
  const int B = split ? int(d->colblocks.size()) : 1;
  auto block = [&](int b) -> const stan_cuda_data* { return split ? d->colblocks[size_t(b)] : d; };
  const int kpad_last = block(B - 1)->t.kpad;
  const int kpad_total = (B - 1) * kColBlock + kpad_last;
  const int* y = static_cast<const int*>(io.d_y ? io.d_y : block(B - 1)->y);

  // warps of pass 2: two CTAs per SM unless the bins would outgrow their budget
  const int sms = sm_count(d->device);
  int64_t ctas = int64_t(sms) * 2;
  const int64_t by_rows = (n + int64_t(kOrdWideWarps) * 32 - 1) / (int64_t(kOrdWideWarps) * 32);
  if (ctas > by_rows) ctas = by_rows;
  const int64_t by_bins = int64_t(kOrdWideBinDoubles / size_t(ncuts)) / kOrdWideWarps;
  if (ctas > by_bins) ctas = by_bins < 1 ? 1 : by_bins;
  const int64_t nwarps = ctas * kOrdWideWarps;
  int64_t rows_per_warp = (n + nwarps - 1) / nwarps;
  rows_per_warp = (rows_per_warp + 31) / 32 * 32;

  // one allocation: [partials of the running fused launch | parameter blocks | result blocks | location | d | warp
  // partials | bins]; the sub-launches find their partials region already large enough
  const size_t np_stride = size_t(kColBlock) + 2, nr_stride = size_t(kColBlock) + 4;
  const size_t rows_al = (size_t(n) + 1) & ~size_t(1);
  const size_t partials = size_t(2 * sms + 64) * nr_stride + 2;
  const bool want_dcuts = (io.flags & STAN_CUDA_NEED_DCUTS) != 0;
  const size_t bin_doubles = want_dcuts ? size_t(nwarps) * size_t(ncuts) : 0;
  const size_t extras = size_t(B) * (np_stride + nr_stride) + 2 * rows_al + size_t(nwarps) * 4 + bin_doubles + 8;
  SC_CUDA(cudaSetDevice(d->device));
  SC_TRY(workspace_reserve(ws, d->device, partials + extras + 4));
  double* base = ws->scratch + ((partials + 1) & ~size_t(1));
  double* par = base;
  double* res = par + size_t(B) * np_stride;
  double* loc = res + size_t(B) * nr_stride;
  double* drow = loc + rows_al;
  double* wpart = drow + rows_al;
  double* bins = wpart + size_t(nwarps) * 4;

  // parameter block of every column block (the cut-points stay in io.d_params)
  ordered_wide_params_kernel<<<B, 256, 0, stream>>>(io.d_params, par, B, kpad_last, np_stride);
  SC_CUDA(cudaGetLastError());
  // 1. location = x * beta, block after block into the carry vector
  for (int b = 0; b < B; ++b) {
    LaunchIO sub{0u, par + size_t(b) * np_stride, b == 0 ? nullptr : loc, nullptr,
                 res + size_t(b) * nr_stride, nullptr, loc};
    sub.family_override = kFamilyLinearBlock;
    sub.shared_gpu = io.shared_gpu;
    SC_TRY(launch_glm_warp(block(b), sub, ws, stream));
  }
  // 2. the rows
  if (want_dcuts) SC_CUDA(cudaMemsetAsync(bins, 0, bin_doubles * sizeof(double), stream));
  ordered_wide_rows_kernel<<<unsigned(ctas), kOrdWideThreads, 0, stream>>>(loc, y, n, ncuts, io.d_params + kpad_total + 2,
                                                                          rows_per_warp, io.flags, drow, wpart, bins);
  SC_CUDA(cudaGetLastError());
  // 3. dbeta_b = x_b' * d
  if (io.flags & STAN_CUDA_NEED_DBETA) {
    for (int b = 0; b < B; ++b) {
      LaunchIO sub{0u, par + size_t(b) * np_stride, nullptr, drow, res + size_t(b) * nr_stride, nullptr, nullptr};
      sub.family_override = kFamilyLinearBlock;
      sub.shared_gpu = io.shared_gpu;
      SC_TRY(launch_glm_warp(block(b), sub, ws, stream));
    }
  }
  // 4. the result block
  const int64_t cols = int64_t(4) + kpad_total + ncuts;
  ordered_wide_finish_kernel<<<unsigned((cols + 255) / 256), 256, 0, stream>>>(wpart, bins, nwarps, ncuts, res, nr_stride,
                                                                              kpad_total, io.flags, io.d_result);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(3, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
