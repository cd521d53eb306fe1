// normal_lpdf.cu -- vectorised normal_lpdf value + partials in one pass
// (stan/math/prim/scal/prob/normal_lpdf.hpp:78-110).  Elementwise map + fixed-order reduction;
// each operand is either a device vector or a scalar passed by value.  Per-row partials are
// written only for vector operands; scalar operands get the sum (the reference accumulates them
// through broadcast_array, prim/scal/meta/broadcast_array.hpp:9-33).
#include <math.h>

#include "common.cuh"

namespace stancuda {

struct NormalKernelParams {
  NormalIO io;
  double* scratch;       // [grid][4]
  unsigned int* ticket;
};

template <int THREADS>
__global__ void __launch_bounds__(THREADS) normal_lpdf_kernel(const NormalKernelParams p) {
  const NormalIO& io = p.io;
  const double NEG_LOG_SQRT_TWO_PI = -0.91893853320467274178032973640561763986;
  const bool lp_const = io.flags & STAN_CUDA_LP_CONST, lp_scale = io.flags & STAN_CUDA_LP_SCALE,
             lp_main = io.flags & STAN_CUDA_LP_MAIN;
  double a_lp = 0.0, a_sd = 0.0, a_ds = 0.0;  // lp, sum scaled_diff, sum dsigma
  // domain of the operands, checked here instead of in three O(N) host loops
  // (check_not_nan(y), check_finite(mu), check_positive(sigma): normal_lpdf.hpp:51-53)
  unsigned bad = 0;  // bit 0: y is NaN, bit 1: mu not finite, bit 2: sigma not > 0
  const double inv_sigma0 = 1.0 / io.sigma0, log_sigma0 = log(io.sigma0);
  for (int64_t i = blockIdx.x * int64_t(THREADS) + threadIdx.x; i < io.n; i += int64_t(gridDim.x) * THREADS) {
    const double y = io.y_vec ? io.y[i] : io.y0;
    const double mu = io.mu_vec ? io.mu[i] : io.mu0;
    double inv_sigma = inv_sigma0, log_sigma = log_sigma0;
    if (io.sigma_vec) {
      const double sg = io.sigma[i];
      inv_sigma = 1.0 / sg;
      if (lp_scale) log_sigma = log(sg);
    }
    bad |= (isnan(y) ? 1u : 0u) | (isfinite(mu) ? 0u : 2u);
    if (io.sigma_vec) bad |= (io.sigma[i] > 0.0) ? 0u : 4u;
    const double z = (y - mu) * inv_sigma;           // :82-83
    const double z2 = z * z;
    double lp = 0.0;
    if (lp_const) lp += NEG_LOG_SQRT_TWO_PI;         // :89-91
    if (lp_scale) lp -= log_sigma;                   // :92-94
    if (lp_main) lp += -0.5 * z2;                    // :95-97
    a_lp += lp;
    const double scaled_diff = inv_sigma * z;        // :99
    const double ds = -inv_sigma + inv_sigma * z2;   // :108
    a_sd += scaled_diff;
    a_ds += ds;
    if (io.dmu_row) io.dmu_row[i] = scaled_diff;     // :104
    if (io.dy_row) io.dy_row[i] = -scaled_diff;      // :101
    if (io.dsigma_row) io.dsigma_row[i] = ds;
  }
  // block reduction in fixed order
  __shared__ double sh[3][THREADS / 32];
  __shared__ int s_is_last;
  __shared__ unsigned s_bad;
  if (threadIdx.x == 0) s_bad = (io.sigma_vec || io.sigma0 > 0.0) ? 0u : 4u;
  __syncthreads();
  bad = __reduce_or_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicOr(&s_bad, bad);
  double v[3] = {a_lp, a_sd, a_ds};
#pragma unroll
  for (int q = 0; q < 3; ++q) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], off);
    if ((threadIdx.x & 31) == 0) sh[q][threadIdx.x >> 5] = v[q];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double s = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) s += sh[threadIdx.x][w];
    p.scratch[size_t(blockIdx.x) * 4 + threadIdx.x] = s;
  }
  if (threadIdx.x == 3) p.scratch[size_t(blockIdx.x) * 4 + 3] = double(s_bad);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(p.ticket, 1u);
    s_is_last = (t == gridDim.x - 1u);
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  // warp q sums column q of the per-CTA partials: lane l takes CTAs l, l + 32, ... in order, then a shuffle tree
  // (a fixed order: bit-reproducible from call to call); warp 3 ORs the bad-operand bits.  A single thread walking
  // all 1184 partials cost more than the map itself at N = 1e6.
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (warp < 3u) {
    double s = 0.0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) s += __ldcg(p.scratch + size_t(b) * 4 + warp);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) {
      io.result[warp] = s;
      if (warp == 0) io.result[3] = s;  // finiteness witness = logp
    }
  } else if (warp == 3u) {
    unsigned all_bad = 0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) all_bad |= unsigned(__ldcg(p.scratch + size_t(b) * 4 + 3));
    all_bad = __reduce_or_sync(0xffffffffu, all_bad);
    if (lane == 0) io.result[4] = double(all_bad);
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.ticket = 0u;
}

int launch_normal_lpdf(const NormalIO& io, Workspace* ws, int device, cudaStream_t stream) {
  if (io.n <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_normal_lpdf: empty");
  constexpr int THREADS = 256;
  int64_t want = (io.n + THREADS * 4 - 1) / (THREADS * 4);
  int64_t cap = int64_t(sm_count(device)) * 8;
  const int grid = int(want < cap ? (want < 1 ? 1 : want) : cap);
  SC_TRY(workspace_reserve(ws, device, size_t(grid) * 4));
  NormalKernelParams p{io, ws->scratch, ws->ticket};
  normal_lpdf_kernel<THREADS><<<grid, THREADS, 0, stream>>>(p);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
