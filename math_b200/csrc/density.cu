// density.cu -- vectorised elementwise densities with an integer outcome and one real operand: value +
// partials in one pass.  The per-element terms are the row epilogues of the GLM kernels
// with the linear predictor given directly:
//   bernoulli_logit_lpmf(n | theta)   stan/math/prim/scal/prob/bernoulli_logit_lpmf.hpp:51-80
//   poisson_log_lpmf(n | alpha)       stan/math/prim/scal/prob/poisson_log_lpmf.hpp:51-93
//   poisson_lpmf(n | lambda)          stan/math/prim/scal/prob/poisson_lpmf.hpp:45-71
//   bernoulli_lpmf(n | theta)         stan/math/prim/scal/prob/bernoulli_lpmf.hpp:54-107
// Same shape as normal_lpdf.cu: grid-stride map, fixed-order reduction (warp shuffle tree -> warps in order ->
// CTAs in order by the last CTA), each operand a device vector or a scalar passed by value, per-element
// partials written only for a vector operand (a scalar operand gets the sum: what the reference accumulates
// through broadcast_array, prim/scal/meta/broadcast_array.hpp:9-33).  The eager checks of the reference
// (check_bounded / check_nonnegative on n, check_not_nan on the operand) and the LOG_ZERO early returns of
// poisson_log_lpmf are answered by bits the kernel collects on its way through, not by O(N) host loops.
#include <math.h>

#include "common.cuh"

namespace stancuda {

struct DensityKernelParams {
  DensityIO io;
  double* scratch;       // [grid][4]
  unsigned int* ticket;
};

template <int DENSITY, int THREADS>
__global__ void __launch_bounds__(THREADS) density_kernel(const DensityKernelParams p) {
  const DensityIO& io = p.io;
  const bool lp_const = io.flags & STAN_CUDA_LP_CONST, lp_main = io.flags & STAN_CUDA_LP_MAIN;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  double a_lp = 0.0, a_d = 0.0;
  unsigned bits = 0;  // 1: outcome outside the support, 2: operand NaN, 4: LOG_ZERO condition
  for (int64_t i = blockIdx.x * int64_t(THREADS) + threadIdx.x; i < io.n; i += int64_t(gridDim.x) * THREADS) {
    const int y = io.y_vec ? io.y[i] : io.y0;
    const double th = io.theta_vec ? io.theta[i] : io.theta0;
    bits |= (th != th) ? 2u : 0u;
    double lp = 0.0, d = 0.0;
    if (DENSITY == STAN_CUDA_DENSITY_BERNOULLI_LOGIT) {
      bits |= (y < 0 || y > 1) ? 1u : 0u;                                      // :42 check_bounded(n, 0, 1)
      const double sign = double(2 * y - 1);                                   // :58
      const double ntheta = sign * th;
      const double e = exp(-ntheta);                                           // :60
      // cutoff = 20, the three-way select of :63-69 (value) and :72-79 (partial)
      lp = ntheta > 20.0 ? -e : (ntheta < -20.0 ? ntheta : -log1p(e));
      d = ntheta > 20.0 ? -e : (ntheta < -20.0 ? sign : sign * e / (e + 1.0));
      if (!lp_main) lp = 0.0;
    } else if (DENSITY == STAN_CUDA_DENSITY_BERNOULLI) {
      bits |= (y < 0 || y > 1) ? 1u : 0u;                                      // :41 check_bounded(n, 0, 1)
      bits |= (!(fabs(th) <= 1.79769313486231570815e308) || th < 0.0 || th > 1.0) ? 2u : 0u;  // :42-43 finite, in [0, 1]
      if (io.theta_vec) {                                                      // :92-107
        lp = y == 1 ? log(th) : log1p(-th);
        d = y == 1 ? 1.0 / th : 1.0 / (th - 1.0);
        if (!lp_main) lp = 0.0;
      } else {
        d = double(y);  // scalar probability: only the number of successes is needed (:63-91; closed form on the host, api.cu)
      }
    } else if (DENSITY == STAN_CUDA_DENSITY_POISSON) {
      bits |= (y < 0) ? 1u : 0u;                                               // :31 check_nonnegative(n)
      bits |= (th < 0.0) ? 2u : 0u;                                            // :33 check_nonnegative(lambda) (NaN: above)
      bits |= (th == inf || (th == 0.0 && y != 0)) ? 4u : 0u;                  // :45-54 LOG_ZERO
      const double yd = double(y);
      if (!(th == 0.0 && y == 0)) {                                            // :58-66
        if (lp_const && !io.lgamma_cached) lp -= lgamma(yd + 1.0);
        if (lp_main) lp += (y == 0 ? 0.0 : yd * log(th)) - th;                 // multiply_log(n, lambda) - lambda
      }
      d = yd / th - 1.0;                                                       // :69-70
    } else {
      bits |= (y < 0) ? 1u : 0u;                                               // :38 check_nonnegative(n)
      // :51-62: +inf anywhere, or -inf next to a non-zero count -> the whole density is LOG_ZERO
      bits |= (th == inf || (th == -inf && y != 0)) ? 4u : 0u;
      const double yd = double(y);
      const double e = exp(th);                                                // :70-74
      if (!(th == -inf && y == 0)) {                                           // :77-86
        if (lp_const && !io.lgamma_cached) lp -= lgamma(yd + 1.0);
        if (lp_main) lp += yd * th - e;
      }
      d = yd - e;                                                              // :88-90
    }
    a_lp += lp;
    a_d += d;
    if (io.dtheta_row) io.dtheta_row[i] = d;
  }
  // block reduction in fixed order
  __shared__ double sh[2][THREADS / 32];
  __shared__ int s_is_last;
  __shared__ unsigned s_bits;
  if (threadIdx.x == 0) s_bits = 0u;
  __syncthreads();
  bits = __reduce_or_sync(0xffffffffu, bits);
  if ((threadIdx.x & 31) == 0 && bits) atomicOr(&s_bits, bits);
  double v[2] = {a_lp, a_d};
#pragma unroll
  for (int q = 0; q < 2; ++q) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], off);
    if ((threadIdx.x & 31) == 0) sh[q][threadIdx.x >> 5] = v[q];
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    double s = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) s += sh[threadIdx.x][w];
    p.scratch[size_t(blockIdx.x) * 4 + threadIdx.x] = s;
  }
  if (threadIdx.x == 2) p.scratch[size_t(blockIdx.x) * 4 + 2] = double(s_bits);
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
  // (a fixed order: bit-reproducible from call to call); warp 2 ORs the status bits
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (warp < 2u) {
    double s = 0.0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) s += __ldcg(p.scratch + size_t(b) * 4 + warp);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) {
      if (warp == 0 && (DENSITY == STAN_CUDA_DENSITY_POISSON_LOG || DENSITY == STAN_CUDA_DENSITY_POISSON) && lp_const &&
          io.lgamma_cached)
        s -= io.lgamma_total;
      io.result[warp] = s;
      if (warp == 0) {
        io.result[2] = 0.0;
        io.result[3] = s;  // witness = logp
      }
    }
  } else if (warp == 2u) {
    unsigned all = 0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) all |= unsigned(__ldcg(p.scratch + size_t(b) * 4 + 2));
    all = __reduce_or_sync(0xffffffffu, all);
    if (lane == 0) io.result[4] = double(all);
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.ticket = 0u;
}

int launch_density(const DensityIO& io, Workspace* ws, int device, cudaStream_t stream) {
  if (io.n <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_density: empty");
  constexpr int THREADS = 256;
  int64_t want = (io.n + THREADS * 4 - 1) / (THREADS * 4);
  int64_t cap = int64_t(sm_count(device)) * 8;
  const int grid = int(want < cap ? (want < 1 ? 1 : want) : cap);
  SC_TRY(workspace_reserve(ws, device, size_t(grid) * 4));
  DensityKernelParams p{io, ws->scratch, ws->ticket};
  switch (io.density) {
    case STAN_CUDA_DENSITY_BERNOULLI_LOGIT:
      density_kernel<STAN_CUDA_DENSITY_BERNOULLI_LOGIT, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_DENSITY_POISSON_LOG:
      density_kernel<STAN_CUDA_DENSITY_POISSON_LOG, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_DENSITY_POISSON:
      density_kernel<STAN_CUDA_DENSITY_POISSON, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_DENSITY_BERNOULLI:
      density_kernel<STAN_CUDA_DENSITY_BERNOULLI, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    default:
      return fail(STAN_CUDA_ERR_ARG, "launch_density: unknown density");
  }
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
