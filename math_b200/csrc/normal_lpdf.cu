// normal_lpdf.cu -- vectorised normal_lpdf (and, same operands, cauchy_lpdf / lognormal_lpdf) value + partials in one
// pass (stan/math/prim/scal/prob/normal_lpdf.hpp:78-110).  Elementwise map + fixed-order reduction;
// each operand is either a device vector or a scalar passed by value.  Per-row partials are
// written only for vector operands; scalar operands get the sum (the reference accumulates them
// through broadcast_array, prim/scal/meta/broadcast_array.hpp:9-33).
#include <math.h>

#include "glm_device.cuh"  // digamma_pos

namespace stancuda {

struct NormalKernelParams {
  NormalIO io;
  double* scratch;       // [grid][8]
  unsigned int* ticket;
};

// FAMILY: STAN_CUDA_LOCSCALE_NORMAL | CAUCHY | LOGNORMAL -- the same three operands, the same operand handling, a
// different per-element term:
//   cauchy_lpdf     stan/math/prim/scal/prob/cauchy_lpdf.hpp:60-108
//   lognormal_lpdf  stan/math/prim/scal/prob/lognormal_lpdf.hpp:41-131
//   gamma_lpdf      stan/math/prim/scal/prob/gamma_lpdf.hpp:69-145        (alpha in the mu slot, beta in the sigma slot)
//   exponential_lpdf stan/math/prim/scal/prob/exponential_lpdf.hpp:55-90  (beta in the sigma slot, mu unused)
//   beta_lpdf       stan/math/prim/scal/prob/beta_lpdf.hpp:48-177         (alpha in the mu slot, beta in the sigma slot)
//   weibull_lpdf    stan/math/prim/scal/prob/weibull_lpdf.hpp:38-139      (alpha in the mu slot)
//   inv_gamma_lpdf  stan/math/prim/scal/prob/inv_gamma_lpdf.hpp:58-140    (the slots of gamma_lpdf)
template <int FAMILY, int THREADS>
__global__ void __launch_bounds__(THREADS) normal_lpdf_kernel(const NormalKernelParams p) {
  const NormalIO& io = p.io;
  const double NEG_LOG_SQRT_TWO_PI = -0.91893853320467274178032973640561763986;
  const double NEG_LOG_PI = -1.14472988584940017414342735135;
  const bool lp_const = io.flags & STAN_CUDA_LP_CONST, lp_scale = io.flags & STAN_CUDA_LP_SCALE,
             lp_main = io.flags & STAN_CUDA_LP_MAIN, lp_y = io.flags & STAN_CUDA_LP_Y;
  double a_lp = 0.0, a_sd = 0.0, a_ds = 0.0, a_dy = 0.0;  // lp, sum dmu, sum dsigma, sum dy
  // domain of the operands, checked here instead of in three O(N) host loops
  // (check_not_nan(y), check_finite(mu), check_positive(sigma): normal_lpdf.hpp:51-53; cauchy / lognormal:
  // check_positive_finite(sigma), lognormal also check_nonnegative(y))
  unsigned bad = 0;  // bit 0: y is NaN (lognormal: or negative), bit 1: mu not finite, bit 2: sigma outside its domain,
                     // bit 3: lognormal with an outcome equal to zero (LOG_ZERO, lognormal_lpdf.hpp:41-45)
  const double inv_sigma0 = 1.0 / io.sigma0, log_sigma0 = log(io.sigma0);
  for (int64_t i = blockIdx.x * int64_t(THREADS) + threadIdx.x; i < io.n; i += int64_t(gridDim.x) * THREADS) {
    const double y = io.y_vec ? io.y[i] : io.y0;
    const double mu = io.mu_vec ? io.mu[i] : io.mu0;
    double sg = io.sigma0, inv_sigma = inv_sigma0, log_sigma = log_sigma0;
    if (io.sigma_vec) {
      sg = io.sigma[i];
      inv_sigma = 1.0 / sg;
      if (lp_scale) log_sigma = log(sg);
    }
    bad |= isnan(y) ? 1u : 0u;
    if (FAMILY == STAN_CUDA_LOCSCALE_GAMMA || FAMILY == STAN_CUDA_LOCSCALE_BETA || FAMILY == STAN_CUDA_LOCSCALE_WEIBULL ||
        FAMILY == STAN_CUDA_LOCSCALE_INV_GAMMA)
      bad |= (mu > 0.0 && isfinite(mu)) ? 0u : 2u;  // the shape: positive and finite
    else if (FAMILY != STAN_CUDA_LOCSCALE_EXPONENTIAL) bad |= isfinite(mu) ? 0u : 2u;
    if (io.sigma_vec) bad |= (sg > 0.0 && (FAMILY == STAN_CUDA_LOCSCALE_NORMAL || isfinite(sg))) ? 0u : 4u;
    double lp = 0.0, d_mu, d_sigma, d_y;
    if (FAMILY == STAN_CUDA_LOCSCALE_NORMAL) {
      const double z = (y - mu) * inv_sigma;           // :82-83
      const double z2 = z * z;
      if (lp_const) lp += NEG_LOG_SQRT_TWO_PI;         // :89-91
      if (lp_scale) lp -= log_sigma;                   // :92-94
      if (lp_main) lp += -0.5 * z2;                    // :95-97
      d_mu = inv_sigma * z;                            // :99, :104
      d_sigma = -inv_sigma + inv_sigma * z2;           // :108
      d_y = -d_mu;                                     // :101
    } else if (FAMILY == STAN_CUDA_LOCSCALE_CAUCHY) {
      const double sigma_squared = sg * sg;            // :69
      const double y_minus_mu = y - mu;                // :81-85
      const double y_minus_mu_squared = y_minus_mu * y_minus_mu;
      const double z = y_minus_mu * inv_sigma;
      if (lp_const) lp += NEG_LOG_PI;                  // :87-89
      if (lp_scale) lp -= log_sigma;                   // :90-92
      if (lp_main) lp -= log1p(z * z);                 // :93-95
      d_mu = 2.0 * y_minus_mu / (sigma_squared + y_minus_mu_squared);                                   // :101-104
      d_y = -d_mu;                                                                                      // :97-100
      d_sigma = (y_minus_mu_squared - sigma_squared) * inv_sigma / (sigma_squared + y_minus_mu_squared);  // :105-109
    } else if (FAMILY == STAN_CUDA_LOCSCALE_GAMMA) {
      // gamma_lpdf.hpp:69-145 with alpha = mu, beta = sg; the four summands have a flag each (include/stan_cuda.h)
      bad |= (y < 0.0) ? 8u : 0u;                      // :69-74 LOG_ZERO
      const double log_y = y > 0.0 ? log(y) : 0.0;     // :86-90 (an outcome of zero leaves the builder's zero)
      const double log_beta = log(sg);                 // :109-113
      if (io.flags & STAN_CUDA_LP_DOF) lp -= lgamma(mu);               // :118-120
      if (lp_scale) lp += mu * log_beta;                                 // :121-123
      if (lp_y) lp += (mu - 1.0) * log_y;                                // :124-126
      if (lp_main) lp -= sg * y;                                         // :127-129
      d_y = (mu - 1.0) / y - sg;                                         // :131-133
      d_mu = (io.flags & STAN_CUDA_NEED_DALPHA) ? -digamma_pos(mu) + log_beta + log_y : 0.0;  // :134-137
      d_sigma = mu / sg - y;                                             // :138-140
    } else if (FAMILY == STAN_CUDA_LOCSCALE_INV_GAMMA) {
      // inv_gamma_lpdf.hpp:58-140 with alpha = mu, beta = sg
      bad |= (y <= 0.0) ? 8u : 0u;                     // :58-63 LOG_ZERO
      const double log_y = y > 0.0 ? log(y) : 0.0, inv_y = 1.0 / y;      // :76-86
      const double log_beta = log(sg);                 // :106-110
      if (io.flags & STAN_CUDA_LP_DOF) lp -= lgamma(mu);               // :115-117
      if (lp_scale) lp += mu * log_beta;                                 // :118-120
      if (lp_y) lp -= (mu + 1.0) * log_y;                                // :121-123
      if (lp_main) lp -= sg * inv_y;                                     // :124-126
      d_y = -(mu + 1.0) * inv_y + sg * inv_y * inv_y;                    // :128-131
      d_mu = (io.flags & STAN_CUDA_NEED_DALPHA) ? -digamma_pos(mu) + log_beta - log_y : 0.0;  // :132-135
      d_sigma = mu / sg - inv_y;                                         // :136-138
    } else if (FAMILY == STAN_CUDA_LOCSCALE_BETA) {
      // beta_lpdf.hpp:48-176 with alpha = mu, beta = sg; five summands, a flag each (include/stan_cuda.h)
      bad |= (y < 0.0 || y > 1.0) ? 1u : 0u;           // :53-54 check_nonnegative / check_less_or_equal(y, 1)
      const double log_y = log(y), log1m_y = log1p(-y);                  // :87-94
      const bool want_dig = (io.flags & (STAN_CUDA_NEED_DALPHA | STAN_CUDA_NEED_DSIGMA)) != 0;
      const double dig_ab = want_dig ? digamma_pos(mu + sg) : 0.0;       // :131-141
      if (lp_scale) lp += lgamma(mu + sg);                               // :150-152
      if (io.flags & STAN_CUDA_LP_DOF) lp -= lgamma(mu);                 // :153-155
      if (io.flags & STAN_CUDA_LP_LINEAR) lp -= lgamma(sg);              // :156-158
      if (lp_y) lp += (mu - 1.0) * log_y;                                // :159-161
      if (lp_main) lp += (sg - 1.0) * log1m_y;                           // :162-164
      d_y = (mu - 1.0) / y + (sg - 1.0) / (y - 1.0);                     // :166-169
      d_mu = (io.flags & STAN_CUDA_NEED_DALPHA) ? log_y + dig_ab - digamma_pos(mu) : 0.0;      // :170-173
      d_sigma = (io.flags & STAN_CUDA_NEED_DSIGMA) ? log1m_y + dig_ab - digamma_pos(sg) : 0.0;  // :174-177
    } else if (FAMILY == STAN_CUDA_LOCSCALE_WEIBULL) {
      // weibull_lpdf.hpp:38-139 with alpha = mu, sigma = sg
      bad |= isfinite(y) ? 0u : 1u;                    // :38 check_finite(y)
      bad |= (y < 0.0) ? 8u : 0u;                      // :56-61 LOG_ZERO
      const double log_y = log(y), log_sg = log(sg), inv_y = 1.0 / y;
      const double pw = pow(y * inv_sigma, mu);        // :95-103
      if (io.flags & STAN_CUDA_LP_DOF) lp += log(mu);                    // :111-113
      if (lp_y) lp += (mu - 1.0) * log_y;                                // :114-116
      if (lp_scale) lp -= mu * log_sg;                                   // :117-119
      if (lp_main) lp -= pw;                                             // :120-122
      d_y = (mu - 1.0) * inv_y - mu * pw * inv_y;                        // :125-130
      d_mu = 1.0 / mu + (1.0 - pw) * (log_y - log_sg);                   // :131-135
      d_sigma = mu * inv_sigma * (pw - 1.0);                             // :136-139
    } else if (FAMILY == STAN_CUDA_LOCSCALE_EXPONENTIAL) {
      bad |= (y < 0.0) ? 1u : 0u;                      // exponential_lpdf.hpp:55 check_nonnegative(y)
      if (lp_scale) lp += log(sg);                     // :76-78
      if (lp_main) lp -= sg * y;                       // :79-81
      d_y = -sg;                                       // :83-85
      d_mu = 0.0;
      d_sigma = 1.0 / sg - y;                          // :86-88
    } else {
      bad |= (y < 0.0) ? 1u : 0u;                      // :25 check_nonnegative(y)
      bad |= (y == 0.0) ? 8u : 0u;                     // :41-45
      const double log_y = log(y), inv_y = 1.0 / y;    // :78-90
      const double inv_sigma_sq = inv_sigma * inv_sigma;
      const double logy_m_mu = log_y - mu;             // :99-102
      const double logy_m_mu_div_sigma = logy_m_mu * inv_sigma_sq;  // :105-108
      if (lp_const) lp += NEG_LOG_SQRT_TWO_PI;         // :92-94 (N times the constant)
      if (lp_scale) lp -= log_sigma;                   // :110-112
      if (lp_y) lp -= log_y;                           // :113-115
      if (lp_main) lp -= 0.5 * (logy_m_mu * logy_m_mu) * inv_sigma_sq;  // :116-118
      d_y = -(1.0 + logy_m_mu_div_sigma) * inv_y;      // :120-122
      d_mu = logy_m_mu_div_sigma;                      // :123-125
      d_sigma = (logy_m_mu_div_sigma * logy_m_mu - 1.0) * inv_sigma;  // :126-129
    }
    a_lp += lp;
    a_sd += d_mu;
    a_ds += d_sigma;
    a_dy += d_y;
    if (io.dmu_row) io.dmu_row[i] = d_mu;
    if (io.dy_row) io.dy_row[i] = d_y;
    if (io.dsigma_row) io.dsigma_row[i] = d_sigma;
  }
  // block reduction in fixed order
  __shared__ double sh[4][THREADS / 32];
  __shared__ int s_is_last;
  __shared__ unsigned s_bad;
  if (threadIdx.x == 0)
    s_bad = (io.sigma_vec || (io.sigma0 > 0.0 && (FAMILY == STAN_CUDA_LOCSCALE_NORMAL || isfinite(io.sigma0)))) ? 0u : 4u;
  __syncthreads();
  bad = __reduce_or_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicOr(&s_bad, bad);
  double v[4] = {a_lp, a_sd, a_ds, a_dy};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], off);
    if ((threadIdx.x & 31) == 0) sh[q][threadIdx.x >> 5] = v[q];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double s = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) s += sh[threadIdx.x][w];
    p.scratch[size_t(blockIdx.x) * 8 + threadIdx.x] = s;
  }
  if (threadIdx.x == 4) p.scratch[size_t(blockIdx.x) * 8 + 4] = double(s_bad);
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
  // (a fixed order: bit-reproducible from call to call); warp 4 ORs the status bits.  A single thread walking
  // all 1184 partials cost more than the map itself at N = 1e6.
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (warp < 4u) {
    double s = 0.0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) s += __ldcg(p.scratch + size_t(b) * 8 + warp);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) {
      io.result[warp == 3u ? 5 : warp] = s;  // [0] lp, [1] sum dmu, [2] sum dsigma, [5] sum dy
      if (warp == 0) io.result[3] = s;       // finiteness witness = logp
    }
  } else if (warp == 4u) {
    unsigned all_bad = 0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) all_bad |= unsigned(__ldcg(p.scratch + size_t(b) * 8 + 4));
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
  SC_TRY(workspace_reserve(ws, device, size_t(grid) * 8));
  NormalKernelParams p{io, ws->scratch, ws->ticket};
  switch (io.family) {
    case STAN_CUDA_LOCSCALE_NORMAL:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_NORMAL, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_CAUCHY:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_CAUCHY, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_LOGNORMAL:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_LOGNORMAL, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_GAMMA:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_GAMMA, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_EXPONENTIAL:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_EXPONENTIAL, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_BETA:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_BETA, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_WEIBULL:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_WEIBULL, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    case STAN_CUDA_LOCSCALE_INV_GAMMA:
      normal_lpdf_kernel<STAN_CUDA_LOCSCALE_INV_GAMMA, THREADS><<<grid, THREADS, 0, stream>>>(p);
      break;
    default:
      return fail(STAN_CUDA_ERR_ARG, "launch_normal_lpdf: unknown density");
  }
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
