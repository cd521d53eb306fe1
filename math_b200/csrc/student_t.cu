// student_t.cu -- vectorised student_t_lpdf(y | nu, mu, sigma): value + the partials of all four operands in one
// pass (stan/math/prim/scal/prob/student_t_lpdf.hpp:85-199).  Same shape as normal_lpdf.cu: an elementwise map over
// n with a fixed-order reduction; each operand is a device vector or a scalar passed by value.  Per-element partials
// are written only for vector operands, scalar operands get the sum over n (the reference accumulates them through
// broadcast_array, prim/scal/meta/broadcast_array.hpp:9-33).
//
// The reference precomputes lgamma / digamma of nu/2 and nu/2 + 1/2 once per element of nu (:94-112): for a scalar
// nu that is once per call, and so it is here (on the host, student_t_scalar_dof_terms, with the same libm lgamma
// the reference calls); for a vector nu they are per-element work in both.  HBM-side this is 8 bytes per vector
// operand in and 8 per vector partial out, but the kernel is bound by FP64 issue (B200: ~40 TFLOP/s), not by HBM:
// log1p and the divisions for a scalar nu, the lgamma / digamma pairs on top for a vector nu.
#include <math.h>

#include <cmath>

#include "glm_device.cuh"

namespace stancuda {

struct StudentTKernelParams {
  StudentTIO io;
  double* scratch;       // [grid][8]
  unsigned int* ticket;
};

constexpr int kStSums = 5;  // lp, dy, dnu, dmu, dsigma

template <int THREADS>  // at least (kStSums + 1) warps
__global__ void __launch_bounds__(THREADS) student_t_lpdf_kernel(const StudentTKernelParams p) {
  const StudentTIO& io = p.io;
  const double NEG_LOG_SQRT_PI = -0.5723649429247000870717135;  // prim/scal/fun/constants.hpp:162
  const uint32_t f = io.flags;
  const bool lp_const = f & STAN_CUDA_LP_CONST, lp_dof = f & STAN_CUDA_LP_DOF, lp_scale = f & STAN_CUDA_LP_SCALE,
             lp_main = f & STAN_CUDA_LP_MAIN;
  const bool want_y = f & STAN_CUDA_NEED_DY, want_nu = f & STAN_CUDA_NEED_DNU, want_mu = f & STAN_CUDA_NEED_DALPHA,
             want_sigma = f & STAN_CUDA_NEED_DSIGMA;
  double acc[kStSums] = {0.0, 0.0, 0.0, 0.0, 0.0};
  // domain of the operands, checked here instead of in four O(N) host loops (:67-70)
  unsigned bad = 0;  // bit 0: y NaN, 1: nu not positive finite, 2: mu not finite, 3: sigma not positive finite
  // The loop is FP64-issue-bound, not HBM-bound (a division is ~20 issue slots, log1p ~50): everything that
  // depends only on a scalar operand is hoisted, and the quotients of :145-146, :171-173, :186-188 are written
  // with the two reciprocals 1/sigma and 1/nu (one division each, none for a scalar) and ONE more division,
  // 1/(1 + sq).  d/dy and d/dmu share their factor: the reference's two expressions differ in rounding only.
  const bool nu_vec = io.nu != nullptr, sigma_vec = io.sigma != nullptr;
  const double inv_sigma0 = 1.0 / io.sigma0, inv_nu0 = 1.0 / io.nu0;
  const double log_sigma0 = lp_scale && !sigma_vec ? log(io.sigma0) : 0.0;                               // :125
  for (int64_t i = blockIdx.x * int64_t(THREADS) + threadIdx.x; i < io.n; i += int64_t(gridDim.x) * THREADS) {
    const double y = io.y ? io.y[i] : io.y0;
    const double nu = nu_vec ? io.nu[i] : io.nu0;
    const double mu = io.mu ? io.mu[i] : io.mu0;
    const double sg = sigma_vec ? io.sigma[i] : io.sigma0;
    bad |= (isnan(y) ? 1u : 0u) | ((nu > 0.0 && isfinite(nu)) ? 0u : 2u) | (isfinite(mu) ? 0u : 4u) |
           ((sg > 0.0 && isfinite(sg)) ? 0u : 8u);
    const double inv_sigma = sigma_vec ? 1.0 / sg : inv_sigma0;
    const double inv_nu = nu_vec ? 1.0 / nu : inv_nu0;
    const double hp = 0.5 * nu + 0.5;                                                                    // :90, half_nu + 0.5
    const double diff = y - mu;
    const double zs = diff * inv_sigma;
    const double sq = zs * zs * inv_nu;                                                                  // :145-146
    const double l1p = log1p(sq);                                                                        // :147
    const double inv_1p = 1.0 / (1.0 + sq);
    double lp = 0.0;
    if (lp_const) lp += NEG_LOG_SQRT_PI;                                                                 // :155-157
    if (lp_dof) lp += nu_vec ? lgamma(hp) - lgamma(0.5 * nu) - 0.5 * log(nu) : io.dof_term0;             // :158-160
    if (lp_scale) lp -= sigma_vec ? log(sg) : log_sigma0;                                                // :161-163
    if (lp_main) lp -= hp * l1p;                                                                         // :164-166
    acc[0] += lp;
    if (want_y || want_mu) {                                                                             // :168-173, :183-188
      const double d = hp * inv_1p * (2.0 * diff * (inv_sigma * inv_sigma) * inv_nu);                    // = d/dmu = -d/dy
      if (want_y) {
        acc[1] -= d;
        if (io.dy_row) io.dy_row[i] = -d;
      }
      if (want_mu) {
        acc[3] += d;
        if (io.dmu_row) io.dmu_row[i] = d;
      }
    }
    if (want_nu) {                                                                                       // :174-182
      double c = io.dnu_const0;
      if (nu_vec) c = nu > 0.0 ? 0.5 * digamma_pos(hp) - 0.5 * digamma_pos(0.5 * nu) - 0.5 * inv_nu : 0.0;
      const double d = c - 0.5 * l1p + hp * (inv_1p * sq * inv_nu);
      acc[2] += d;
      if (io.dnu_row) io.dnu_row[i] = d;
    }
    if (want_sigma) {                                                                                    // :189-196
      const double d = -inv_sigma + (nu + 1.0) * inv_1p * (sq * inv_sigma);
      acc[4] += d;
      if (io.dsigma_row) io.dsigma_row[i] = d;
    }
  }
  // block reduction in fixed order
  __shared__ double sh[kStSums][THREADS / 32];
  __shared__ int s_is_last;
  __shared__ unsigned s_bad;
  if (threadIdx.x == 0) s_bad = 0u;
  __syncthreads();
  bad = __reduce_or_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicOr(&s_bad, bad);
#pragma unroll
  for (int q = 0; q < kStSums; ++q) {
    double v = acc[q];
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) sh[q][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < kStSums) {
    double s = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) s += sh[threadIdx.x][w];
    p.scratch[size_t(blockIdx.x) * 8 + threadIdx.x] = s;
  }
  if (threadIdx.x == kStSums) p.scratch[size_t(blockIdx.x) * 8 + kStSums] = double(s_bad);
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
  // (a fixed order: the result is bit-reproducible from call to call); warp kStSums ORs the bad-operand bits
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (warp < unsigned(kStSums)) {
    double s = 0.0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) s += __ldcg(p.scratch + size_t(b) * 8 + warp);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) {
      io.result[warp] = s;
      if (warp == 0) io.result[5] = s;  // finiteness witness = logp
    }
  } else if (warp == unsigned(kStSums)) {
    unsigned all_bad = 0;
    for (unsigned b = lane; b < gridDim.x; b += 32u) all_bad |= unsigned(__ldcg(p.scratch + size_t(b) * 8 + kStSums));
    all_bad = __reduce_or_sync(0xffffffffu, all_bad);
    if (lane == 0) {
      io.result[6] = double(all_bad);
      io.result[7] = 0.0;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.ticket = 0u;
}

void student_t_scalar_dof_terms(StudentTIO* io) {
  io->dof_term0 = 0.0;
  io->dnu_const0 = 0.0;
  const double nu = io->nu0;
  if (io->nu || !(nu > 0.0) || !std::isfinite(nu)) return;  // vector nu: per element in the kernel; bad nu: the kernel reports it
  const double h = 0.5 * nu;
  int sign = 0;  // lgamma_r: no write to the global signgam (the reference's lgamma under STAN_THREADS, prim/scal/fun/lgamma.hpp)
  io->dof_term0 = ::lgamma_r(h + 0.5, &sign) - ::lgamma_r(h, &sign) - 0.5 * std::log(nu);                              // :98-99, :117, :158-160
  io->dnu_const0 = 0.5 * digamma_pos(h + 0.5) - 0.5 * digamma_pos(h) - 0.5 * (1.0 / nu);             // :108-109, :177-180
}

int launch_student_t_lpdf(const StudentTIO& io, Workspace* ws, int device, cudaStream_t stream) {
  if (io.n <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_student_t_lpdf: empty");
  constexpr int THREADS = 256;
  // one element per thread per trip: the per-element work (log1p, up to two lgamma and two digamma, four divisions)
  // is heavy enough that more CTAs beat longer loops until every SM holds its 8
  int64_t want = (io.n + THREADS - 1) / THREADS;
  int64_t cap = int64_t(sm_count(device)) * 8;
  const int grid = int(want < cap ? (want < 1 ? 1 : want) : cap);
  SC_TRY(workspace_reserve(ws, device, size_t(grid) * 8));
  StudentTKernelParams p{io, ws->scratch, ws->ticket};
  student_t_lpdf_kernel<THREADS><<<grid, THREADS, 0, stream>>>(p);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda
