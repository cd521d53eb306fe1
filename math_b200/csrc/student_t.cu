// student_t.cu -- vectorised student_t_lpdf(y | nu, mu, sigma): value + the partials of all four operands in one
// pass (stan/math/prim/scal/prob/student_t_lpdf.hpp:85-199).  Same shape as normal_lpdf.cu: an elementwise map over
// n with a fixed-order reduction; each operand is a device vector or a scalar passed by value.  Per-element partials
// are written only for vector operands, scalar operands get the sum over n (the reference accumulates them through
// broadcast_array, prim/scal/meta/broadcast_array.hpp:9-33).
//
// The reference precomputes lgamma / digamma of nu/2 and nu/2 + 1/2 once per element of nu (:94-112): for a scalar
// nu that is once per call, and so it is here (each thread computes them before its loop); for a vector nu they are
// per-element work in both.  HBM-side this is 8 bytes per vector operand in and 8 per vector partial out; with a
// vector nu the lgamma / digamma pairs make it FP64-issue-bound instead.
#include <math.h>

#include "glm_device.cuh"

namespace stancuda {

struct StudentTKernelParams {
  StudentTIO io;
  double* scratch;       // [grid][8]
  unsigned int* ticket;
};

constexpr int kStSums = 5;  // lp, dy, dnu, dmu, dsigma

template <int THREADS>
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
  // per-nu terms, hoisted for a scalar nu
  double dof_term = 0.0, dnu_const = 0.0, log_sigma0 = 0.0;
  if (!io.nu) {
    const double h = 0.5 * io.nu0;
    if (lp_dof) dof_term = lgamma(h + 0.5) - lgamma(h) - 0.5 * log(io.nu0);                              // :98-99, :117, :158-160
    if (want_nu && io.nu0 > 0.0) dnu_const = 0.5 * digamma_pos(h + 0.5) - 0.5 * digamma_pos(h) - 0.5 * (1.0 / io.nu0);  // :108-109, :177-180
  }
  if (!io.sigma && lp_scale) log_sigma0 = log(io.sigma0);                                                // :125
  for (int64_t i = blockIdx.x * int64_t(THREADS) + threadIdx.x; i < io.n; i += int64_t(gridDim.x) * THREADS) {
    const double y = io.y ? io.y[i] : io.y0;
    const double nu = io.nu ? io.nu[i] : io.nu0;
    const double mu = io.mu ? io.mu[i] : io.mu0;
    const double sg = io.sigma ? io.sigma[i] : io.sigma0;
    bad |= (isnan(y) ? 1u : 0u) | ((nu > 0.0 && isfinite(nu)) ? 0u : 2u) | (isfinite(mu) ? 0u : 4u) |
           ((sg > 0.0 && isfinite(sg)) ? 0u : 8u);
    const double half_nu = 0.5 * nu;                                                                     // :90
    const double zs = (y - mu) / sg;
    const double sq = zs * zs / nu;                                                                      // :145-146
    const double l1p = log1p(sq);                                                                        // :147
    const double inv_1p = 1.0 / (1.0 + sq);
    double lp = 0.0;
    if (lp_const) lp += NEG_LOG_SQRT_PI;                                                                 // :155-157
    if (lp_dof) lp += io.nu ? lgamma(half_nu + 0.5) - lgamma(half_nu) - 0.5 * log(nu) : dof_term;        // :158-160
    if (lp_scale) lp -= io.sigma ? log(sg) : log_sigma0;                                                 // :161-163
    if (lp_main) lp -= (half_nu + 0.5) * l1p;                                                            // :164-166
    acc[0] += lp;
    if (want_y) {                                                                                        // :168-173
      const double d = -(half_nu + 0.5) * 1.0 * inv_1p * (2.0 * (y - mu) / (sg * sg) / nu);
      acc[1] += d;
      if (io.dy_row) io.dy_row[i] = d;
    }
    if (want_nu) {                                                                                       // :174-182
      const double inv_nu = 1.0 / nu;
      double c = dnu_const;
      if (io.nu) c = nu > 0.0 ? 0.5 * digamma_pos(half_nu + 0.5) - 0.5 * digamma_pos(half_nu) - 0.5 * inv_nu : 0.0;
      const double d = c - 0.5 * l1p + (half_nu + 0.5) * (inv_1p * sq * inv_nu);
      acc[2] += d;
      if (io.dnu_row) io.dnu_row[i] = d;
    }
    if (want_mu) {                                                                                       // :183-188
      const double d = -((half_nu + 0.5) * inv_1p * (2.0 * (mu - y) / (sg * sg * nu)));
      acc[3] += d;
      if (io.dmu_row) io.dmu_row[i] = d;
    }
    if (want_sigma) {                                                                                    // :189-196
      const double inv_sigma = 1.0 / sg;
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
  if (threadIdx.x < kStSums) {
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; ++b) s += __ldcg(p.scratch + size_t(b) * 8 + threadIdx.x);
    io.result[threadIdx.x] = s;
    if (threadIdx.x == 0) io.result[5] = s;  // finiteness witness = logp
  }
  if (threadIdx.x == kStSums) {
    unsigned all_bad = 0;
    for (unsigned b = 0; b < gridDim.x; ++b) all_bad |= unsigned(__ldcg(p.scratch + size_t(b) * 8 + kStSums));
    io.result[6] = double(all_bad);
    io.result[7] = 0.0;
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.ticket = 0u;
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
