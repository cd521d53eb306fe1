// glm_fused.cu -- the fused GLM value+gradient kernel for sm_100a (FP64, HBM-bound).
//
// One pass over X computes, for a block of rows,
//     eta = X*beta + alpha,  the per-row log-likelihood and d = dL/deta,  and  g = X^T d
// i.e. the whole numeric body of
//     bernoulli_logit_glm_lpmf  stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp:88-135
//     poisson_log_glm_lpmf      stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp:91-129
//     normal_id_glm_lpdf        stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp:100-172
// which the reference does with two Eigen GEMVs (two reads of X) plus N-length temporaries, and
// its OpenCL backend with two kernels (opencl/kernels/*_glm_*.hpp + row_vector_matrix_multiply,
// opencl/kernels/matrix_multiply.hpp:354-386) and a round trip of d through global memory.
//
// Structure:
//  * persistent CTAs, static round-robin tile schedule (deterministic sums);
//  * a tile = R rows x kpad columns, contiguous in HBM, staged into shared memory by ONE 1-D TMA
//    bulk copy (cp.async.bulk ... mbarrier::complete_tx) per tile into a ring of `stages` buffers;
//  * thread (cg, r) pulls its 16 doubles of row r from shared memory into registers ONCE, uses
//    them for the partial dot product (eta) and again for the gradient FMA after d is known:
//    X crosses HBM once and shared memory once;
//  * partial dots of the C column groups meet in shared memory, the row owner (cg == 0) runs the
//    family epilogue, d goes back through shared memory;
//  * per-thread gradient accumulators live in registers for the whole kernel; at the end a
//    fixed-order warp-shuffle + block reduction writes one partial block per CTA, and the last
//    CTA to finish (atomic ticket) adds the CTA partials in CTA order -> bit-reproducible.
#include <math.h>

#include "glm_device.cuh"

namespace stancuda {

// The CTA-synchronous kernel was superseded by glm_warp.cu in round 1 and is kept for A/B measurements only: it is
// compiled into the library when the build defines STAN_CUDA_WITH_CTA_KERNEL (math_b200/build.py, environment
// STAN_CUDA_WITH_CTA_KERNEL=1); the product library carries the warp kernel alone.
#ifdef STAN_CUDA_WITH_CTA_KERNEL
// Sum `v` over the `width` consecutive lanes that share a column group (width = min(R, 32)).
__device__ __forceinline__ double group_reduce(double v, int width) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const double o = __shfl_xor_sync(0xffffffffu, v, off);
    if (off < width) v += o;
  }
  return v;
}

template <int FAMILY, int THREADS, bool ROWIO>
__global__ void __launch_bounds__(THREADS, (THREADS >= 512 ? 1 : (THREADS >= 256 ? 2 : 4)))
glm_fused_kernel(const GlmKernelParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int WARPS = THREADS / 32;
  const int tid = threadIdx.x;
  const int R = p.R, C = p.C, kc = p.kc, kpad = p.kpad;
  const int r = tid & (R - 1);
  const int cg = tid >> p.log2R;
  const bool owner = cg == 0;
  const bool multi = C > 1;

  // ---- shared memory carve-up -------------------------------------------------------------
  unsigned char* stage0 = smem;
  double* part = reinterpret_cast<double*>(smem + size_t(p.stages) * p.stage_bytes);  // [THREADS]
  double* dsm = part + THREADS;                                                       // [THREADS]
  double* red = dsm + THREADS;                             // [2 * WARPS][kMaxKc + 4] leader slots
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 2 * WARPS * (kMaxKc + 4));       // [stages]
  __shared__ int s_is_last;

  const int grid = gridDim.x;
  const int ntl = (p.tiles - int(blockIdx.x) + grid - 1) / grid;  // tiles of this CTA
  const uint32_t tx_bytes = p.tile_bytes + p.y_bytes;
  const unsigned char* xbytes = reinterpret_cast<const unsigned char*>(p.x);
  const unsigned char* ybytes = reinterpret_cast<const unsigned char*>(p.y);
  uint64_t policy = 0;

  auto issue = [&](int it) {  // thread 0 only: stage tile `it` of this CTA
    const int s = it % p.stages;
    const int64_t tile = int64_t(blockIdx.x) + int64_t(it) * grid;
    const uint32_t bar = smem_u32(&bars[s]);
    const uint32_t dst = smem_u32(stage0 + size_t(s) * p.stage_bytes);
    mbar_expect_tx(bar, tx_bytes);
    tma_bulk_g2s(dst, xbytes + tile * p.tile_bytes, p.tile_bytes, bar, policy);
    tma_bulk_g2s(dst + p.tile_bytes, ybytes + tile * p.y_bytes, p.y_bytes, bar, policy);
  };

  PHASE_STAMP(0);
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) mbar_init(smem_u32(&bars[s]), 1);
    mbar_fence_init();
    policy = l2_evict_first_policy();
  }
  __syncthreads();
  if (tid == 0) {
    const int pre = ntl < p.stages ? ntl : p.stages;
    for (int it = 0; it < pre; ++it) issue(it);
  }

  // ---- per-thread constants ---------------------------------------------------------------
  double bj[kMaxKc], acc[kMaxKc];
#pragma unroll
  for (int j = 0; j < kMaxKc; ++j) {
    bj[j] = j < kc ? kernel_param(p, cg * kc + j) : 0.0;
    acc[j] = 0.0;
  }
  const double alpha0 = kernel_param(p, kpad);
  const double sigma0 = kernel_param(p, kpad + 1);
  const double inv_sigma0 = FAMILY == STAN_CUDA_NORMAL_ID ? 1.0 / sigma0 : 1.0;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;  // owner accumulators: lp main, sum d, sum s^2, sum log sigma

  // ---- main loop over this CTA's tiles ------------------------------------------------------
  for (int it = 0; it < ntl; ++it) {
    const int s = it % p.stages;
    const uint32_t parity = (it / p.stages) & 1;
    const int64_t tile = int64_t(blockIdx.x) + int64_t(it) * grid;
    const int64_t row = tile * R + r;
    mbar_wait(smem_u32(&bars[s]), parity);
#ifdef STAN_CUDA_PHASE_TIMERS
    if (it == 0) PHASE_STAMP(1);
#endif

    const unsigned char* st = stage0 + size_t(s) * p.stage_bytes;
    const double* xs = reinterpret_cast<const double*>(st) + (cg * kc) * R + r;
    double xr[kMaxKc];
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j) xr[j] = j < kc ? xs[j * R] : 0.0;

    double yv = 0.0, alpha = alpha0, inv_sigma = inv_sigma0, sigma_row = sigma0;
    if (owner) {
      if (FAMILY == STAN_CUDA_NORMAL_ID) {
        yv = reinterpret_cast<const double*>(st + p.tile_bytes)[r];
      } else {
        yv = double(reinterpret_cast<const int*>(st + p.tile_bytes)[r]);
      }
      if (ROWIO && row < p.n) {
        if (p.alpha_vec) alpha = p.alpha_vec[row];
        if (FAMILY == STAN_CUDA_NORMAL_ID && p.sigma_vec) sigma_row = p.sigma_vec[row];
      }
    }

    double p0 = 0.0, p1 = 0.0;
#pragma unroll
    for (int j = 0; j < kMaxKc; j += 2) {
      p0 = fma(xr[j], bj[j], p0);
      p1 = fma(xr[j + 1], bj[j + 1], p1);
    }
    double eta = p0 + p1;
    if (multi) part[tid] = eta;
    __syncthreads();  // every thread has drained stage s; partial dots are visible

    if (tid == 0 && it + p.stages < ntl) issue(it + p.stages);

    double d = 0.0;
    if (owner) {
      for (int c = 1; c < C; ++c) eta += part[c * R + r];  // fixed order over column groups
      if (ROWIO && FAMILY == STAN_CUDA_NORMAL_ID && p.sigma_vec) inv_sigma = 1.0 / sigma_row;
      RowTerms o = row_epilogue<FAMILY>(eta, alpha, yv, inv_sigma);
      if (row < p.n) {
        d = o.d;
        a0 += o.t0;
        a1 += o.d;
        a2 += o.t2;
        if (ROWIO) {
          if (FAMILY == STAN_CUDA_NORMAL_ID && p.sigma_vec) {
            a3 += log(sigma_row);
            if (p.row_dsigma) p.row_dsigma[row] = (o.t2 - 1.0) * inv_sigma;
          }
          if (p.row_dalpha) p.row_dalpha[row] = o.d;
        }
      }
      if (multi) dsm[r] = d;
    }
    if (multi) {
      __syncthreads();
      d = dsm[r];
    }
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j) acc[j] = fma(xr[j], d, acc[j]);
  }

  PHASE_STAMP(2);
  // ---- CTA reduction: lanes of a column group -> warps of a column group -> CTA partial -----
  const int width = R < 32 ? R : 32;
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int j = 0; j < kMaxKc; ++j) acc[j] = group_reduce(acc[j], width);
  a0 = group_reduce(a0, width);
  a1 = group_reduce(a1, width);
  a2 = group_reduce(a2, width);
  a3 = group_reduce(a3, width);
  __syncthreads();
  // slot (warp, sub) -> one leader lane per column group inside the warp
  if ((lane & (width - 1)) == 0) {
    // red is laid out [THREADS / width][kMaxKc + 4]
    double* dst = red + size_t(tid / width) * (kMaxKc + 4);
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j) dst[j] = acc[j];
    dst[kMaxKc + 0] = a0;
    dst[kMaxKc + 1] = a1;
    dst[kMaxKc + 2] = a2;
    dst[kMaxKc + 3] = a3;
  }
  __syncthreads();
  const int cols = kpad + 4;
  double* my_scratch = p.scratch + size_t(blockIdx.x) * cols;
  const int slots_per_cg = R / width;  // leader slots per column group (>= 1)
  for (int c = tid; c < cols; c += THREADS) {
    double sum = 0.0;
    if (c < 4) {  // scalars live in column group 0
      for (int q = 0; q < slots_per_cg; ++q) sum += red[size_t(q) * (kMaxKc + 4) + kMaxKc + c];
    } else {
      const int col = c - 4, g = col / kc, j = col - g * kc;
      const int first = g * slots_per_cg;  // leader slots are ordered by tid, i.e. by (cg, r)
      for (int q = 0; q < slots_per_cg; ++q) sum += red[size_t(first + q) * (kMaxKc + 4) + j];
    }
    my_scratch[c] = sum;
  }

  // ---- two-level fixed-order reduction of the CTA partials, finalisation, all-reduce ----------
  double* fin = reinterpret_cast<double*>(stage0);  // [cols]; every stage has been consumed by now
  finish_reduction<FAMILY, THREADS>(p, tid, cols, part, fin, &s_is_last, [] { __syncthreads(); });
}

#endif  // STAN_CUDA_WITH_CTA_KERNEL

}  // namespace stancuda

// =============================================================================================
// host side: tiling choice and launch
// =============================================================================================
namespace stancuda {

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  return atoi(v);
}

static Tuning& tuning_mut() {
  static Tuning t = [] {
    Tuning u;
    u.threads = env_int("STAN_CUDA_THREADS", 0);
    u.stages = env_int("STAN_CUDA_STAGES", 0);
    u.ctas_per_sm = env_int("STAN_CUDA_CTAS_PER_SM", 0);
    const char* sv = getenv("STAN_CUDA_SCHEDULE");
    u.blocked = (sv && (sv[0] == 'b' || sv[0] == 'B')) ? 1 : 0;
    const char* kv = getenv("STAN_CUDA_KERNEL");
#ifdef STAN_CUDA_WITH_CTA_KERNEL
    u.layout = (kv && (kv[0] == 'c' || kv[0] == 'C')) ? kLayoutCta : kLayoutWarp;
#else
    (void)kv;
    u.layout = kLayoutWarp;
#endif
    const char* cv = getenv("STAN_CUDA_CATEGORICAL");
    u.cat_mma = (cv && (cv[0] == 'w' || cv[0] == 'W' || cv[0] == 'v' || cv[0] == 'V')) ? 0 : 1;
    return u;
  }();
  return t;
}

const Tuning& tuning() { return tuning_mut(); }

int sm_count(int device) {
  static int cached[64] = {0};
  if (device < 0 || device >= 64) return 148;
  if (!cached[device]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0) v = 148;
    cached[device] = v;
  }
  return cached[device];
}

bool choose_tiling(int64_t n, int64_t k, int threads_override, Tiling* t, int force_layout) {
  if (k < 1) k = 1;
  const int layout = force_layout >= 0 ? force_layout : tuning().layout;
  if (layout == kLayoutPlain) {
    t->layout = kLayoutPlain;
    t->threads = 256;
    t->C = 1;
    t->kc = int(k);
    t->kpad = int(k);
    t->R = int((n + 1) & ~int64_t(1));
    t->tiles = 1;
    return true;
  }
  if (layout == kLayoutMma) {
    // fragments of 8 rows x 4 columns (glm_categorical_mma.cu); tiles of at most 32 KiB up to 64 columns
    if (k > 256) return false;
    t->layout = kLayoutMma;
    t->threads = 256;
    t->kpad = int((k + 7) / 8) * 8;
    t->C = 1;
    t->kc = t->kpad;
    t->R = t->kpad <= 64 ? 64 : 32;
    t->tiles = (n + t->R - 1) / t->R;
    return true;
  }
  if (layout == kLayoutWarp) {
    // 8 consumer warps of 32 lanes; smallest power-of-two C <= 32 with at most kMaxKc columns per lane
    int C = 1;
    while ((k + C - 1) / C > kMaxKc) C *= 2;
    if (C > 32) return false;  // k > 512: outside the fused kernel
    t->layout = kLayoutWarp;
    t->threads = 256;
    t->C = C;
    t->kc = int((k + C - 1) / C);
    t->R = 256 / C;
    t->kpad = C * t->kc;
    t->tiles = (n + t->R - 1) / t->R;
    return true;
  }
  t->layout = kLayoutCta;
  int threads = threads_override ? threads_override : tuning().threads;
  if (threads != 128 && threads != 256 && threads != 512) threads = 256;
  if (k < 1) k = 1;
  // smallest power-of-two number of column groups with at most kMaxKc columns per thread
  int C = 1;
  while ((k + C - 1) / C > kMaxKc) C *= 2;
  while (threads / C < 16 && threads < 512) threads *= 2;  // keep >= 16 rows per tile
  if (threads / C < 16) return false;                     // k > 512: outside the fused kernel
  t->threads = threads;
  t->C = C;
  t->kc = int((k + C - 1) / C);
  t->R = threads / C;
  t->kpad = C * t->kc;
  t->tiles = (n + t->R - 1) / t->R;
  return true;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef STAN_CUDA_WITH_CTA_KERNEL
template <int FAMILY, int THREADS, bool ROWIO>
static int launch_inst(const GlmKernelParams& p, int grid, size_t smem, cudaStream_t stream) {
  auto kern = glm_fused_kernel<FAMILY, THREADS, ROWIO>;
  static std::atomic<size_t> configured[64];
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && configured[dev].load(std::memory_order_relaxed) < smem) {
    SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev].store(smem, std::memory_order_relaxed);
  }
  kern<<<grid, THREADS, smem, stream>>>(p);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

template <int FAMILY, bool ROWIO>
static int launch_threads(int threads, const GlmKernelParams& p, int grid, size_t smem, cudaStream_t s) {
  switch (threads) {
    case 128: return launch_inst<FAMILY, 128, ROWIO>(p, grid, smem, s);
    case 256: return launch_inst<FAMILY, 256, ROWIO>(p, grid, smem, s);
    default: return launch_inst<FAMILY, 512, ROWIO>(p, grid, smem, s);
  }
}

template <int FAMILY>
static int launch_family(bool rowio, int threads, const GlmKernelParams& p, int grid, size_t smem,
                         cudaStream_t s) {
  return rowio ? launch_threads<FAMILY, true>(threads, p, grid, smem, s)
               : launch_threads<FAMILY, false>(threads, p, grid, smem, s);
}

// Shared-memory footprint of one CTA with `stages` stages.
static size_t smem_bytes(const Tiling& t, int stages, size_t* stage_bytes_out) {
  const size_t stage = align_up(t.tile_bytes() + size_t(t.R) * sizeof(double), 128);
  if (stage_bytes_out) *stage_bytes_out = stage;
  const size_t fixed = (2 * size_t(t.threads) + 2 * size_t(t.threads / 32) * (kMaxKc + 4)) * sizeof(double)
                       + size_t(stages) * sizeof(uint64_t);
  return size_t(stages) * stage + fixed;
}

#endif  // STAN_CUDA_WITH_CTA_KERNEL

#ifdef STAN_CUDA_PHASE_TIMERS
static unsigned long long* g_phase_timers = nullptr;  // debugging build only (tools/phase_timers.py)
#endif

// Everything of the kernel parameter block that does not depend on the kernel variant.
int fill_kernel_params(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, GlmKernelParams* p) {
  const Tiling& t = d->t;
  p->stages = 0;
  p->group_size = 1;
  p->stage_bytes = 0;
  p->blocked = 0;
  p->nb_sums = nullptr;
  p->nb_rowfun = 1;
  const int family = io.family_override >= 0 ? io.family_override : d->family;
  p->ncuts = family == STAN_CUDA_ORDERED_LOGISTIC ? io.ncuts : 0;
  p->ord_private = 0;
  p->pair = 0;
  p->pdl = 0;
  p->ncls = 0;
  p->cat_groups = 1;
  if (p->ncuts < 0 || p->ncuts > kMaxCuts) return fail(STAN_CUDA_ERR_ARG, "launch_glm: ordered_logistic supports at most 1024 cut-points on the device");
  p->x = d->x;
  p->y = io.d_y ? io.d_y : d->y;
  p->n = d->n;
  p->tiles = int(t.tiles);
  p->R = t.R; p->C = t.C; p->kc = t.kc; p->kpad = t.kpad;
  p->log2R = 0;
  while ((1 << p->log2R) < t.R) ++p->log2R;
  p->tile_bytes = uint32_t(t.tile_bytes());
  p->y_bytes = uint32_t(size_t(t.R) * (d->y_is_double ? sizeof(double) : sizeof(int)));
  p->params = io.d_params;
  p->alpha_vec = io.d_alpha_vec;
  p->sigma_vec = io.d_sigma_vec;
  p->row_dalpha = io.d_row_dalpha;
  p->row_dsigma = io.d_row_dsigma;
  p->scratch = ws->scratch;
  p->ticket = ws->ticket;
  p->result = io.d_result;
  p->flags = io.flags;
  p->lgamma_sum = io.d_y ? io.y_lgamma_sum : d->lgamma_sum;
  p->inline_n = 0;
  p->host_flag = io.host_flag;
  p->host_seq = io.host_seq;
  const int np = t.kpad + 2 + p->ncuts;
  if (io.h_params_inline && np <= kInlineParams && family != STAN_CUDA_CATEGORICAL_LOGIT) {
    p->inline_n = np;
    for (int i = 0; i < np; ++i) p->inline_params[i] = io.h_params_inline[i];
  } else if (!io.d_params) {
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  }
  p->xchg_world = 0;
  p->xchg_rank = 0;
  p->xchg_seq = io.xchg_seq;
  p->xchg_local = nullptr;
  for (int q = 0; q < kXchgMaxWorld; ++q) p->xchg_peers[q] = nullptr;
  if (io.xchg && io.xchg->world > 1) {
    if (result_cols(d, io) > size_t(kXchgCols)) return fail(STAN_CUDA_ERR_ARG, "launch_glm: result block exceeds the exchange slot");
    p->xchg_world = io.xchg->world;
    p->xchg_rank = io.xchg->rank;
    p->xchg_local = io.xchg->local;
    for (int q = 0; q < io.xchg->world; ++q) p->xchg_peers[q] = io.xchg->peers[q];
  }
#ifdef STAN_CUDA_PHASE_TIMERS
  p->timers = g_phase_timers;
#endif
  return STAN_CUDA_OK;
}

int launch_glm(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  if (d->family == STAN_CUDA_ORDERED_LOGISTIC && io.ncuts > kMaxCuts) return launch_glm_ordered_wide(d, io, ws, stream);
  if (!d->colblocks.empty()) return launch_glm_colblocks(d, io, ws, stream);  // k > 512
  const Tiling& t = d->t;
  if (d->family == STAN_CUDA_CATEGORICAL_LOGIT && t.layout == kLayoutPlain) return launch_glm_categorical_wide(d, io, ws, stream);
  if (d->family == STAN_CUDA_CATEGORICAL_LOGIT)
    return t.layout == kLayoutMma ? launch_glm_categorical_mma(d, io, ws, stream) : launch_glm_categorical(d, io, ws, stream);
  if (t.layout == kLayoutWarp) return launch_glm_warp(d, io, ws, stream);
  if (d->family >= STAN_CUDA_NEG_BINOMIAL_2_LOG)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: neg_binomial_2_log and ordered_logistic run on the warp kernel only");
#ifndef STAN_CUDA_WITH_CTA_KERNEL
  return fail(STAN_CUDA_ERR_STATE, "launch_glm: this library was built without the CTA-synchronous kernel variant "
                                   "(STAN_CUDA_WITH_CTA_KERNEL)");
#else
  if (d->n <= 0 || t.tiles <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_glm: empty data block");
  const int sms = sm_count(d->device);
  int ctas_per_sm = tuning().ctas_per_sm;
  if (ctas_per_sm <= 0) ctas_per_sm = t.threads >= 512 ? 1 : (t.threads >= 256 ? 2 : 4);
  // stages: as many as fit next to the other resident CTAs (228 KB per SM, 1 KB reserved per CTA)
  const size_t budget = (size_t(228) * 1024) / ctas_per_sm - 1024 - 64;
  int stages = tuning().stages;
  size_t stage_bytes = 0;
  if (stages <= 0) {
    stages = 4;
    while (stages > 1 && smem_bytes(t, stages, nullptr) > budget) --stages;
  }
  size_t smem = smem_bytes(t, stages, &stage_bytes);
  while (stages > 1 && smem > size_t(227) * 1024) smem = smem_bytes(t, --stages, &stage_bytes);
  if (smem > size_t(227) * 1024) return fail(STAN_CUDA_ERR_ARG, "launch_glm: tile does not fit shared memory");

  int64_t grid64 = int64_t(sms) * ctas_per_sm;
  if (grid64 > t.tiles) grid64 = t.tiles;
  const int grid = int(grid64);
  int group_size = 1;  // ~sqrt(grid): both levels of the final sum have about the same length
  while (group_size * group_size < grid) ++group_size;
  const int ngroups = (grid + group_size - 1) / group_size;  // <= 25 for grid <= 592 (ticket block holds 63)
  if (ngroups > 60) return fail(STAN_CUDA_ERR_ARG, "launch_glm: grid too large for the ticket block");
  SC_TRY(workspace_reserve(ws, d->device, size_t(grid + ngroups) * (t.kpad + 4)));

  GlmKernelParams p;
  SC_TRY(fill_kernel_params(d, io, ws, &p));
  p.stages = stages;
  p.group_size = group_size;
  p.stage_bytes = uint32_t(stage_bytes);
  const bool rowio = io.d_alpha_vec || io.d_sigma_vec || io.d_row_dalpha || io.d_row_dsigma;
  switch (d->family) {
    case STAN_CUDA_BERNOULLI_LOGIT:
      return launch_family<STAN_CUDA_BERNOULLI_LOGIT>(rowio, t.threads, p, grid, smem, stream);
    case STAN_CUDA_POISSON_LOG:
      return launch_family<STAN_CUDA_POISSON_LOG>(rowio, t.threads, p, grid, smem, stream);
    case STAN_CUDA_NORMAL_ID:
      return launch_family<STAN_CUDA_NORMAL_ID>(rowio, t.threads, p, grid, smem, stream);
  }
  return fail(STAN_CUDA_ERR_ARG, "launch_glm: unknown family");
#endif  // STAN_CUDA_WITH_CTA_KERNEL
}

}  // namespace stancuda

#ifdef STAN_CUDA_PHASE_TIMERS
extern "C" STAN_CUDA_API void stan_cuda_debug_set_timers(unsigned long long* device_buf) {
  stancuda::g_phase_timers = device_buf;
}
#endif

extern "C" void stan_cuda_set_kernel(int variant) {
#ifdef STAN_CUDA_WITH_CTA_KERNEL
  stancuda::tuning_mut().layout = variant == 0 ? stancuda::kLayoutCta : stancuda::kLayoutWarp;
#else
  (void)variant;  // built without the superseded variant: the warp kernel is the only one
  stancuda::tuning_mut().layout = stancuda::kLayoutWarp;
#endif
}

extern "C" void stan_cuda_set_categorical_kernel(int variant) { stancuda::tuning_mut().cat_mma = variant == 1 ? 1 : 0; }

extern "C" void stan_cuda_set_tuning(int threads, int stages, int ctas_per_sm) {
  stancuda::Tuning& t = stancuda::tuning_mut();
  if (threads >= 0) t.threads = threads;
  if (stages >= 0) t.stages = stages;
  if (ctas_per_sm >= 0) t.ctas_per_sm = ctas_per_sm;
}
