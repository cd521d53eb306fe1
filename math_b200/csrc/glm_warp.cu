// glm_warp.cu -- warp-synchronous variant of the fused GLM value+gradient kernel (sm_100a, FP64).
//
// Same contract as glm_fused.cu (one pass over X: eta = X*beta + alpha, per-row log-likelihood,
// d = dL/deta, g = X^T d;  bernoulli_logit_glm_lpmf.hpp:88-135, poisson_log_glm_lpmf.hpp:91-129,
// normal_id_glm_lpdf.hpp:100-172), different decomposition:
//
//  * a CTA is 8 warps.  Every stage of the TMA ring has a "full" mbarrier (cp.async.bulk
//    complete_tx) and an arrival counter in shared memory: a warp bumps the counter once its part
//    of the tile sits in registers, and the LAST warp to do so re-arms the stage with the tile
//    `stages` ahead -- the earliest possible moment, and no warp ever waits for another one.
//    There is NO CTA-wide barrier in the tile loop.  (Tried and dropped: a dedicated producer warp
//    -- 288 threads x 2 CTAs leave 112 registers per thread and the 48 live doubles spill; a fixed
//    producer lane in warp 0 -- CTAs whose warp 0 fell behind degenerated to one tile in flight and
//    ran 2x slower than the rest.)
//  * all C column groups of a row live in ONE warp: lane = row_in_warp * C + column_group.  The
//    partial dot products meet through log2(C) xor-shuffles (every lane of the row ends up with the
//    bit-identical eta), every lane runs the family epilogue for its row, and goes straight on
//    to its 16 gradient FMAs.  Warps never wait for each other, so the long dependent FP64 chains of
//    the epilogues (exp / log1p / division) of 16 resident warps overlap each other and the loads;
//    in glm_fused.cu one owner warp per CTA ran the epilogue while the other seven sat at a barrier.
//  * tile layout in HBM (see data.cu): x_dev[tile][j][warp][lane], so the j-th load of a warp is one
//    contiguous 256-byte shared-memory read, conflict-free.
//  * two tiles per iteration when a row is shared by C >= 2 lanes (PAIR, below): each row's epilogue is
//    evaluated by half of its lanes instead of all of them, and x is re-read from the still-resident
//    stage for the gradient FMAs instead of being held in registers (a stage is then released after
//    the gradient, not after the load).
//  * reductions are fixed-order (shuffle trees, then warps in order, then finish_reduction):
//    results are bit-reproducible.
// Families: bernoulli_logit, poisson_log, normal_id, neg_binomial_2_log
// (neg_binomial_2_log_glm_lpmf.hpp:108-197) and ordered_logistic (ordered_logistic_glm_lpmf.hpp:93-199);
// categorical_logit has its own kernel (glm_categorical.cu).
#include "glm_device.cuh"

namespace stancuda {

constexpr int kWarpConsumers = 8;                 // warps per CTA
constexpr int kWarpThreads = kWarpConsumers * 32;

__device__ __forceinline__ void consumer_sync() { __syncthreads(); }

// FULLKC: every column group has all kMaxKc columns (k_pad is a multiple of 16, the case whenever K >= 16):
// the `j < kc` guards of the register loops vanish at compile time.
// PAIR (needs FULLKC and C >= 2): two tiles per iteration; the even column-group lanes of a row evaluate
// the epilogue of tile A's row, the odd ones tile B's, `d` is exchanged by shuffle and x is re-read from
// the still-resident stages for the gradient FMAs -- the epilogue runs C/2 instead of C times per row.
// The body is shared by two entry points: glm_warp_kernel (one problem, parameter block by value in the kernel
// arguments) and glm_warp_batch_kernel (blockIdx.y selects one of several independent problems whose parameter
// blocks sit in device memory: map_rect shards of one device in ONE launch).
template <int FAMILY, bool ROWIO, bool FULLKC, bool PAIR>
__device__ __forceinline__ void glm_warp_body(const GlmKernelParams& p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int NT = kWarpConsumers * 32;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int C = p.C, kc = FULLKC ? kMaxKc : p.kc, kpad = p.kpad, R = p.R;
  const int rpw = 32 / C;  // rows per warp
  const int cg = lane & (C - 1);
  const int rl = lane / C;

  unsigned char* stage0 = smem;
  const int ncuts = FAMILY == STAN_CUDA_ORDERED_LOGISTIC ? p.ncuts : 0;
  const int cols = kpad + 4 + ncuts;  // width of a partial: [lp, sum d, a2, a3, dbeta(kpad), dcuts(ncuts)]
  double* red = reinterpret_cast<double*>(smem + size_t(p.stages) * p.stage_bytes);  // [WARPS][cols]
  double* part = red + size_t(kWarpConsumers) * cols;                                // [8]
  double* ord_tab = part + 8;                                                        // [kOrdTabRows][ncuts + 1]
  const int ord_doubles = FAMILY == STAN_CUDA_ORDERED_LOGISTIC ? kOrdTabRows * (ncuts + 1) : 0;
  const bool ord_private = FAMILY == STAN_CUDA_ORDERED_LOGISTIC && p.ord_private;
  double* ord_bins = ord_tab + ord_doubles;                                          // [ncuts][R] when ord_private
  const int RS = PAIR ? 2 * R : R;  // row slots of d/dcuts bins (PAIR: tile A slots, then tile B slots)
  uint64_t* full = reinterpret_cast<uint64_t*>(ord_bins + (ord_private ? ncuts * RS : 0));  // [stages]
  int* arrived = reinterpret_cast<int*>(full + p.stages);                            // [stages]
  __shared__ int s_flag;

  const int grid = gridDim.x;
  // static tile schedule: blocked (CTA b owns a contiguous run of tiles) or round-robin
  const int per = (p.tiles + grid - 1) / grid;
  const int64_t tile0 = p.blocked ? int64_t(blockIdx.x) * per : int64_t(blockIdx.x);
  const int64_t tstride = p.blocked ? 1 : grid;
  int ntl;  // tiles of this CTA
  if (p.blocked) {
    const int64_t left = int64_t(p.tiles) - tile0;
    ntl = left <= 0 ? 0 : (left < per ? int(left) : per);
  } else {
    ntl = (p.tiles - int(blockIdx.x) + grid - 1) / grid;
  }

  // Programmatic dependent launch (p.pdl): the NEXT evaluation on this stream may start while this one is still in
  // its reduction / exchange tail.  Nothing a GLM kernel writes (per-CTA partials, tickets, result block, exchange
  // slots) is touched before the griddepcontrol.wait further down, and everything it reads before that point (X, y,
  // the parameter block) is never written by a GLM kernel -- so the tile loop of evaluation n + 1 overlaps the tail of
  // evaluation n.  Kernels with per-row outputs (ROWIO, the column-block chain) are launched without the attribute.
  if (p.pdl) pdl_launch_dependents();
  PHASE_STAMP(0);
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(smem_u32(&full[s]), 1);
      arrived[s] = 0;
    }
    mbar_fence_init();
  }
  if (FAMILY == STAN_CUDA_ORDERED_LOGISTIC) {
    // per-class table into shared memory; the per-warp bins of d/dcuts (tail of each warp's row of `red`) start at zero
    ordered_class_table<NT>(p, ord_tab, tid);
    for (int j = tid; j < kWarpConsumers * ncuts; j += NT) red[size_t(j / ncuts) * cols + kpad + 4 + j % ncuts] = 0.0;
    if (ord_private)
      for (int j = tid; j < ncuts * RS; j += NT) ord_bins[j] = 0.0;
  }
  __syncthreads();

  // ---- producer role: lane 0 of warp 0 ------------------------------------------------------------
  const uint32_t tx_bytes = p.tile_bytes + p.y_bytes;
  const unsigned char* xbytes = reinterpret_cast<const unsigned char*>(p.x);
  const unsigned char* ybytes = reinterpret_cast<const unsigned char*>(p.y);
  uint64_t policy = 0;
  auto issue = [&](int t) {  // stage tile number t of this CTA
    const int s = t % p.stages;
    const int64_t tile = tile0 + int64_t(t) * tstride;
    const uint32_t bar = smem_u32(&full[s]);
    const uint32_t dst = smem_u32(stage0 + size_t(s) * p.stage_bytes);
    mbar_expect_tx(bar, tx_bytes);
    tma_bulk_g2s(dst, xbytes + tile * p.tile_bytes, p.tile_bytes, bar, policy);
    tma_bulk_g2s(dst + p.tile_bytes, ybytes + tile * p.y_bytes, p.y_bytes, bar, policy);
  };
  if (lane == 0) policy = l2_evict_first_policy();
  if (tid == 0) {
    const int pre = ntl < p.stages ? ntl : p.stages;
    for (int t = 0; t < pre; ++t) issue(t);
  }

  // ---- consumer warps --------------------------------------------------------------------------
  double bj[kMaxKc], acc[kMaxKc];
#pragma unroll
  for (int j = 0; j < kMaxKc; ++j) {
    bj[j] = j < kc ? kernel_param(p, cg * kc + j) : 0.0;
    acc[j] = 0.0;
  }
  const double alpha0 = kernel_param(p, kpad);
  const double sigma0 = kernel_param(p, kpad + 1);
  double inv_sigma0 = FAMILY == STAN_CUDA_NORMAL_ID ? 1.0 / sigma0 : 1.0;
  const double log_sigma0 = FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG ? log(sigma0) : 0.0;  // log phi
  if (FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG) inv_sigma0 = 1.0 / sigma0;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;  // lp main, sum d, sum s^2, sum log sigma (cg == 0 lanes)
  const int r_in_tile = warp * rpw + rl;

  // One row's family epilogue: per-row operands, log-density terms and d = dL/deta; the owner lane of
  // the row (`own`) adds the terms to its accumulators and writes the per-row outputs.  `slot` is the
  // row's slot in the d/dcuts bins, `u` the PAIR role (0: tile A, 1: tile B; always 0 otherwise).
  auto row_terms = [&](double eta, double yv, int64_t row, bool valid, bool own, int slot, int u) -> double {
    if (FAMILY == kFamilyLinearBlock) {
      // column-block path (k > 512): this block's share of the linear predictor goes into the per-row carry
      // (row_dsigma = carry out, alpha_vec | alpha0 = carry in), or a given per-row d (sigma_vec) drives X_b^T d
      double dd = 0.0;
      if (valid) {
        if (own && p.row_dsigma) p.row_dsigma[row] = eta + (p.alpha_vec ? p.alpha_vec[row] : alpha0);
        if (p.sigma_vec) dd = p.sigma_vec[row];
        if (own) a1 += dd;
      }
      return dd;
    }
    double alpha = alpha0, inv_sigma = inv_sigma0, sigma_row = sigma0, log_sigma_row = log_sigma0;
    if (ROWIO && valid) {
      if (p.alpha_vec) alpha = p.alpha_vec[row];
      if (FAMILY >= STAN_CUDA_NORMAL_ID && p.sigma_vec) {
        sigma_row = p.sigma_vec[row];
        inv_sigma = 1.0 / sigma_row;
        if (FAMILY != STAN_CUDA_NORMAL_ID) log_sigma_row = log(sigma_row);
      }
    }
    RowTerms o;
    if (FAMILY == STAN_CUDA_ORDERED_LOGISTIC) {
      const int c = valid ? int(yv) : 1;  // 1 <= c <= ncuts + 1 was checked from the range of y cached at fill
      if (ROWIO && p.alpha_vec) eta += alpha;  // column-block path: the other blocks' share of the location
      o = ordered_row_epilogue(eta, c, ncuts + 1, ord_tab);
      if (valid && own) {
        a0 += o.t0;
        a1 += o.d;
        a2 += eta;  // ordered_logistic_glm_lpmf.hpp:119: finiteness of sum(location)
        if (ROWIO && p.row_dalpha) p.row_dalpha[row] = o.d;
      }
      if ((p.flags & STAN_CUDA_NEED_DCUTS) && ord_private) {
        // d/dcuts (:192-200), few classes: every row slot of the tile owns a private row of bins
        // ([class][row slot]: conflict-free, no ordering question), summed in slot order at the end
        if (valid && own) {
          if (c != ncuts + 1) ord_bins[(c - 1) * RS + slot] += o.t3;
          if (c != 1) ord_bins[(c - 2) * RS + slot] -= o.t2;
        }
      } else if ((p.flags & STAN_CUDA_NEED_DCUTS) && ncuts > 0) {
        // many classes: one owner lane per row adds into this warp's bins, rows in order, so the sum is
        // the same on every run (no floating-point atomics)
        // (warp_bin_add: rows of the warp that hit the same bin are grouped with match.any.)  Two rounds -- all +d2 terms,
        // then all -d1 terms (:192-200) -- instead of one predicated read-modify-write per row of the warp (32 serial
        // steps at K = 32: 1e6 x 32 x 64 classes 73.9 -> see profiles).
        double* bins = red + size_t(warp) * cols + kpad + 4;
        const bool mine = own && valid;
        warp_bin_add(bins, lane, mine && c != ncuts + 1, c - 1, o.t3);
        warp_bin_add(bins, lane, mine && c != 1, c - 2, -o.t2);
      }
    } else if (FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG) {
      o = nb_row_epilogue(eta, alpha, yv, sigma_row, log_sigma_row, inv_sigma, p.flags, ROWIO && p.nb_rowfun != 0, ROWIO);
    } else {
      o = row_epilogue<FAMILY>(eta, alpha, yv, inv_sigma);
    }
    if (FAMILY != STAN_CUDA_ORDERED_LOGISTIC && valid && own) {
      a0 += o.t0;
      a1 += o.d;
      a2 += o.t2;
      if (FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG) a3 += o.t3;
      if (ROWIO) {
        if (FAMILY == STAN_CUDA_NORMAL_ID && p.sigma_vec) {
          a3 += log(sigma_row);
          if (p.row_dsigma) p.row_dsigma[row] = (o.t2 - 1.0) * inv_sigma;
        }
        if (FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG && p.sigma_vec && p.row_dsigma)
          p.row_dsigma[row] = 1.0 + o.t3;                                     // :180-183
        if (p.row_dalpha) p.row_dalpha[row] = o.d;
      }
    }
    return valid ? o.d : 0.0;
  };
  // this warp is done with stage s (tile number `it` of the CTA): count it; the last warp re-arms the stage
  auto release = [&](int s, int it) {
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      if (atomicAdd(&arrived[s], 1) == kWarpConsumers - 1) {
        arrived[s] = 0;  // the next arrivals on this stage come only after the refill below has landed
        __threadfence_block();
        if (it + p.stages < ntl) issue(it + p.stages);
      }
    }
  };
  auto outcome = [&](const unsigned char* st) -> double {
    if (FAMILY == STAN_CUDA_NORMAL_ID) return reinterpret_cast<const double*>(st + p.tile_bytes)[r_in_tile];
    return double(reinterpret_cast<const int*>(st + p.tile_bytes)[r_in_tile]);
  };

  if (PAIR) {
    const int u = cg & 1;  // this lane's role: epilogue of tile A's row (0) or of tile B's row (1)
    for (int it = 0; it < ntl; it += 2) {
      const bool have_b = it + 1 < ntl;
      const int sa = it % p.stages, sb = (it + 1) % p.stages;
      const unsigned char* sta = stage0 + size_t(sa) * p.stage_bytes;
      const unsigned char* stb = stage0 + size_t(sb) * p.stage_bytes;
      const double* xa = reinterpret_cast<const double*>(sta) + warp * 32 + lane;
      const double* xb = reinterpret_cast<const double*>(stb) + warp * 32 + lane;
      double eta_a, eta_b = 0.0, y_a, y_b = 0.0;
      mbar_wait(smem_u32(&full[sa]), (it / p.stages) & 1);
#ifdef STAN_CUDA_PHASE_TIMERS
      if (it == 0) PHASE_STAMP(1);
#endif
      {
        double p0 = 0.0, p1 = 0.0;
#pragma unroll
        for (int j = 0; j < kMaxKc; j += 2) {
          p0 = fma(xa[j * NT], bj[j], p0);
          p1 = fma(xa[(j + 1) * NT], bj[j + 1], p1);
        }
        eta_a = p0 + p1;
        y_a = outcome(sta);
      }
      if (have_b) {
        mbar_wait(smem_u32(&full[sb]), ((it + 1) / p.stages) & 1);
        double p0 = 0.0, p1 = 0.0;
#pragma unroll
        for (int j = 0; j < kMaxKc; j += 2) {
          p0 = fma(xb[j * NT], bj[j], p0);
          p1 = fma(xb[(j + 1) * NT], bj[j + 1], p1);
        }
        eta_b = p0 + p1;
        y_b = outcome(stb);
      }
      for (int off = 1; off < C; off <<= 1) {
        eta_a += __shfl_xor_sync(0xffffffffu, eta_a, off);
        eta_b += __shfl_xor_sync(0xffffffffu, eta_b, off);
      }
      const int64_t tile = tile0 + int64_t(it + u) * tstride;
      const int64_t row = tile * R + r_in_tile;
      const bool valid = row < p.n && (u == 0 || have_b);
      const double d_mine = row_terms(u ? eta_b : eta_a, u ? y_b : y_a, row, valid, cg == u, r_in_tile + u * R, u);
      const double d_a = __shfl_sync(0xffffffffu, d_mine, lane & ~1);
      const double d_b = __shfl_sync(0xffffffffu, d_mine, lane | 1);
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j) acc[j] = fma(xa[j * NT], d_a, acc[j]);
      release(sa, it);
      if (have_b) {
#pragma unroll
        for (int j = 0; j < kMaxKc; ++j) acc[j] = fma(xb[j * NT], d_b, acc[j]);
        release(sb, it + 1);
      }
    }
    // the owners of tile B rows (cg == 1) hand their sums to the cg == 0 lanes; other lanes hold zeros
    a0 += __shfl_xor_sync(0xffffffffu, a0, 1);
    a1 += __shfl_xor_sync(0xffffffffu, a1, 1);
    a2 += __shfl_xor_sync(0xffffffffu, a2, 1);
    a3 += __shfl_xor_sync(0xffffffffu, a3, 1);
  } else {
    for (int it = 0; it < ntl; ++it) {
      const int s = it % p.stages;
      const int64_t tile = tile0 + int64_t(it) * tstride;
      const int64_t row = tile * R + r_in_tile;
      mbar_wait(smem_u32(&full[s]), (it / p.stages) & 1);
#ifdef STAN_CUDA_PHASE_TIMERS
      if (it == 0) PHASE_STAMP(1);
      if (p.timers && (it & 7) == 0 && (it >> 3) < 32 && lane == 0 && (warp == 0 || warp == 7))
        p.timers[8 * 2048 + (size_t(blockIdx.x) * 2 + (warp == 7)) * 32 + (it >> 3)] = gtimer();
#endif
      const unsigned char* st = stage0 + size_t(s) * p.stage_bytes;
      const double* xs = reinterpret_cast<const double*>(st) + warp * 32 + lane;
      double xr[kMaxKc];
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j) xr[j] = j < kc ? xs[j * NT] : 0.0;
      const double yv = outcome(st);
      double p0 = 0.0, p1 = 0.0;
#pragma unroll
      for (int j = 0; j < kMaxKc; j += 2) {
        p0 = fma(xr[j], bj[j], p0);
        p1 = fma(xr[j + 1], bj[j + 1], p1);
      }
      double eta = p0 + p1;
      release(s, it);  // this warp's part of the tile sits in registers
      // partial dots of the C column groups of a row: xor butterfly, identical result in every lane
      for (int off = 1; off < C; off <<= 1) eta += __shfl_xor_sync(0xffffffffu, eta, off);
      const double d = row_terms(eta, yv, row, row < p.n, cg == 0, r_in_tile, 0);
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j) acc[j] = fma(xr[j], d, acc[j]);
    }
  }

  PHASE_STAMP(2);
  // ---- CTA reduction: rows of a warp (xor offsets C .. 16), then the 8 warps in order ----------
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j) {
      const double o = __shfl_xor_sync(0xffffffffu, acc[j], off);
      if (off >= C) acc[j] += o;
    }
    const double o0 = __shfl_xor_sync(0xffffffffu, a0, off), o1 = __shfl_xor_sync(0xffffffffu, a1, off);
    const double o2 = __shfl_xor_sync(0xffffffffu, a2, off), o3 = __shfl_xor_sync(0xffffffffu, a3, off);
    if (off >= C) {
      a0 += o0;
      a1 += o1;
      a2 += o2;
      a3 += o3;
    }
  }
  double* myred = red + size_t(warp) * cols;
  if (rl == 0) {  // lanes 0 .. C-1 hold the sums of column group cg = lane
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j)
      if (j < kc) myred[4 + cg * kc + j] = acc[j];
    if (cg == 0) {
      myred[0] = a0;
      myred[1] = a1;
      myred[2] = a2;
      myred[3] = a3;
    }
  }
  consumer_sync();
  // first write to memory shared with the previous evaluation: wait until that grid has completed and flushed
  if (p.pdl) pdl_wait();
  double* my_scratch = p.scratch + size_t(blockIdx.x) * cols;
  for (int c = tid; c < cols; c += NT) {
    double sum = 0.0;
    if (ord_private && c >= kpad + 4) {
      const double* b = ord_bins + (c - kpad - 4) * RS;
      for (int r = 0; r < RS; ++r) sum += b[r];
    } else {
#pragma unroll
      for (int w = 0; w < kWarpConsumers; ++w) sum += red[size_t(w) * cols + c];
    }
    my_scratch[c] = sum;
  }
  double* fin = red;  // [cols]: the warps' partials were summed into my_scratch above (finish_reduction starts with a barrier);
                      // `red` holds 8 x cols doubles whatever the tile shape, the stages may be smaller than cols
  finish_reduction<FAMILY, NT>(p, tid, cols, part, fin, &s_flag, [] { consumer_sync(); });
}

template <int FAMILY, bool ROWIO, bool FULLKC, bool PAIR>
__global__ void __launch_bounds__(kWarpThreads, 2) glm_warp_kernel(const GlmKernelParams p) {
  glm_warp_body<FAMILY, ROWIO, FULLKC, PAIR>(p);
}

// grid = (CTAs per problem, problems): every problem has its own scratch, tickets and result block, and
// gridDim.x / blockIdx.x mean inside the body what they mean in a single launch.
template <int FAMILY, bool FULLKC, bool PAIR>
__global__ void __launch_bounds__(kWarpThreads, 2) glm_warp_batch_kernel(const GlmKernelParams* __restrict__ batch) {
  glm_warp_body<FAMILY, false, FULLKC, PAIR>(batch[blockIdx.y]);
}

// =============================================================================================
// host side
// =============================================================================================
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

bool pdl_enabled() {
  static const int on = [] {
    const char* v = getenv("STAN_CUDA_PDL");
    return v ? atoi(v) : 1;
  }();
  return on != 0;
}
int pdl_slack() {
  static const int slack = [] {
    const char* v = getenv("STAN_CUDA_PDL_SLACK");
    return v ? atoi(v) : 1;
  }();
  return slack;
}

static bool pair_mode(const Tiling& t) {
  static const int on = [] {
    const char* v = getenv("STAN_CUDA_PAIR");  // STAN_CUDA_PAIR=0: one tile per iteration (A/B)
    return v ? atoi(v) : 1;
  }();
  return on != 0 && t.kc == kMaxKc && t.C >= 2;
}
static bool ord_private_bins(const Tiling& t, int ncuts) {
  return ncuts > 0 && size_t(ncuts) * t.R * (pair_mode(t) ? 2 : 1) <= size_t(kOrdPrivateDoubles);
}

static size_t warp_smem_bytes(const Tiling& t, int stages, int ncuts, size_t* stage_bytes_out) {
  const size_t stage = align_up(t.tile_bytes() + size_t(t.R) * sizeof(double), 128);
  if (stage_bytes_out) *stage_bytes_out = stage;
  const size_t fixed = (size_t(kWarpConsumers) * (t.kpad + 4 + ncuts) + 8 + size_t(kOrdTabRows) * (ncuts + 1) + (ord_private_bins(t, ncuts) ? size_t(ncuts) * t.R * (pair_mode(t) ? 2 : 1) : 0)) * sizeof(double) + size_t(stages) * (sizeof(uint64_t) + sizeof(int)) + 8;
  return size_t(stages) * stage + fixed;
}

template <int FAMILY, bool ROWIO, bool FULLKC, bool PAIR = false>
static int launch_warp_inst2(const GlmKernelParams& p, int grid, size_t smem, cudaStream_t stream) {
  auto kern = glm_warp_kernel<FAMILY, ROWIO, FULLKC, PAIR>;
  static std::atomic<size_t> configured[64];
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && configured[dev].load(std::memory_order_relaxed) < smem) {
    SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev].store(smem, std::memory_order_relaxed);
  }
  SC_CUDA(launch_maybe_pdl(kern, dim3(unsigned(grid)), dim3(kWarpThreads), smem, stream, p.pdl != 0, p));
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

template <int FAMILY, bool ROWIO>
static int launch_warp_inst(const GlmKernelParams& p, int grid, size_t smem, cudaStream_t stream) {
  // measured: the specialisation gains 1-4 % for the cheap epilogues; for neg_binomial_2_log it costs registers
  // (more spills, 208 -> 242 us at 1.25e6 x 128), so that family stays on the generic instantiation
  if (p.pair) return launch_warp_inst2<FAMILY, ROWIO, true, true>(p, grid, smem, stream);
  if (FAMILY != STAN_CUDA_NEG_BINOMIAL_2_LOG && p.kc == kMaxKc)
    return launch_warp_inst2<FAMILY, ROWIO, FAMILY != STAN_CUDA_NEG_BINOMIAL_2_LOG>(p, grid, smem, stream);
  return launch_warp_inst2<FAMILY, ROWIO, false>(p, grid, smem, stream);
}

// neg_binomial_2_log with a scalar phi: sum_i lgamma(y_i + phi) and sum_i digamma(y_i + phi) depend on
// the data only through the histogram of y, which is built once at cache fill.  One small CTA turns
// the histogram into the two sums per evaluation (O(y_max) special-function calls instead of O(N)
// inside the hot loop); fixed-order tree reduction, so the sums are bit-reproducible.
__global__ void __launch_bounds__(256) nb_hist_sums_kernel(const unsigned int* __restrict__ hist, int len,
                                                           const double* d_params, int phi_index, double phi_inline,
                                                           int use_inline, double* __restrict__ out) {
  // In the dependent-launch chain GLM(n) -> this kernel -> GLM(n + 1): let GLM(n + 1) start right away; the two sums go
  // to memory GLM(n) reads at its very end, so they are written only behind griddepcontrol.wait (GLM(n) complete).
  pdl_launch_dependents();
  __shared__ double sh[2][256];
  const double phi = use_inline ? phi_inline : d_params[phi_index];
  double s_lg = 0.0, s_dg = 0.0;
  for (int v = threadIdx.x; v < len; v += 256) {
    const unsigned int c = hist[v];
    if (c) {
      s_lg += double(c) * lgamma(double(v) + phi);
      s_dg += double(c) * digamma_pos(double(v) + phi);
    }
  }
  sh[0][threadIdx.x] = s_lg;
  sh[1][threadIdx.x] = s_dg;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (int(threadIdx.x) < off) {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + off];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    pdl_wait();
    out[0] = sh[0][0];
    out[1] = sh[1][0];
  }
}

static int prepare_neg_binomial(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, GlmKernelParams* p,
                                cudaStream_t stream) {
  p->nb_sums = nullptr;
  p->nb_rowfun = 1;
  if (io.d_sigma_vec || !d->hist || io.d_y) return STAN_CUDA_OK;  // vector phi, outcomes too large for a histogram, or per-call outcomes
  double* out = ws->scratch + ws->scratch_doubles - 2;  // last two doubles of the workspace
  const bool inl = p->inline_n > 0;
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(256);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p->pdl ? 1 : 0;
    SC_CUDA(cudaLaunchKernelEx(&cfg, nb_hist_sums_kernel, static_cast<const unsigned int*>(d->hist), d->hist_len, io.d_params,
                               d->t.kpad + 1, inl ? p->inline_params[d->t.kpad + 1] : 0.0, inl ? 1 : 0, out));
  }
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  p->nb_sums = out;
  p->nb_rowfun = 0;
  return STAN_CUDA_OK;
}

int launch_glm_warp(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  const Tiling& t = d->t;
  const int family = io.family_override >= 0 ? io.family_override : d->family;
  if (d->n <= 0 || t.tiles <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_glm: empty data block");
  const int sms = sm_count(d->device);
  int ctas_per_sm = tuning().ctas_per_sm;
  if (ctas_per_sm <= 0 || ctas_per_sm > 2) ctas_per_sm = 2;
  const size_t budget = (size_t(228) * 1024) / ctas_per_sm - 1024 - 64;
  const int ncuts = family == STAN_CUDA_ORDERED_LOGISTIC ? io.ncuts : 0;
  if (ncuts < 0 || ncuts > kMaxCuts)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: ordered_logistic supports at most 1024 cut-points on the device");
  const size_t cols = size_t(t.kpad) + 4 + ncuts;
  int stages = tuning().stages;
  size_t stage_bytes = 0;
  if (stages <= 0) {
    stages = 6;
    while (stages > 2 && warp_smem_bytes(t, stages, ncuts, nullptr) > budget) --stages;
  }
  size_t smem = warp_smem_bytes(t, stages, ncuts, &stage_bytes);
  while (stages > 1 && smem > size_t(227) * 1024) smem = warp_smem_bytes(t, --stages, ncuts, &stage_bytes);
  if (smem > size_t(227) * 1024) return fail(STAN_CUDA_ERR_ARG, "launch_glm: tile does not fit shared memory");

  int64_t grid64 = int64_t(sms) * ctas_per_sm;
  if (grid64 > t.tiles) grid64 = t.tiles;
  // Launches that run next to each other (map_rect shards on the streams of concurrent host threads, io.shared_gpu)
  // take a slice of the SMs each, with enough tiles per CTA to amortise ramp and tail: kernels of different
  // shards then overlap instead of queueing behind each other's full-width grids.
  static const int shared_tiles_per_cta = [] {
    const char* v = getenv("STAN_CUDA_SHARED_TILES_PER_CTA");  // 0: full-width grids; measured at C5: 8 is best
    return v ? atoi(v) : 8;
  }();
  if (io.shared_gpu && shared_tiles_per_cta > 0) {
    int64_t g = (t.tiles + shared_tiles_per_cta - 1) / shared_tiles_per_cta;
    if (g < 8) g = 8;
    if (g < grid64) grid64 = g;
  }
  // Programmatic dependent launch for the plain instantiations of the three hot-path families (no per-row operands,
  // not a column block, not next to other shards' launches).  A full-width grid then leaves `slack` CTA slots free:
  // the previous evaluation's last CTA may still sit in its exchange when this grid starts, and with a static tile
  // schedule a CTA that starts late finishes late.  STAN_CUDA_PDL=0 switches it off (A/B).
  const bool rowio = io.d_alpha_vec || io.d_sigma_vec || io.d_row_dalpha || io.d_row_dsigma;
  // neg_binomial_2_log qualifies on its histogram path (its pre-kernel joins the chain, prepare_neg_binomial);
  // ordered_logistic as it is (class table and bins live in shared memory)
  const bool nb_hist = family == STAN_CUDA_NEG_BINOMIAL_2_LOG && !io.d_sigma_vec && d->hist && !io.d_y;
  const bool pdl = pdl_enabled() && !rowio && !io.shared_gpu &&
                   (family <= STAN_CUDA_NORMAL_ID || family == STAN_CUDA_ORDERED_LOGISTIC || nb_hist);
  if (pdl && grid64 == int64_t(sms) * ctas_per_sm && grid64 > pdl_slack() + 8) grid64 -= pdl_slack();
  const int grid = int(grid64);
  int group_size = 1;
  while (group_size * group_size < grid) ++group_size;
  const int ngroups = (grid + group_size - 1) / group_size;
  if (ngroups > 60) return fail(STAN_CUDA_ERR_ARG, "launch_glm: grid too large for the ticket block");
  SC_TRY(workspace_reserve(ws, d->device, size_t(grid + ngroups) * cols + 2));

  GlmKernelParams p;
  SC_TRY(fill_kernel_params(d, io, ws, &p));
  p.stages = stages;
  p.group_size = group_size;
  p.stage_bytes = uint32_t(stage_bytes);
  p.blocked = tuning().blocked;
  p.ord_private = ord_private_bins(t, ncuts) ? 1 : 0;
  p.pair = pair_mode(t) && stages >= 3 ? 1 : 0;
  p.pdl = pdl ? 1 : 0;
  switch (family) {
    case STAN_CUDA_BERNOULLI_LOGIT:
      return rowio ? launch_warp_inst<STAN_CUDA_BERNOULLI_LOGIT, true>(p, grid, smem, stream)
                   : launch_warp_inst<STAN_CUDA_BERNOULLI_LOGIT, false>(p, grid, smem, stream);
    case STAN_CUDA_POISSON_LOG:
      return rowio ? launch_warp_inst<STAN_CUDA_POISSON_LOG, true>(p, grid, smem, stream)
                   : launch_warp_inst<STAN_CUDA_POISSON_LOG, false>(p, grid, smem, stream);
    case STAN_CUDA_NORMAL_ID:
      return rowio ? launch_warp_inst<STAN_CUDA_NORMAL_ID, true>(p, grid, smem, stream)
                   : launch_warp_inst<STAN_CUDA_NORMAL_ID, false>(p, grid, smem, stream);
    case STAN_CUDA_NEG_BINOMIAL_2_LOG:
      SC_TRY(prepare_neg_binomial(d, io, ws, &p, stream));
      // per-row lgamma / digamma (vector phi, or outcomes beyond the histogram) live in the ROWIO instantiation only
      return (rowio || p.nb_rowfun) ? launch_warp_inst<STAN_CUDA_NEG_BINOMIAL_2_LOG, true>(p, grid, smem, stream)
                   : launch_warp_inst<STAN_CUDA_NEG_BINOMIAL_2_LOG, false>(p, grid, smem, stream);
    case STAN_CUDA_ORDERED_LOGISTIC:
      // per-row operands of ordered_logistic exist only inside the column-block path (carry in, d out)
      return rowio ? launch_warp_inst<STAN_CUDA_ORDERED_LOGISTIC, true>(p, grid, smem, stream)
                   : launch_warp_inst<STAN_CUDA_ORDERED_LOGISTIC, false>(p, grid, smem, stream);
    case kFamilyLinearBlock:
      return launch_warp_inst<kFamilyLinearBlock, true>(p, grid, smem, stream);
  }
  return fail(STAN_CUDA_ERR_ARG, "launch_glm: unknown family");
}

// ---------------------------------------------------------------------------------------------
// Batched launch: `count` independent problems of one family and one tile shape on one device in ONE launch
// (grid.y = count).  The caller provides, per problem, a scratch / ticket / result region through the
// Workspace array `wss` (tickets zeroed), pinned host memory for the parameter blocks (h_batch) and its
// device copy (d_batch).  Scalar intercepts only (no per-row operands), families bernoulli_logit /
// poisson_log / normal_id, no exchange.
// ---------------------------------------------------------------------------------------------
template <int FAMILY>
static int launch_warp_batch_family(bool fullkc, bool pair, const GlmKernelParams* d_batch, dim3 grid, size_t smem,
                                    cudaStream_t stream) {
  // (the three kernels share one function-pointer type, so the shared-memory opt-in is tracked per variant)
  auto go = [&](auto kern, int variant) -> int {
    static std::atomic<size_t> configured[3][64];
    int dev = 0;
    SC_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && configured[variant][dev].load(std::memory_order_relaxed) < smem) {
      SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      configured[variant][dev].store(smem, std::memory_order_relaxed);
    }
    kern<<<grid, kWarpThreads, smem, stream>>>(d_batch);
    SC_CUDA(cudaGetLastError());
    counters().launches.fetch_add(1, std::memory_order_relaxed);
    return STAN_CUDA_OK;
  };
  if (pair) return go(glm_warp_batch_kernel<FAMILY, true, true>, 0);
  if (fullkc) return go(glm_warp_batch_kernel<FAMILY, true, false>, 1);
  return go(glm_warp_batch_kernel<FAMILY, false, false>, 2);
}

int glm_warp_batch_plan(const stan_cuda_data* d0, int count, int64_t min_tiles, int* grid_x, int* group_size,
                        size_t* scratch_doubles_per_problem) {
  const Tiling& t = d0->t;
  if (t.layout != kLayoutWarp) return fail(STAN_CUDA_ERR_ARG, "glm batch: needs the warp tile layout");
  const int sms = sm_count(d0->device);
  int64_t gx = (int64_t(sms) * 2) / (count > 0 ? count : 1);  // one wave of 2 CTAs per SM over all problems
  if (gx < 1) gx = 1;
  if (gx > min_tiles) gx = min_tiles;
  int gs = 1;
  while (gs * gs < gx) ++gs;
  const int ngroups = int((gx + gs - 1) / gs);
  if (ngroups > 60) return fail(STAN_CUDA_ERR_ARG, "glm batch: grid too large for the ticket block");
  *grid_x = int(gx);
  *group_size = gs;
  *scratch_doubles_per_problem = size_t(gx + ngroups) * (size_t(t.kpad) + 4) + 2;
  return STAN_CUDA_OK;
}

int launch_glm_warp_batch(const stan_cuda_data* const* ds, const LaunchIO* ios, Workspace* wss, int count, int grid_x,
                          int group_size, GlmKernelParams* h_batch, const GlmKernelParams* d_batch, cudaStream_t stream) {
  if (count < 1) return fail(STAN_CUDA_ERR_ARG, "glm batch: empty");
  const Tiling& t0 = ds[0]->t;
  const int family = ds[0]->family;
  if (family > STAN_CUDA_NORMAL_ID) return fail(STAN_CUDA_ERR_ARG, "glm batch: bernoulli_logit, poisson_log and normal_id only");
  const int sms = sm_count(ds[0]->device);
  (void)sms;
  const size_t budget = (size_t(228) * 1024) / 2 - 1024 - 64;
  int stages = 6;
  size_t stage_bytes = 0;
  while (stages > 2 && warp_smem_bytes(t0, stages, 0, nullptr) > budget) --stages;
  size_t smem = warp_smem_bytes(t0, stages, 0, &stage_bytes);
  while (stages > 1 && smem > size_t(227) * 1024) smem = warp_smem_bytes(t0, --stages, 0, &stage_bytes);
  if (smem > size_t(227) * 1024) return fail(STAN_CUDA_ERR_ARG, "glm batch: tile does not fit shared memory");
  const bool pair = pair_mode(t0) && stages >= 3;
  for (int i = 0; i < count; ++i) {
    const stan_cuda_data* d = ds[i];
    const Tiling& t = d->t;
    if (d->family != family || d->device != ds[0]->device || t.layout != kLayoutWarp || t.C != t0.C || t.kc != t0.kc ||
        t.R != t0.R || t.kpad != t0.kpad || d->y_is_double != ds[0]->y_is_double)
      return fail(STAN_CUDA_ERR_ARG, "glm batch: problems must share family, device and tile shape");
    if (d->n <= 0 || t.tiles < grid_x) return fail(STAN_CUDA_ERR_ARG, "glm batch: a problem has fewer tiles than CTAs");
    if (ios[i].d_alpha_vec || ios[i].d_sigma_vec || ios[i].d_row_dalpha || ios[i].d_row_dsigma || ios[i].xchg)
      return fail(STAN_CUDA_ERR_ARG, "glm batch: scalar intercept / scale only, no exchange");
    GlmKernelParams* p = &h_batch[i];
    SC_TRY(fill_kernel_params(d, ios[i], &wss[i], p));
    p->stages = stages;
    p->group_size = group_size;
    p->stage_bytes = uint32_t(stage_bytes);
    p->blocked = tuning().blocked;
    p->ord_private = 0;
    p->pair = pair ? 1 : 0;
  }
  SC_CUDA(cudaMemcpyAsync(const_cast<GlmKernelParams*>(d_batch), h_batch, size_t(count) * sizeof(GlmKernelParams),
                          cudaMemcpyHostToDevice, stream));
  const dim3 grid(unsigned(grid_x), unsigned(count), 1);
  const bool fullkc = t0.kc == kMaxKc;
  switch (family) {
    case STAN_CUDA_BERNOULLI_LOGIT: return launch_warp_batch_family<STAN_CUDA_BERNOULLI_LOGIT>(fullkc, pair, d_batch, grid, smem, stream);
    case STAN_CUDA_POISSON_LOG: return launch_warp_batch_family<STAN_CUDA_POISSON_LOG>(fullkc, pair, d_batch, grid, smem, stream);
    default: return launch_warp_batch_family<STAN_CUDA_NORMAL_ID>(fullkc, pair, d_batch, grid, smem, stream);
  }
}

// =============================================================================================
// k > 512: column blocks.  The reference accepts any number of columns (bernoulli_logit_glm_lpmf.hpp:60-69 only
// checks consistency); the fused kernel keeps a row's columns in one warp (<= 32 lanes x 16), so wider matrices
// are split into blocks of kColBlock columns, each tiled like a matrix of its own:
//   1. blocks 0 .. B-2 (kFamilyLinearBlock): carry[row] = X_b beta_b + carry[row]   (carry starts as alpha)
//   2. block B-1 (the real family, per-row intercept = carry): log-density, d = dL/deta per row, dbeta_{B-1}
//   3. blocks 0 .. B-2 again: dbeta_b = X_b^T d
// Blocks 0 .. B-2 cross HBM twice, the last one once: (2 - 1/B) x the bytes of one read of X.  Every launch is the
// same glm_warp_kernel with its fixed-order reductions, so results stay bit-reproducible.
// =============================================================================================
__global__ void colblock_split_params_kernel(const double* __restrict__ full, double* __restrict__ blocks, int nblocks,
                                             int kpad_last, int ncuts, int alpha_to_first) {
  // full = [beta(kpad_total), alpha, sigma, cuts(ncuts)];  blocks[b] = [beta_b(kpad_b), alpha_b, sigma, cuts] at stride
  const int b = blockIdx.x;
  const int kpad_total = (nblocks - 1) * kColBlock + kpad_last;
  const int kpad_b = b == nblocks - 1 ? kpad_last : kColBlock;
  const size_t stride = size_t(kColBlock) + 2 + ncuts;
  double* dst = blocks + size_t(b) * stride;
  for (int j = threadIdx.x; j < kpad_b; j += blockDim.x) dst[j] = full[size_t(b) * kColBlock + j];
  if (threadIdx.x == 0) {
    // the scalar intercept enters the carry chain at block 0; the last block then reads it from the carry vector
    dst[kpad_b] = (alpha_to_first ? b == 0 : b == nblocks - 1) ? full[kpad_total] : 0.0;
    dst[kpad_b + 1] = full[kpad_total + 1];
  }
  for (int j = threadIdx.x; j < ncuts; j += blockDim.x) dst[kpad_b + 2 + j] = full[kpad_total + 2 + j];
}

__global__ void colblock_gather_result_kernel(const double* __restrict__ blocks, double* __restrict__ result, int nblocks,
                                              int kpad_last, int ncuts, int have_dbeta) {
  // blocks[b] = [lp, sum d, a2, witness, dbeta_b(kpad_b), dcuts] at stride; result = [4][dbeta(kpad_total)][dcuts]
  const int b = blockIdx.x;
  const int kpad_total = (nblocks - 1) * kColBlock + kpad_last;
  const int kpad_b = b == nblocks - 1 ? kpad_last : kColBlock;
  const size_t stride = size_t(kColBlock) + 4 + ncuts;
  const double* src = blocks + size_t(b) * stride;
  for (int j = threadIdx.x; j < kpad_b; j += blockDim.x)
    result[4 + size_t(b) * kColBlock + j] = (have_dbeta || b == nblocks - 1) ? src[4 + j] : 0.0;
  if (b == nblocks - 1) {
    if (threadIdx.x < 4) result[threadIdx.x] = src[threadIdx.x];
    for (int j = threadIdx.x; j < ncuts; j += blockDim.x) result[4 + kpad_total + j] = src[4 + kpad_b + j];
  }
}

int launch_glm_colblocks(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  const int B = int(d->colblocks.size());
  if (B < 2) return fail(STAN_CUDA_ERR_STATE, "launch_glm: column-block data without blocks");
  if (io.xchg || io.host_flag || io.h_params_inline)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: k > 512 runs as several launches: no in-kernel all-reduce / inline parameters "
                                   "(use stan_cuda_allreduce behind the launch)");
  if (!io.d_params) return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  const stan_cuda_data* last = d->colblocks[size_t(B) - 1];
  const int ncuts = d->family == STAN_CUDA_ORDERED_LOGISTIC ? io.ncuts : 0;
  if (ncuts < 0 || ncuts > kMaxCuts)
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: ordered_logistic supports at most 1024 cut-points on the device");
  const int kpad_last = last->t.kpad;
  const size_t np_stride = size_t(kColBlock) + 2 + ncuts, nr_stride = size_t(kColBlock) + 4 + ncuts;
  const size_t rows_al = (size_t(d->n) + 1) & ~size_t(1);
  // one allocation: [partials of the running launch | parameter blocks | result blocks | carry | d] (+ the 2-double
  // tail neg_binomial_2_log uses); the sub-launches find their partials region already large enough
  const size_t partials = size_t(2 * sm_count(d->device) + 64) * nr_stride + 2;
  const size_t extras = size_t(B) * (np_stride + nr_stride) + 2 * rows_al + 8;
  SC_CUDA(cudaSetDevice(d->device));
  SC_TRY(workspace_reserve(ws, d->device, partials + extras + 4));
  double* base = ws->scratch + ((partials + 1) & ~size_t(1));
  double* par = base;
  double* res = par + size_t(B) * np_stride;
  double* carry = res + size_t(B) * nr_stride;
  double* drow = io.d_row_dalpha ? io.d_row_dalpha : carry + rows_al;

  colblock_split_params_kernel<<<B, 256, 0, stream>>>(io.d_params, par, B, kpad_last, ncuts, io.d_alpha_vec ? 0 : 1);
  SC_CUDA(cudaGetLastError());
  // 1. carry chain over the leading blocks
  for (int b = 0; b < B - 1; ++b) {
    LaunchIO sub{0u, par + size_t(b) * np_stride, nullptr, nullptr, res + size_t(b) * nr_stride, nullptr, carry};
    sub.d_alpha_vec = b == 0 ? io.d_alpha_vec : carry;
    sub.shared_gpu = io.shared_gpu;
    SC_TRY(launch_glm_warp(d->colblocks[size_t(b)], sub, ws, stream));
  }
  // 2. the family's own block: intercept per row = carry, d per row out
  {
    LaunchIO sub{io.flags, par + size_t(B - 1) * np_stride, carry, io.d_sigma_vec, res + size_t(B - 1) * nr_stride, drow,
                 io.d_row_dsigma};
    sub.ncuts = io.ncuts;
    sub.d_y = io.d_y;
    sub.y_lgamma_sum = io.y_lgamma_sum;
    sub.shared_gpu = io.shared_gpu;
    SC_TRY(launch_glm_warp(last, sub, ws, stream));
  }
  // 3. gradient of the leading blocks from d
  const bool have_dbeta = (io.flags & STAN_CUDA_NEED_DBETA) != 0;
  for (int b = 0; have_dbeta && b < B - 1; ++b) {
    LaunchIO sub{0u, par + size_t(b) * np_stride, nullptr, drow, res + size_t(b) * nr_stride, nullptr, nullptr};
    sub.shared_gpu = io.shared_gpu;
    SC_TRY(launch_glm_warp(d->colblocks[size_t(b)], sub, ws, stream));
  }
  colblock_gather_result_kernel<<<B, 256, 0, stream>>>(res, io.d_result, B, kpad_last, ncuts, have_dbeta ? 1 : 0);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(2, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda

// =============================================================================================
// stan_cuda_glm_launch_noalloc -- the fused GLM evaluation for callers that cannot allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The kernel is the same glm_warp_kernel.  What differs is where its reduction scratch lives: the per-CTA
// partials, the group sums and the ticket words of finish_reduction sit in the two file-scope __device__ arrays
// below instead of a cudaMalloc'ed Workspace.  Every GPU that loads the library holds the arrays (1.27 MiB + 256 B)
// until the process exits.  In host code the array NAMES are host-side stand-ins, not device addresses: the
// addresses the kernel receives are resolved with cudaGetSymbolAddress, once per device, with that device current.
// ALL calls on a device share the arrays, which keep the previous call's contents: the kernel overwrites every
// partial it later reads, the ticket words (which must start at zero) are cleared on the stream in front of
// every launch; two calls that OVERLAP on one GPU (different streams, concurrent host threads) race on the
// scratch -- serialise them.
// =============================================================================================
namespace stancuda {

constexpr size_t kNoAllocScratchDoubles = size_t(320) * 520;  // (296 CTAs + 18 group sums) x (k_pad 512 + 4) and the 2-double tail
constexpr size_t kNoAllocTicketWords = 64;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocScratchDoubles];
static __device__ __align__(256) unsigned int g_noalloc_ticket[kNoAllocTicketWords];

struct NoAllocDevice {
  bool ready = false;
  double* scratch = nullptr;       // device address of g_noalloc_scratch on this device
  unsigned int* ticket = nullptr;  // device address of g_noalloc_ticket on this device
};
static std::mutex g_noalloc_mu;
static NoAllocDevice g_noalloc_dev[64];

// once per device, with that device current
static int noalloc_setup(int device, NoAllocDevice** out) {
  if (device < 0 || device >= 64) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  NoAllocDevice& nd = g_noalloc_dev[device];
  if (!nd.ready) {
    SC_CUDA(cudaSetDevice(device));
    void* ps = nullptr;
    void* pt = nullptr;
    SC_CUDA(cudaGetSymbolAddress(&ps, g_noalloc_scratch));
    SC_CUDA(cudaGetSymbolAddress(&pt, g_noalloc_ticket));
    if (!ps || !pt) return fail(STAN_CUDA_ERR_CUDA, "stan_cuda_glm_launch_noalloc: cudaGetSymbolAddress returned null");
    nd.scratch = static_cast<double*>(ps);
    nd.ticket = static_cast<unsigned int*>(pt);
    nd.ready = true;
  }
  *out = &nd;
  return STAN_CUDA_OK;
}

}  // namespace stancuda

extern "C" int stan_cuda_glm_launch_noalloc(stan_cuda_data_t d, uint32_t flags, const double* d_params, double* d_result,
                                            void* stream) {
  using namespace stancuda;
  if (!d || !d_params || !d_result) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: null argument");
  if (d->family > STAN_CUDA_NORMAL_ID)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: bernoulli_logit, poisson_log and normal_id only");
  if (!d->colblocks.empty())
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: k <= 512 only (a wider matrix is evaluated as several launches "
                                   "with per-row temporaries in an allocated workspace)");
  if (d->t.layout != kLayoutWarp)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: the data block was created for another kernel variant");
  if (flags & STAN_CUDA_FUSED_ALLREDUCE)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: no collective on this entry point");
  // largest problem the fixed scratch serves: (CTAs + reduction groups) x (k_pad + 4) + 2 doubles
  const int64_t sms = sm_count(d->device);
  int64_t grid = sms * 2 < d->t.tiles ? sms * 2 : d->t.tiles;
  int64_t gs = 1;
  while (gs * gs < grid) ++gs;
  const size_t need = size_t(grid + (grid + gs - 1) / gs) * (size_t(d->t.kpad) + 4) + 2;
  if (need > kNoAllocScratchDoubles)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_launch_noalloc: the problem needs " + std::to_string(need) +
                                       " doubles of reduction scratch, the fixed scratch holds " +
                                       std::to_string(kNoAllocScratchDoubles) + " (k_pad <= 512 on at most 160 SMs)");
  NoAllocDevice* nd = nullptr;
  SC_TRY(noalloc_setup(d->device, &nd));
  SC_CUDA(cudaSetDevice(d->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the ticket words must start at zero; a launch leaves them at zero, but the array is shared and outlives aborted runs
  SC_CUDA(cudaMemsetAsync(nd->ticket, 0, kNoAllocTicketWords * sizeof(unsigned int), st));
  Workspace view;  // not owned: never freed, never grown
  view.device = d->device;
  view.scratch = nd->scratch;
  view.ticket = nd->ticket;
  view.scratch_doubles = kNoAllocScratchDoubles;
  view.fixed = true;
  LaunchIO io{flags, d_params, nullptr, nullptr, d_result, nullptr, nullptr};
  return launch_glm_warp(d, io, &view, st);
}

