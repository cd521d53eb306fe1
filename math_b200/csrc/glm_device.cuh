// glm_device.cuh -- device-side pieces shared by the GLM kernels: kernel parameters, mbarrier / TMA
// PTX helpers, the family epilogues, the two-level fixed-order final reduction and the all-reduce
// over peer memory.
#pragma once
#include <math.h>

#include "context.cuh"

namespace stancuda {

struct GlmKernelParams {
  const double* x;
  const void* y;
  int64_t n;          // valid rows
  int tiles;
  int R, C, kc, kpad;
  int log2R;
  int stages;
  int group_size;        // CTAs per reduction group (two-level final sum)
  int blocked;           // warp kernel: 1 = CTA b owns tiles [b*per, (b+1)*per), 0 = round-robin
  uint32_t stage_bytes;  // per stage in shared memory (tile + y, 128-aligned)
  uint32_t tile_bytes;
  uint32_t y_bytes;      // per tile
  const double* params;
  const double* alpha_vec;
  const double* sigma_vec;
  double* row_dalpha;
  double* row_dsigma;
  double* scratch;
  unsigned int* ticket;
  double* result;
  uint32_t flags;
  double lgamma_sum;
  const double* nb_sums;  // neg_binomial_2_log with a histogram of y: [sum lgamma(y+phi), sum digamma(y+phi)]
  int nb_rowfun;          // neg_binomial_2_log: per-row lgamma/digamma (vector phi or no histogram)
  int ncuts;              // ordered_logistic: number of cut-points (classes - 1); 0 otherwise
  int ncls, cat_groups;   // categorical_logit: classes; class groups of the gradient phase (glm_categorical.cu)
  int pair;               // warp kernel: two tiles per iteration (glm_warp.cu, PAIR)
  int ord_private;        // ordered_logistic: d/dcuts bins private to each row slot of the tile ([ncuts][R] in shared memory)
  int pdl;                // launched with programmatic stream serialisation: see glm_warp.cu (early start of the next evaluation)
  // low-latency host path: parameters by value, completion flag in mapped host memory
  int inline_n;                        // > 0: read the parameter block from inline_params
  unsigned long long* host_flag;       // non-null: store host_seq there after the result block
  unsigned long long host_seq;
  double inline_params[kInlineParams];
  // all-reduce fused into the kernel (peer exchange over NVLink, comm.cu); world <= 1: off
  int xchg_world, xchg_rank;
  unsigned long long xchg_seq;
  double* xchg_local;
  double* xchg_peers[kXchgMaxWorld];
#ifdef STAN_CUDA_PHASE_TIMERS
  unsigned long long* timers;  // [grid][4] globaltimer stamps: start, first tile landed, loop end, CTA partial written; [grid*4] = end of last CTA
#endif
};

// Programmatic dependent launch (see glm_warp.cu): STAN_CUDA_PDL=0 switches it off, STAN_CUDA_PDL_SLACK = CTA slots a
// full-width grid leaves free for the previous evaluation's last CTA (default 1).
bool pdl_enabled();
int pdl_slack();
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename Kern, typename Params>
inline cudaError_t launch_maybe_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool pdl,
                                    const Params& p) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, p);
}

__device__ __forceinline__ double kernel_param(const GlmKernelParams& p, int i) {
  return p.inline_n > 0 ? p.inline_params[i] : p.params[i];
}

int fill_kernel_params(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, GlmKernelParams* p);
// batched launch (glm_warp.cu): plan the grid, then launch `count` problems of one device in one kernel
int glm_warp_batch_plan(const stan_cuda_data* d0, int count, int64_t min_tiles, int* grid_x, int* group_size,
                        size_t* scratch_doubles_per_problem);
int launch_glm_warp_batch(const stan_cuda_data* const* ds, const LaunchIO* ios, Workspace* wss, int count, int grid_x,
                          int group_size, GlmKernelParams* h_batch, const GlmKernelParams* d_batch, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + 1-D TMA bulk copy (SASS: UBLKCP / SYNCS)
// ---------------------------------------------------------------------------------------------
#ifdef STAN_CUDA_PHASE_TIMERS
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned int smid() {
  unsigned int v;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
  return v;
}
#define PHASE_STAMP(slot) do { if (tid == 0 && p.timers) { p.timers[size_t(blockIdx.x) * 4 + (slot)] = gtimer(); \
  if ((slot) == 0) p.timers[4 * 2048 + 8 + blockIdx.x] = smid(); } } while (0)
#else
#define PHASE_STAMP(slot) do { } while (0)
#endif
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                             uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// All-reduce of the result block fused into the kernel (replaces kernel -> ncclAllReduce).
// Runs in the last CTA only.  fin[0..cols) holds this device's block on entry and the sum over
// the world, added in rank order (bit-identical on every rank), on exit.
//   push : fin -> slot [seq & 1][rank] of EVERY device's exchange block (stores to peer-mapped
//          memory travel over NVLink), __threadfence_system, then one release-store of `seq` into
//          that device's flag [seq & 1][rank];
//   wait : acquire-loads of the world's flags in the own block until they show `seq`;
//   sum  : system-scope loads of the world's slots in rank order.
// Two slots suffice: a rank can be at most one collective ahead of the slowest rank, because it
// cannot finish collective s before every rank has pushed its block of collective s.
// A rank that does not show up within 30 s makes the result NaN (no hang).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long* ptr, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(ptr), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* ptr) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(ptr) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double* ptr) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(ptr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <int THREADS, typename Sync>
__device__ __forceinline__ void xchg_allreduce(const GlmKernelParams& p, double* fin, int cols, int tid,
                                               int* s_flag, Sync sync) {
  const int world = p.xchg_world, rank = p.xchg_rank;
  const unsigned long long seq = p.xchg_seq;
  const size_t slot = size_t(seq & 1ull) * world;
  const size_t flag0 = size_t(2) * world * kXchgCols;  // flags follow the slots (as 8-byte words)
  for (int idx = tid; idx < world * cols; idx += THREADS) {
    const int q = idx / cols, c = idx - q * cols;
    p.xchg_peers[q][(slot + rank) * kXchgCols + c] = fin[c];
  }
  __threadfence_system();
  sync();
  if (tid < world)
    st_release_sys(reinterpret_cast<unsigned long long*>(p.xchg_peers[tid] + flag0) + slot + rank, seq);
  if (tid == 0) *s_flag = 0;
  sync();
  if (tid < world) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(p.xchg_local + flag0) + slot + tid;
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(f) != seq) {
      if (global_timer_ns() - t0 > 30000000000ull) {
        atomicExch(s_flag, 1);
        break;
      }
    }
  }
  sync();
  const bool timed_out = *s_flag != 0;
  for (int c = tid; c < cols; c += THREADS) {
    double sum = 0.0;
    for (int q = 0; q < world; ++q) sum += ld_relaxed_sys(p.xchg_local + (slot + q) * kXchgCols + c);
    fin[c] = timed_out ? __longlong_as_double(0x7ff8dead00000000ll) : sum;
  }
  sync();
}

// ---------------------------------------------------------------------------------------------
// Lean special functions for the argument ranges the epilogues actually use.  The libdevice exp / log1p spend
// more instructions on materialising their constants (UMOV / IMAD.MOV pairs per coefficient) and on range
// handling these call sites cannot reach than on arithmetic -- measured on the categorical kernel: 64 % of its
// instructions were the softmax phase, 175 of 559 per thread and tile constant moves
// (profiles/r01_cat_mma_*_details.txt).  Coefficients come from the constant bank as DFMA operands.
// ---------------------------------------------------------------------------------------------
static __constant__ double kExpTaylor[14] = {1.0, 1.0, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040,
                                             1.0 / 40320, 1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800,
                                             1.0 / 479001600, 1.0 / 6227020800.0};
static __constant__ double kAtanhOdd[11] = {1.0, 1.0 / 3, 1.0 / 5, 1.0 / 7, 1.0 / 9, 1.0 / 11, 1.0 / 13, 1.0 / 15,
                                            1.0 / 17, 1.0 / 19, 1.0 / 21};

// exp(x) for x <= 0: x = n ln2 + r with |r| <= ln2 / 2, Taylor polynomial of degree 13 in r (remainder below
// 5e-18), 2^n by an add into the exponent field; <= 1 ulp against glibc on [-708, 0].  x < -708 (a result below
// 1e-307) gives 0; NaN stays NaN (the callers' finiteness witnesses rely on it).
__device__ __forceinline__ double exp_nonpos(double x) {
  const double t = fma(x, 1.4426950408889634074, 6755399441055744.0);  // low word of t = rint(x / ln2)
  const int n = __double2loint(t);
  const double fn = t - 6755399441055744.0;
  double r = fma(fn, -6.93147180369123816490e-01, x);  // ln2 split in two (fdlibm)
  r = fma(fn, -1.90821492927058770002e-10, r);
  double p = kExpTaylor[13];
#pragma unroll
  for (int i = 12; i >= 0; --i) p = fma(p, r, kExpTaylor[i]);
  const double v = __hiloint2double(__double2hiint(p) + n * 1048576, __double2loint(p));
  return x != x ? x : (x < -708.0 ? 0.0 : v);
}

// log1p(e) for 0 <= e <= 1 (e = exp(-|.|)): 1 + e = 2^k (1 + f) with k in {0, 1} and f in [-0.293, 0.414] formed
// exactly, log1p(f) = 2 atanh(s), s = f / (2 + f), |s| <= 0.172: odd series up to s^21 (next term below 3e-17);
// <= 2 ulp against glibc, exact 0 at 0, NaN stays NaN.
__device__ __forceinline__ double log1p_unit(double e) {
  const bool hi = e > 0.41421356237309503;
  const double f = hi ? (e - 1.0) * 0.5 : e;
  const double s = f / (2.0 + f);
  const double z = s * s;
  double r = kAtanhOdd[10];
#pragma unroll
  for (int i = 9; i >= 0; --i) r = fma(r, z, kAtanhOdd[i]);
  const double l = 2.0 * s * r;
  return hi ? (l + 1.90821492927058770002e-10) + 6.93147180369123816490e-01 : l;
}

// ---------------------------------------------------------------------------------------------
// family epilogues: per-row log-likelihood terms and d = dL/deta, same expressions and the same
// branch structure as the reference.
// ---------------------------------------------------------------------------------------------
struct RowTerms {
  double d;   // derivative wrt the linear predictor
  double t0;  // -> lp main term
  double t2;  // normal: s^2; neg_binomial_2_log: theta (finiteness witness)
  double t3;  // neg_binomial_2_log: per-row part of d/dphi
};

// digamma for x > 0: recurrence up to x >= 10, then the asymptotic series
// ln x - 1/(2x) - sum B_2k / (2k x^2k) with 7 terms (error < 1e-17 at x = 10).  The reference calls
// boost::math::digamma (stan/math/prim/scal/fun/digamma.hpp); agreement is pinned in the tests.
__host__ __device__ __forceinline__ double digamma_pos(double x) {
  double r = 0.0;
  while (x < 10.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  const double f = 1.0 / (x * x);
  const double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 +
                   f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + log(x) - 0.5 / x + t;
}

// neg_binomial_2_log_glm_lpmf.hpp:118-190, one row.  `rowfun`: lgamma / digamma of (y + phi) and
// phi are evaluated per row (vector phi, or no histogram of y); otherwise the caller adds their
// sums from the histogram of y (nb_hist_sums_kernel).  t0 is the row's contribution to the
// log-density under `flags`; t3 the row's contribution to d/dphi.
__device__ __forceinline__ RowTerms nb_row_epilogue(double eta, double alpha, double yv, double phi,
                                                    double log_phi, double inv_phi, uint32_t flags, bool rowfun,
                                                    bool ref_order) {
  RowTerms o;
  const double theta = eta + alpha;                                           // :118-119
  // one exponential serves the whole row: with u = theta - log phi and e = exp(-|u|),
  //   logsumexp(theta, log phi) = max + log1p(e)                               (:122-125)
  //   exp(theta) / (exp(theta) + phi) = sigmoid(u)  = (u > 0 ? 1 : e) / (1 + e)  (:159-162)
  //   phi / (exp(theta) + phi)        = sigmoid(-u) = (u > 0 ? e : 1) / (1 + e)  (:185-190)
  const double u = theta - log_phi;
  const bool pos = u > 0.0;
  const double e = exp_nonpos(-fabs(u));
  const double lse = (pos ? theta : log_phi) + log1p_unit(e);
  const double inv1pe = 1.0 / (1.0 + e);
  const double ypp = yv + phi;                                                // :127
  double q;                                                                   // (y + phi) / (exp(theta) + phi)
  if (ref_order) {
    // per-row outputs (d/dalpha_i, d/dphi_i) are differences of O(1) terms: keep the reference's own
    // operation order so that components that nearly cancel round the same way (:159-162, :180-183)
    const double te = exp(theta);
    q = ypp / (te + phi);
    o.d = yv - te * q;
  } else {
    o.d = yv - ypp * ((pos ? 1.0 : e) * inv1pe);                              // :162
    q = ypp * inv_phi * ((pos ? e : 1.0) * inv1pe);
  }
  double lp = 0.0;
  if (flags & STAN_CUDA_LP_MAIN) lp -= ypp * lse;                             // :145-147
  if (flags & STAN_CUDA_LP_LINEAR) lp += yv * theta;                          // :148-150
  double t3 = -q + log_phi - lse;                                             // :185-190
  if (rowfun) {
    if (flags & STAN_CUDA_LP_SCALE) lp += phi * log_phi - lgamma(phi) + lgamma(ypp);  // :137, :152
    if (flags & STAN_CUDA_NEED_DSIGMA) t3 += digamma_pos(ypp) - digamma_pos(phi);
  }
  o.t0 = lp;
  o.t2 = theta;
  o.t3 = t3;
  return o;
}

// ordered_logistic_glm_lpmf.hpp:104-171.  Everything that depends on the class alone is tabulated once
// per CTA in shared memory (ordered_class_table): for class c (1 .. ncuts + 1), with cy1 = cuts[c-1]
// (+inf for the last class) and cy2 = cuts[c-2] (-inf for the first),
//   tab[0][c-1] = cy1, tab[1][c-1] = cy2,
//   tab[2][c-1] = ecd = exp(cy2 - cy1)                         (:164; 0 at the two ends)
//   tab[3][c-1] = 1 / (1 - ecd)                                (:165-171)
//   tab[4][c-1] = log1m_exp(cy2 - cy1)                         (:141; the reference evaluates it at
//                 (loc - cy1) - (loc - cy2), the same number up to rounding of the location; 0 at the ends)
// which leaves two exponentials, two log1p and two divisions per row.
constexpr int kOrdTabRows = 5;
__device__ __forceinline__ double log1m_exp_dev(double a) {  // prim/scal/fun/log1m_exp.hpp:44-54
  if (a >= 0) return nan("");
  return a > -0.693147 ? log(-expm1(a)) : log1p(-exp(a));
}
template <int NT>
__device__ __forceinline__ void ordered_class_table(const GlmKernelParams& p, double* tab, int tid) {
  const int ncls = p.ncuts + 1;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  for (int c = 1 + tid; c <= ncls; c += NT) {
    const bool first = c == 1, last = c == ncls;
    const double cy1 = last ? inf : kernel_param(p, p.kpad + 2 + c - 1);
    const double cy2 = first ? -inf : kernel_param(p, p.kpad + 2 + c - 2);
    const bool inner = !first && !last;
    const double ecd = inner ? exp(cy2 - cy1) : 0.0;
    tab[0 * ncls + c - 1] = cy1;
    tab[1 * ncls + c - 1] = cy2;
    tab[2 * ncls + c - 1] = ecd;
    tab[3 * ncls + c - 1] = 1.0 / (1.0 - ecd);
    tab[4 * ncls + c - 1] = inner ? log1m_exp_dev(cy2 - cy1) : 0.0;
  }
}
// one row of a class with cut-points cy1 above and cy2 below it (and ecd, r = 1 / (1 - ecd), l1m = log1m_exp(cy2 - cy1) of
// the class): t0 = log-density of the row, d = d1 - d2 (location derivative), t2 = d1, t3 = d2 (the row's contributions
// to d/dcuts[c-2] and d/dcuts[c-1], :192-200)
__device__ __forceinline__ RowTerms ordered_row_terms(double loc, bool first, bool last, double cy1, double cy2, double ecd,
                                                      double r, double l1m) {
  RowTerms o;
  const double cut1 = loc - cy1, cut2 = loc - cy2;  // :124-125
  // exp(-|cut|) serves both the log-density (:128-131) and the derivative: for cut > 0 the reference
  // uses exp(-cut) / (1 + exp(-cut)), else 1 / (1 + exp(cut)) (:165-171) -- the same exponential
  const double e1 = exp_nonpos(-fabs(cut1)), e2 = exp_nonpos(-fabs(cut2));
  const double m1 = (cut1 > 0.0 ? -cut1 : 0.0) - log1p_unit(e1);                     // :128-129
  const double m2 = (cut2 <= 0.0 ? cut2 : 0.0) - log1p_unit(e2);                     // :130-131
  o.t0 = first ? m1 : (last ? m2 : m2 + l1m + m1);                                    // :137-144
  const double s2 = (cut2 > 0 ? e2 : 1.0) / (1 + e2);                                 // :165-167
  const double s1 = (cut1 > 0 ? e1 : 1.0) / (1 + e1);                                 // :168-171
  const double d1 = s2 + ecd * r;  // s2 - ecd / (ecd - 1)
  const double d2 = r - s1;        // 1 / (1 - ecd) - s1
  o.d = d1 - d2;                                                                      // :173
  o.t2 = d1;
  o.t3 = d2;
  return o;
}
// the same from the per-CTA class table
__device__ __forceinline__ RowTerms ordered_row_epilogue(double loc, int c, int ncls, const double* tab) {
  return ordered_row_terms(loc, c == 1, c == ncls, tab[0 * ncls + c - 1], tab[1 * ncls + c - 1], tab[2 * ncls + c - 1],
                           tab[3 * ncls + c - 1], tab[4 * ncls + c - 1]);
}

// d/dcuts bins owned by ONE warp: the active lanes add `val` into bins[key].  Lanes of the warp that hit the same bin are
// found with match.any; every member adds the group's values in lane order (so the sum does not depend on timing) and
// the lowest lane of the group makes the one update of that bin -- no floating-point atomics.  All 32 lanes must call.
__device__ __forceinline__ void warp_bin_add(double* bins, int lane, bool active, int key, double val) {
  const int mykey = active ? key : -1 - lane;  // inactive lanes form groups of their own
  const unsigned grp = __match_any_sync(0xffffffffu, mykey);
  unsigned rem = grp;
  double acc = 0.0;
  while (__any_sync(0xffffffffu, rem != 0u)) {
    const int src = rem ? __ffs(rem) - 1 : lane;
    const double v = __shfl_sync(0xffffffffu, val, src);
    if (rem) {
      acc += v;
      rem &= rem - 1;
    }
  }
  if (active && lane == __ffs(grp) - 1) bins[key] += acc;
  __syncwarp();
}

template <int FAMILY>
__device__ __forceinline__ RowTerms row_epilogue(double eta, double alpha, double yv, double inv_sigma) {
  RowTerms o;
  if (FAMILY == STAN_CUDA_BERNOULLI_LOGIT) {
    // bernoulli_logit_glm_lpmf.hpp:90-104,115-121 (cutoff = 20, the three-way select)
    const double sign = 2.0 * yv - 1.0;
    const double ytheta = sign * (eta + alpha);
    const double e = exp(-ytheta);
    o.t0 = ytheta > 20.0 ? -e : (ytheta < -20.0 ? ytheta : -log1p(e));
    o.d = ytheta > 20.0 ? -e : (ytheta < -20.0 ? sign : sign * e / (e + 1.0));
    o.t2 = 0.0;
  } else if (FAMILY == STAN_CUDA_POISSON_LOG) {
    // poisson_log_glm_lpmf.hpp:91-95,110-111
    const double theta = eta + alpha;
    const double e = exp(theta);
    o.d = yv - e;
    o.t0 = yv * theta - e;
    o.t2 = 0.0;
  } else {
    // normal_id_glm_lpdf.hpp:108-117,138-141
    const double s = (yv - eta - alpha) * inv_sigma;
    o.d = inv_sigma * s;
    o.t2 = s * s;
    o.t0 = 0.0;
  }
  o.t3 = 0.0;
  return o;
}

// ---------------------------------------------------------------------------------------------
// End of kernel, after every CTA has written its (cols)-wide partial to p.scratch[blockIdx.x]:
// two-level fixed-order reduction of the CTA partials, final scalars, optional all-reduce over
// the devices, result block.  `sync` is the CTA-wide barrier of the NT participating threads.
//   CTAs are grouped in blocks of `group_size` consecutive CTAs.  The last CTA of a group to
//   finish adds the group's partials in CTA order; the last group to finish adds the group sums
//   in group order.  The grouping depends on the grid only, so the result is bit-reproducible,
//   and the critical path after the slowest CTA is two short sums instead of one of `grid` terms.
// part: shared scratch of >= 4 doubles; fin: shared scratch of >= cols doubles; s_flag: shared int.
// ---------------------------------------------------------------------------------------------
template <int FAMILY, int NT, typename Sync>
__device__ __forceinline__ void finish_reduction(const GlmKernelParams& p, int tid, int cols, double* part,
                                                 double* fin, int* s_flag, Sync sync) {
  const int grid = gridDim.x;
  __threadfence();
  sync();
  PHASE_STAMP(3);
  const int gsz = p.group_size, ngroups = (grid + gsz - 1) / gsz;
  const int g = int(blockIdx.x) / gsz;
  const int gfirst = g * gsz, glast = (gfirst + gsz < grid) ? gfirst + gsz : grid;
  if (tid == 0) {
    const unsigned int t = atomicAdd(p.ticket + 1 + g, 1u);
    *s_flag = (t == unsigned(glast - gfirst) - 1u);
  }
  sync();
  if (!*s_flag) return;
  __threadfence();
  double* gscratch = p.scratch + size_t(grid) * cols;  // [ngroups][cols]
  for (int c = tid; c < cols; c += NT) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int b = gfirst;
    for (; b + 4 <= glast; b += 4) {
      s0 += __ldcg(p.scratch + size_t(b + 0) * cols + c);
      s1 += __ldcg(p.scratch + size_t(b + 1) * cols + c);
      s2 += __ldcg(p.scratch + size_t(b + 2) * cols + c);
      s3 += __ldcg(p.scratch + size_t(b + 3) * cols + c);
    }
    for (; b < glast; ++b) s0 += __ldcg(p.scratch + size_t(b) * cols + c);
    gscratch[size_t(g) * cols + c] = (s0 + s1) + (s2 + s3);
  }
  __threadfence();
  sync();
  if (tid == 0) {
    p.ticket[1 + g] = 0u;  // every CTA of the group has taken its ticket: ready for the next launch
    const unsigned int t = atomicAdd(p.ticket, 1u);
    *s_flag = (t == unsigned(ngroups) - 1u);
  }
  sync();
  if (!*s_flag) return;
  __threadfence();
  for (int c = tid; c < cols; c += NT) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int b = 0;
    for (; b + 4 <= ngroups; b += 4) {
      s0 += __ldcg(gscratch + size_t(b + 0) * cols + c);
      s1 += __ldcg(gscratch + size_t(b + 1) * cols + c);
      s2 += __ldcg(gscratch + size_t(b + 2) * cols + c);
      s3 += __ldcg(gscratch + size_t(b + 3) * cols + c);
    }
    for (; b < ngroups; ++b) s0 += __ldcg(gscratch + size_t(b) * cols + c);
    const double tot = (s0 + s1) + (s2 + s3);
    if (c >= 4) fin[c] = tot;
    else part[c] = tot;
  }
  sync();
  if (tid == 0) {
    const double A0 = part[0], A1 = part[1], A2 = part[2], A3 = part[3];
    const double nn = double(p.n);
    const double sigma0 = kernel_param(p, p.kpad + 1);
    double lp = 0.0, dsig = 0.0, witness;
    if (FAMILY == STAN_CUDA_BERNOULLI_LOGIT) {
      if (p.flags & STAN_CUDA_LP_MAIN) lp = A0;
      witness = A0;  // bernoulli_logit_glm_lpmf.hpp:106 tests logp
    } else if (FAMILY == STAN_CUDA_POISSON_LOG) {
      if (p.flags & STAN_CUDA_LP_CONST) lp -= p.lgamma_sum;  // :102-108
      if (p.flags & STAN_CUDA_LP_MAIN) lp += A0;             // :109-112
      witness = A1;  // poisson_log_glm_lpmf.hpp:97 tests sum(theta_derivative)
    } else if (FAMILY == kFamilyLinearBlock) {
      witness = 0.0;  // a column block of the k > 512 path: no log-density of its own
    } else if (FAMILY == STAN_CUDA_CATEGORICAL_LOGIT) {
      if (p.flags & STAN_CUDA_LP_MAIN) lp = A0;
      witness = A0;  // categorical_logit_glm_lpmf.hpp:109 tests logp
    } else if (FAMILY == STAN_CUDA_ORDERED_LOGISTIC) {
      if (p.flags & STAN_CUDA_LP_MAIN) lp = A0;
      witness = A2;  // ordered_logistic_glm_lpmf.hpp:119 tests sum(location)
    } else if (FAMILY == STAN_CUDA_NEG_BINOMIAL_2_LOG) {
      // neg_binomial_2_log_glm_lpmf.hpp:130-153, :185-190; A0 already holds the flagged row terms
      const double phi = sigma0;
      lp = A0;
      if (p.flags & STAN_CUDA_LP_CONST) lp -= p.lgamma_sum;
      if (!p.nb_rowfun) {
        if (p.flags & STAN_CUDA_LP_SCALE) lp += nn * (phi * log(phi) - lgamma(phi)) + p.nb_sums[0];
        dsig = nn + A3 + p.nb_sums[1] - nn * digamma_pos(phi);
      } else if (!p.sigma_vec) {
        dsig = nn + A3;
      }
      witness = A2;  // :120 check_finite(theta)
    } else {
      // normal_id_glm_lpdf.hpp:159-172
      const double NEG_LOG_SQRT_TWO_PI = -0.91893853320467274178032973640561763986;
      if (p.flags & STAN_CUDA_LP_CONST) lp += NEG_LOG_SQRT_TWO_PI * nn;
      if (p.flags & STAN_CUDA_LP_SCALE) lp -= p.sigma_vec ? A3 : nn * log(sigma0);
      if (p.flags & STAN_CUDA_LP_MAIN) lp -= 0.5 * A2;
      if (!p.sigma_vec) dsig = (A2 - nn) * (1.0 / sigma0);  // :142-143
      witness = A2;  // :150 tests y_scaled_sq_sum
    }
    fin[0] = lp;
    fin[1] = A1;
    fin[2] = dsig;
    fin[3] = witness;
    *p.ticket = 0u;  // ready for the next launch on this workspace (stream ordered)
  }
  sync();
  // every term of the block is additive over row shards: sum it over the devices right here
  if (p.xchg_world > 1) xchg_allreduce<NT>(p, fin, cols, tid, s_flag, sync);
  for (int c = tid; c < cols; c += NT) p.result[c] = fin[c];
  if (p.host_flag) {  // result block is in mapped host memory: publish it
    __threadfence_system();
    sync();
    if (tid == 0) st_release_sys(p.host_flag, p.host_seq);
  }
#ifdef STAN_CUDA_PHASE_TIMERS
  sync();
  if (tid == 0 && p.timers) p.timers[size_t(grid) * 4] = gtimer();
#endif
}

}  // namespace stancuda
