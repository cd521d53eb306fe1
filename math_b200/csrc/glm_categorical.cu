// glm_categorical.cu -- categorical_logit_glm_lpmf value + gradient in one pass over X (sm_100a, FP64).
//
// Reference: stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:74-190
//   lin = x * beta + alpha'        (N x C; beta is K x C)
//   logp = sum_i [ log(1 / sum_c exp(lin_ic - max_i)) - max_i + lin_i,y_i ]            (:82-104)
//   d_ic = [c == y_i] - softmax(lin_i)_c;  dalpha = colsum(d);  dbeta = x' * d          (:145-181)
// The reference makes two GEMM passes over x and keeps four N x C temporaries.  Here a tile of rows
// is staged once by TMA (same tile layout and ring as glm_warp.cu) and used three times from shared
// memory:
//   A  every warp, for its own rows: lin_c = <x_row, beta_c> + alpha_c for all classes (beta in shared
//      memory, partial dots of the column groups meet by xor-shuffle) -> lin[R][C] in shared memory
//   S  one thread per row: max, exp, sum, log-density term, d = onehot - softmax   (in place)
//   B  gradient x' * d with register accumulators: the 8 warps are G class groups x 8/G row groups;
//      a warp re-reads the x slices of its row group and accumulates kc x CPG partials per lane
//      (CPG = classes per group, 1 or 2: 16 x CPG doubles per lane is what the register file allows)
// Three CTA barriers per tile; two CTAs per SM overlap each other's phases.  All reductions run in a
// fixed order (shuffle trees, warps in order, finish_reduction): bit-reproducible.
// Limits of the device path (loud errors beyond): C <= 16 classes and C * k_pad <= 4096.
#include <algorithm>

#include "glm_device.cuh"

namespace stancuda {

constexpr int kCatWarps = 8;
constexpr int kCatThreads = kCatWarps * 32;

template <int CPG>
__global__ void __launch_bounds__(kCatThreads, 2) glm_categorical_kernel(const GlmKernelParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int NT = kCatThreads;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int C = p.C, kc = p.kc, kpad = p.kpad, R = p.R;
  const int ncls = p.ncls, G = p.cat_groups, NQ = kCatWarps / G;
  const int rpw = 32 / C;
  const int cg = lane & (C - 1);
  const int rl = lane / C;
  const int ls = ncls | 1;          // row stride of lin (odd: conflict-free column walks)
  const int bstride = C * (kc + 1);  // row stride of beta in shared memory (column groups 17 doubles apart)
  const int cols = 4 + ncls + ncls * kpad;

  unsigned char* stage0 = smem;
  double* bs = reinterpret_cast<double*>(smem + size_t(p.stages) * p.stage_bytes);  // [ncls][bstride]
  double* alpha_s = bs + size_t(ncls) * bstride + kMaxKc;                            // [ncls] (16 zeros behind bs)
  double* lin = alpha_s + ncls;                                                      // [R][ls]
  double* part = lin + size_t(R) * ls;                                               // [NT] (lp partials), first 8 reused
  uint64_t* full = reinterpret_cast<uint64_t*>(part + NT);                           // [stages]
  __shared__ int s_flag;

  if (p.pdl) pdl_launch_dependents();  // the next evaluation may start under this one's tail (see glm_warp.cu)
  const int grid = gridDim.x;
  const int64_t tile0 = blockIdx.x;
  const int64_t tstride = grid;
  const int ntl = (p.tiles - int(blockIdx.x) + grid - 1) / grid;

  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) mbar_init(smem_u32(&full[s]), 1);
    mbar_fence_init();
  }
  // weights (class-major rows of k_pad in the parameter block) and intercepts into shared memory
  for (int q = tid; q < ncls * bstride + kMaxKc; q += NT) bs[q] = 0.0;
  __syncthreads();
  for (int q = tid; q < ncls * kpad; q += NT) {
    const int c = q / kpad, k = q - c * kpad;
    bs[size_t(c) * bstride + (k / kc) * (kc + 1) + (k % kc)] = p.params[q];
  }
  for (int c = tid; c < ncls; c += NT) alpha_s[c] = p.params[size_t(ncls) * kpad + c];
  __syncthreads();

  const uint32_t tx_bytes = p.tile_bytes + p.y_bytes;
  const unsigned char* xbytes = reinterpret_cast<const unsigned char*>(p.x);
  const unsigned char* ybytes = reinterpret_cast<const unsigned char*>(p.y);
  uint64_t policy = 0;
  auto issue = [&](int t) {
    const int s = t % p.stages;
    const int64_t tile = tile0 + int64_t(t) * tstride;
    const uint32_t bar = smem_u32(&full[s]);
    const uint32_t dst = smem_u32(stage0 + size_t(s) * p.stage_bytes);
    mbar_expect_tx(bar, tx_bytes);
    tma_bulk_g2s(dst, xbytes + tile * p.tile_bytes, p.tile_bytes, bar, policy);
    tma_bulk_g2s(dst + p.tile_bytes, ybytes + tile * p.y_bytes, p.y_bytes, bar, policy);
  };
  if (tid == 0) {
    policy = l2_evict_first_policy();
    const int pre = ntl < p.stages ? ntl : p.stages;
    for (int t = 0; t < pre; ++t) issue(t);
  }

  const bool want_grad = (p.flags & (STAN_CUDA_NEED_DBETA | STAN_CUDA_NEED_DALPHA)) != 0;
  const int g = warp & (G - 1), q0 = warp / G;  // class group and row group of this warp (phase B)
  double acc[CPG][kMaxKc], da[CPG];
#pragma unroll
  for (int cl = 0; cl < CPG; ++cl) {
    da[cl] = 0.0;
#pragma unroll
    for (int j = 0; j < kMaxKc; ++j) acc[cl][j] = 0.0;
  }
  double lp_acc = 0.0;
  const int r_own = warp * rpw + rl;  // this lane's row of the tile in phase A

  for (int it = 0; it < ntl; ++it) {
    const int s = it % p.stages;
    const int64_t tile = tile0 + int64_t(it) * tstride;
    mbar_wait(smem_u32(&full[s]), (it / p.stages) & 1);
    const unsigned char* st = stage0 + size_t(s) * p.stage_bytes;
    const double* xt = reinterpret_cast<const double*>(st);
    // ---- A + S inside the warp: lin = x * beta + alpha for the rows of this warp, then the softmax of
    // each row split over the C lanes that share it (lane cg takes classes cg, cg + C, ...) --------------
    {
      const double* xs = xt + warp * 32 + lane;
      double xr[kMaxKc];
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j) xr[j] = j < kc ? xs[j * NT] : 0.0;
      const double* bl = bs + cg * (kc + 1);
      double* l = lin + r_own * ls;
      const int64_t row = tile * R + r_own;
      const bool valid = row < p.n;
      const int yc = valid ? reinterpret_cast<const int*>(st + p.tile_bytes)[r_own] - 1 : 0;  // 1 <= y <= ncls: checked at fill
      double mx = -__longlong_as_double(0x7ff0000000000000LL), ly = 0.0;
      // two classes per trip: their dot products are independent chains (the class loop is latency-bound)
      for (int c = 0; c < ncls; c += 2) {
        const bool two = c + 1 < ncls;
        const double* b0 = bl + size_t(c) * bstride;
        const double* b1 = two ? b0 + bstride : b0;
        double p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
#pragma unroll
        for (int j = 0; j < kMaxKc; j += 2) {  // j >= kc: x is 0 and b reads zero padding (or a finite neighbour)
          p0 = fma(xr[j], b0[j], p0);
          q0 = fma(xr[j], b1[j], q0);
          p1 = fma(xr[j + 1], b0[j + 1], p1);
          q1 = fma(xr[j + 1], b1[j + 1], q1);
        }
        double v = p0 + p1, w = q0 + q1;
        for (int off = 1; off < C; off <<= 1) {
          v += __shfl_xor_sync(0xffffffffu, v, off);
          w += __shfl_xor_sync(0xffffffffu, w, off);
        }
        v += alpha_s[c];
        mx = fmax(mx, v);                                                            // :84-85 (every lane of the row)
        if (c == yc) ly = v;
        if (cg == (c & (C - 1))) l[c] = v;
        if (two) {
          w += alpha_s[c + 1];
          mx = fmax(mx, w);
          if (c + 1 == yc) ly = w;
          if (cg == ((c + 1) & (C - 1))) l[c + 1] = w;
        }
      }
      double sum = 0.0;
      for (int c = cg; c < ncls; c += C) {                                           // :88-89
        const double e = exp(l[c] - mx);  // this lane wrote l[c] itself and is the only one to touch it
        sum += e;
        l[c] = e;
      }
      for (int off = 1; off < C; off <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      const double inv = 1.0 / sum;                                                  // :90-91
      if (valid && cg == 0) lp_acc += log(inv) - mx + ly;                            // :93, :98-104
      for (int c = cg; c < ncls; c += C)
        l[c] = valid ? ((c == yc ? 1.0 : 0.0) - l[c] * inv) : 0.0;                   // :147-148 + onehot (:157-159, :172-178)
    }
    __syncthreads();
    // ---- B: gradient x' * d, classes of group g, x slices of row group q0 --------------------------
    if (want_grad) {
      for (int sl = q0; sl < kCatWarps; sl += NQ) {
        const double* xs = xt + sl * 32 + lane;
        double xr[kMaxKc];
#pragma unroll
        for (int j = 0; j < kMaxKc; ++j) xr[j] = j < kc ? xs[j * NT] : 0.0;
        const double* l = lin + (sl * rpw + rl) * ls + g * CPG;
#pragma unroll
        for (int cl = 0; cl < CPG; ++cl) {
          const double dv = (g * CPG + cl < ncls) ? l[cl] : 0.0;
          if (cg == 0) da[cl] += dv;
#pragma unroll
          for (int j = 0; j < kMaxKc; ++j) acc[cl][j] = fma(xr[j], dv, acc[cl][j]);
        }
      }
    }
    __syncthreads();  // stage s and lin are free again
    if (tid == 0 && it + p.stages < ntl) issue(it + p.stages);
  }

  // ---- CTA reduction ------------------------------------------------------------------------------
  // rows of a warp (xor offsets C .. 16); lanes rl == 0 then hold the sums of column group cg
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
    for (int cl = 0; cl < CPG; ++cl) {
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j) {
        const double o = __shfl_xor_sync(0xffffffffu, acc[cl][j], off);
        if (off >= C) acc[cl][j] += o;
      }
      const double o = __shfl_xor_sync(0xffffffffu, da[cl], off);
      if (off >= C) da[cl] += o;
    }
  }
  part[tid] = lp_acc;
  double* red = reinterpret_cast<double*>(stage0);  // [warps][CPG][kpad + 1]; every stage has been consumed
  const int rstride = CPG * (kpad + 1);
  if (rl == 0) {
#pragma unroll
    for (int cl = 0; cl < CPG; ++cl) {
#pragma unroll
      for (int j = 0; j < kMaxKc; ++j)
        if (j < kc) red[warp * rstride + cl * (kpad + 1) + cg * kc + j] = acc[cl][j];
      if (cg == 0) red[warp * rstride + cl * (kpad + 1) + kpad] = da[cl];
    }
  }
  __syncthreads();
  if (p.pdl) pdl_wait();  // first write to memory the previous evaluation also uses
  double* my_scratch = p.scratch + size_t(blockIdx.x) * cols;
  if (tid == 0) {
    double sum = 0.0;
    for (int t = 0; t < NT; ++t) sum += part[t];
    my_scratch[0] = sum;
    my_scratch[1] = 0.0;
    my_scratch[2] = 0.0;
    my_scratch[3] = 0.0;
  }
  for (int idx = tid; idx < ncls * (kpad + 1); idx += NT) {
    const int c = idx / (kpad + 1), k = idx - c * (kpad + 1);  // k == kpad: the class's d/dalpha
    const int gg = c / CPG, cl = c - gg * CPG;
    double sum = 0.0;
    for (int q = 0; q < NQ; ++q) sum += red[(q * G + gg) * rstride + cl * (kpad + 1) + k];
    if (k == kpad) my_scratch[4 + c] = sum;
    else my_scratch[4 + ncls + c * kpad + k] = sum;
  }
  __syncthreads();  // red is consumed before finish_reduction reuses the stage ring
  double* fin = reinterpret_cast<double*>(stage0);  // [cols]
  finish_reduction<STAN_CUDA_CATEGORICAL_LOGIT, NT>(p, tid, cols, part, fin, &s_flag, [] { __syncthreads(); });
}

// =============================================================================================
static inline size_t align_up_cat(size_t v, size_t a) { return (v + a - 1) / a * a; }

static size_t cat_smem_bytes(const Tiling& t, int stages, int ncls, size_t* stage_bytes_out) {
  const size_t stage = align_up_cat(t.tile_bytes() + size_t(t.R) * sizeof(double), 128);
  if (stage_bytes_out) *stage_bytes_out = stage;
  const size_t doubles = size_t(ncls) * t.C * (t.kc + 1) + kMaxKc + ncls + size_t(t.R) * (ncls | 1) + kCatThreads;
  return size_t(stages) * stage + doubles * sizeof(double) + size_t(stages) * sizeof(uint64_t) + 16;
}

template <int CPG>
static int launch_cat_inst(const GlmKernelParams& p, int grid, size_t smem, cudaStream_t stream) {
  auto kern = glm_categorical_kernel<CPG>;
  static std::atomic<size_t> configured[64];
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && configured[dev].load(std::memory_order_relaxed) < smem) {
    SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev].store(smem, std::memory_order_relaxed);
  }
  SC_CUDA(launch_maybe_pdl(kern, dim3(unsigned(grid)), dim3(kCatThreads), smem, stream, p.pdl != 0, p));
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

int launch_glm_categorical(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  const Tiling& t = d->t;
  if (d->n <= 0 || t.tiles <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_glm: empty data block");
  if (t.layout != kLayoutWarp) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit needs the warp tile layout");
  const int ncls = io.ncls;
  if (ncls < 1 || ncls > kMaxClasses || size_t(ncls) * t.kpad > size_t(kMaxClassWeights))
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit supports at most 16 classes and classes * k_pad <= 4096 "
                                   "on the device");
  if (!io.d_params) return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  const int sms = sm_count(d->device);
  const size_t budget = (size_t(228) * 1024) / 2 - 1024 - 64;
  int stages = 4;
  size_t stage_bytes = 0;
  while (stages > 2 && cat_smem_bytes(t, stages, ncls, nullptr) > budget) --stages;
  size_t smem = cat_smem_bytes(t, stages, ncls, &stage_bytes);
  if (smem > size_t(227) * 1024) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit tile does not fit shared memory");
  const size_t cols = 4 + size_t(ncls) * (t.kpad + 1);
  // the final reductions reuse the stage ring: [8][CPG][kpad + 1] partials, then `cols` doubles
  if (size_t(stages) * stage_bytes < std::max(cols, size_t(kCatWarps) * 2 * (t.kpad + 1)) * sizeof(double))
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit result block does not fit the stage ring");

  int64_t grid64 = int64_t(sms) * 2;
  if (grid64 > t.tiles) grid64 = t.tiles;
  const bool pdl = pdl_enabled() && !io.shared_gpu;
  if (pdl && grid64 == int64_t(sms) * 2 && grid64 > pdl_slack() + 8) grid64 -= pdl_slack();
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
  p.pdl = pdl ? 1 : 0;
  p.ncls = ncls;
  int G = 1;  // class groups: power of two <= 8 with G * CPG >= ncls, CPG <= 2
  while (G < kCatWarps && G * 2 < ncls) G <<= 1;
  p.cat_groups = G;
  if (G >= ncls) return launch_cat_inst<1>(p, grid, smem, stream);
  return launch_cat_inst<2>(p, grid, smem, stream);
}

}  // namespace stancuda
