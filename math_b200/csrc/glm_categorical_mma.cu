// glm_categorical_mma.cu -- categorical_logit_glm_lpmf value + gradient in one pass over X with the
// FP64 tensor-core instruction (mma.sync m8n8k4 f64, SASS DMMA) for both skinny GEMMs (sm_100a).
//
// Reference: stan/math/prim/mat/prob/categorical_logit_glm_lpmf.hpp:74-190
//   lin = x * beta + alpha'        (N x C; beta is K x C)                                  (:82-83)
//   logp = sum_i [ log(1 / sum_c exp(lin_ic - max_i)) - max_i + lin_i,y_i ]                (:84-104)
//   d_ic = [c == y_i] - softmax(lin_i)_c;  dalpha = colsum(d);  dbeta = x' * d             (:145-181)
//
// Why a second kernel: glm_categorical.cu feeds every FMA of `lin` with one shared-memory load of a
// weight (16 FMA / clk / SM at best) and ends up at 0.12-0.34 of the HBM roofline.  One DMMA does
// 8 x 8 x 4 = 256 FMAs from one 8-byte operand per lane and side, so the shared-memory traffic per
// FMA drops by 8x and the class count stops mattering below 8 (one n-tile) / 16 (two).
//
// Tile layout in HBM (kLayoutMma, written once at cache fill by data.cu): a tile holds R rows and
// kpad = 8 * ceil(K / 8) columns as R/8 x kpad/4 FRAGMENTS of 8 rows x 4 columns, 32 doubles each,
// element (r, col) at
//     x_dev[tile * R * kpad + ((r / 8) * (kpad / 4) + col / 4) * 32 + (r % 8) * 4 + col % 4]
// i.e. every fragment is stored in the lane order of the A operand of the m8n8k4 instruction
// (lane = row * 4 + k): one TMA bulk copy stages the tile, and
//   phase A  (lin = x * beta)    A fragment = x[8 rows][4 cols]    = fragment[lane]              one 256-byte LDS
//   phase B  (dbeta = x' * d)    A fragment = x'[8 cols][4 rows]   = two half fragments, lanes
//            (col = lane / 4, row = lane % 4) -> word ((col / 4) * 32 + (4 * half + row) * 4 + col % 4):
//            16 distinct 8-byte banks per half warp, conflict-free as well
// so the same staged bytes serve both products without a transpose.
//
// CTA = 8 warps, per tile:
//   A  warp (rb, ks): row block rb (8 rows), k-blocks of slice ks -> partial lin (C fragment) into shared
//      memory; weights sit in shared memory in B-fragment order, zero padded to 8 / 16 classes
//   S  256 / R threads per row: max, exp, sum, log-density term, d = onehot - softmax, written in the
//      B-fragment order of phase B; per-thread d/dalpha partials
//   B  warp -> m-tiles of 8 columns (and a slice of the rows when kpad < 64): accumulators
//      dbeta[8 cols][8 classes] per (m-tile, n-tile) stay in registers over the whole tile loop
// Two CTA barriers per tile; the stage of tile t is refilled after the first barrier of tile t + 1.
// Reductions run in a fixed order (finish_reduction): bit-reproducible.
// Limits (loud errors beyond): C <= 16 classes, kpad <= 256, C * kpad <= 4096.
#include <algorithm>

#include "glm_device.cuh"

namespace stancuda {

constexpr int kMmaThreads = 256;

// D(8x8) += A(8x4) * B(4x8), FP64.  Fragments (PTX ISA, mma.m8n8k4 .f64): a = A[lane / 4][lane % 4],
// b = B[lane % 4][lane / 4], {d0, d1} = D[lane / 4][2 * (lane % 4) + {0, 1}].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// (the softmax terms exp(lin - max) use exp_nonpos of glm_device.cuh)

// NTL: n-tiles of 8 classes (1 | 2).  MPW: m-tiles of 8 columns per warp in phase B (1 | 2 | 4).
// TPR: threads per row in phase S (4: 64 rows per tile, 8: 32 rows).  CPT: classes per thread in phase S,
// ceil(classes / TPR).  All four are functions of (k_pad, classes); eight combinations exist (launch below).
template <int NTL, int MPW, int TPR, int CPT>
__global__ void __launch_bounds__(kMmaThreads, TPR == 4 ? 3 : 2) glm_categorical_mma_kernel(const GlmKernelParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int NT = kMmaThreads;
  constexpr int ldl = NTL == 1 ? 8 : 24;  // row stride of the partial lin block (== 8 mod 16: conflict-free C-fragment stores)
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int R = NT / TPR;  // rows per tile: 64 | 32
  constexpr int RB = R >> 3;   // row blocks: 8 | 4
  constexpr int KS = 8 / RB;   // warps per row block in phase A (1 | 2), each a slice of the k-blocks
  const int kpad = p.kpad, ncls = p.ncls;
  const int kb = kpad >> 2;    // k-blocks of 4 columns
  const int MT = kpad >> 3;    // m-tiles of 8 columns
  const int cols = 4 + ncls + ncls * kpad;

  unsigned char* stage0 = smem;
  double* bfrag = reinterpret_cast<double*>(smem + size_t(p.stages) * p.stage_bytes);  // [kb][NTL][32]
  double* alpha_s = bfrag + size_t(kb) * NTL * 32;                                      // [16]
  double* linp = alpha_s + 16;                                                          // [KS * R = 64][ldl]
  double* dfrag = linp + 64 * ldl;                                                      // [R / 4][NTL][32]
  double* part = dfrag + size_t(R / 4) * NTL * 32;                                      // [NT * 4]
  uint64_t* full = reinterpret_cast<uint64_t*>(part + NT * 4);                          // [stages]
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
  // weights (class-major rows of kpad in the parameter block) into B-fragment order: element
  // (k, c) of k-block k / 4 and n-tile c / 8 sits at lane (c % 8) * 4 + k % 4; classes >= ncls are zero.
  // d starts as zero as well: phase S never writes the padding classes.
  for (int q = tid; q < kb * NTL * 32 + 16; q += NT) bfrag[q] = 0.0;
  for (int q = tid; q < (R / 4) * NTL * 32; q += NT) dfrag[q] = 0.0;
  __syncthreads();
  for (int q = tid; q < ncls * kpad; q += NT) {
    const int c = q / kpad, k = q - c * kpad;
    bfrag[(size_t(k >> 2) * NTL + (c >> 3)) * 32 + (c & 7) * 4 + (k & 3)] = p.params[q];
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
  // phase A: row block and k-slice of this warp
  const int rbk = warp % RB, ks = warp / RB;
  const int kbeg = ks * (kb / KS), kend = (ks == KS - 1) ? kb : kbeg + kb / KS;  // (kb is even)
  // phase S: row and class lane of this thread (class q, q + TPR, ...; at most 4 of them)
  const int srow = tid / TPR, sq = tid % TPR;
  // phase B: m-tiles mt0, mt0 + 8, ... of this warp; with fewer than 8 m-tiles the warps also split the rows
  int mt0, rs0, rs1;
  bool b_active = true;
  if (MT >= 8) {
    mt0 = warp;
    rs0 = 0;
    rs1 = R / 4;
  } else {
    const int RS = 8 / MT;                // row slices (R = 64 here: 16 row steps)
    const int per = (R / 4) / RS;
    const int rpart = warp / MT;
    mt0 = warp % MT;
    b_active = rpart < RS;
    rs0 = rpart * per;
    rs1 = (rpart == RS - 1) ? R / 4 : rs0 + per;
  }
  double g[MPW][NTL][2];
#pragma unroll
  for (int m = 0; m < MPW; ++m)
#pragma unroll
    for (int j = 0; j < NTL; ++j) g[m][j][0] = g[m][j][1] = 0.0;
  double da[CPT];
#pragma unroll
  for (int i = 0; i < CPT; ++i) da[i] = 0.0;
  double lp_acc = 0.0;
  // sum_rows log(1 / sum_c exp(.)) = -log(prod_rows sum_c exp(.)): the sums lie in [1, 16], so a running product of up
  // to 128 of them (< 2^512) replaces 128 logarithms by one
  double prod = 1.0;
  const double neg_inf = -__longlong_as_double(0x7ff0000000000000LL);

  for (int it = 0; it < ntl; ++it) {
    const int s = it % p.stages;
    const int64_t tile = tile0 + int64_t(it) * tstride;
    mbar_wait(smem_u32(&full[s]), (it / p.stages) & 1);
    const unsigned char* st = stage0 + size_t(s) * p.stage_bytes;
    const double* xt = reinterpret_cast<const double*>(st);
    // ---- A: partial lin of row block rbk over the k-blocks [kbeg, kend) -------------------------------
    {
      const double* xa = xt + (size_t(rbk) * kb + kbeg) * 32 + lane;
      const double* bf = bfrag + size_t(kbeg) * NTL * 32 + lane;
      double c[NTL][2][2];  // [n-tile][chain][2]: two independent accumulation chains per n-tile
#pragma unroll
      for (int j = 0; j < NTL; ++j) c[j][0][0] = c[j][0][1] = c[j][1][0] = c[j][1][1] = 0.0;
      int left = kend - kbeg;
#pragma unroll 2
      for (; left >= 2; left -= 2, xa += 64, bf += 2 * NTL * 32) {
        const double a0 = xa[0], a1 = xa[32];
#pragma unroll
        for (int j = 0; j < NTL; ++j) {
          dmma884(c[j][0][0], c[j][0][1], a0, bf[j * 32]);
          dmma884(c[j][1][0], c[j][1][1], a1, bf[(NTL + j) * 32]);
        }
      }
      if (left) {
        const double a0 = xa[0];
#pragma unroll
        for (int j = 0; j < NTL; ++j) dmma884(c[j][0][0], c[j][0][1], a0, bf[j * 32]);
      }
      double* lrow = linp + size_t(ks * R + rbk * 8 + (lane >> 2)) * ldl + 2 * (lane & 3);
#pragma unroll
      for (int j = 0; j < NTL; ++j)
        *reinterpret_cast<double2*>(lrow + j * 8) = make_double2(c[j][0][0] + c[j][1][0], c[j][0][1] + c[j][1][1]);
    }
    __syncthreads();  // partial lin complete; every warp is past phase B of the previous tile
    if (tid == 0 && it >= 1 && it - 1 + p.stages < ntl) issue(it - 1 + p.stages);  // refill the previous tile's stage
    // ---- S: softmax of row srow, classes sq, sq + TPR, ... (branch-free: a slot without a class carries -inf,
    // its exponential is 0 and its d is 0, written into the zero padding of the d fragment) ---------------
    {
      const int64_t left_rows = p.n - tile * R;
      const bool valid = srow < (left_rows < R ? int(left_rows) : R);
      const int yc = valid ? reinterpret_cast<const int*>(st + p.tile_bytes)[srow] - 1 : -1;  // 1 <= y <= ncls: checked at fill
      double v[CPT];
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        const int c = sq + TPR * i;  // < 8 * NTL: inside the (zero padded) row of linp
        double t = linp[srow * ldl + c];
        if (KS == 2) t += linp[(R + srow) * ldl + c];
        t += alpha_s[c];
        v[i] = (i < CPT - 1 || c < ncls) ? t : neg_inf;
      }
      double mx = v[0];                                                              // :84-85
#pragma unroll
      for (int i = 1; i < CPT; ++i) mx = v[i] > mx ? v[i] : mx;
#pragma unroll
      for (int off = 1; off < TPR; off <<= 1) {
        const double o = __shfl_xor_sync(0xffffffffu, mx, off);
        mx = o > mx ? o : mx;
      }
      double sum = 0.0, ly = 0.0;
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        ly = (sq + TPR * i == yc) ? v[i] : ly;
        v[i] = exp_nonpos(v[i] - mx);                                                // :88-89
        sum += v[i];
      }
#pragma unroll
      for (int off = 1; off < TPR; off <<= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, off);
        ly += __shfl_xor_sync(0xffffffffu, ly, off);
      }
      const double inv = 1.0 / sum;                                                  // :90-91
      if (valid && sq == 0) {                                                        // :93, :98-104: log(inv) - mx + ly
        lp_acc += ly - mx;
        prod *= sum;
      }
      if ((it & 127) == 127) {
        lp_acc -= log(prod);
        prod = 1.0;
      }
      // B fragment of phase B: rows are the contraction index (row % 4), classes the n index (c % 8)
      double* drow = dfrag + (srow >> 2) * (NTL * 32) + (srow & 3);
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        const int c = sq + TPR * i;
        const double dv = valid ? ((c == yc ? 1.0 : 0.0) - v[i] * inv) : 0.0;        // :147-148 + onehot (:157-159, :172-178)
        da[i] += dv;
        drow[(c >> 3) * 32 + (c & 7) * 4] = dv;
      }
    }
    __syncthreads();  // d complete
    // ---- B: dbeta[m-tile][n-tile] += x'[8 cols][4 rows] * d[4 rows][8 classes] over the row steps ------
    if (want_grad && b_active) {
      // lane (col = lane / 4, row = lane % 4) of the transposed fragment: k-block (col / 4) of the m-tile,
      // word (4 * (rs & 1) + row) * 4 + col % 4 of the row block rs / 2
      // (rs0 and rs1 are even: a row block of 8 rows is two row steps, always taken together)
      const double* xr = xt + size_t(rs0 >> 1) * kb * 32 + mt0 * 64 + (lane >> 4) * 32 + (lane & 3) * 4 + ((lane >> 2) & 3);
      const double* df = dfrag + size_t(rs0) * NTL * 32 + lane;
      const int xstep = kb * 32;
      for (int rs = rs0; rs < rs1; rs += 2, xr += xstep, df += 2 * NTL * 32) {
        double b0[NTL], b1[NTL];
#pragma unroll
        for (int j = 0; j < NTL; ++j) {
          b0[j] = df[j * 32];
          b1[j] = df[(NTL + j) * 32];
        }
#pragma unroll
        for (int m = 0; m < MPW; ++m) {
          if (m == 0 || mt0 + 8 * m < MT) {
            const double a0 = xr[m * 512], a1 = xr[m * 512 + 16];
#pragma unroll
            for (int j = 0; j < NTL; ++j) {
              dmma884(g[m][j][0], g[m][j][1], a0, b0[j]);
              dmma884(g[m][j][0], g[m][j][1], a1, b1[j]);
            }
          }
        }
      }
    }
  }
  __syncthreads();  // every stage has been consumed: the ring is scratch from here on

  // ---- CTA reduction -----------------------------------------------------------------------------------
  // dbeta: accumulator {g0, g1} of (m-tile mt, n-tile j) is dbeta[mt * 8 + lane / 4][j * 8 + 2 * (lane % 4) + {0, 1}]
  // -> rg[row slice][k][8 * NTL classes]; the row slices are then added in order
  double* rg = reinterpret_cast<double*>(stage0);
  const int RSn = MT >= 8 ? 1 : 8 / MT;
  if (b_active) {
    const int rpart = MT >= 8 ? 0 : warp / MT;
#pragma unroll
    for (int m = 0; m < MPW; ++m) {
      const int mt = mt0 + 8 * m;
      if (mt < MT) {
#pragma unroll
        for (int j = 0; j < NTL; ++j) {
          double* dst = rg + (size_t(rpart) * kpad + mt * 8 + (lane >> 2)) * (8 * NTL) + j * 8 + 2 * (lane & 3);
          dst[0] = g[m][j][0];
          dst[1] = g[m][j][1];
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) part[tid * 4 + i] = i < CPT ? da[i < CPT ? i : 0] : 0.0;
  __syncthreads();
  if (p.pdl) pdl_wait();  // first write to memory the previous evaluation also uses
  double* my_scratch = p.scratch + size_t(blockIdx.x) * cols;
  for (int idx = tid; idx < ncls * kpad; idx += NT) {
    const int c = idx / kpad, k = idx - c * kpad;
    double sum = 0.0;
    for (int q = 0; q < RSn; ++q) sum += rg[(size_t(q) * kpad + k) * (8 * NTL) + c];
    my_scratch[4 + ncls + idx] = sum;
  }
  if (tid < ncls) {  // d/dalpha of class tid: threads (row, tid % TPR), slot tid / TPR, rows in order
    const int q = tid % TPR, i = tid / TPR;
    double sum = 0.0;
    for (int r = 0; r < R; ++r) sum += part[(r * TPR + q) * 4 + i];
    my_scratch[4 + tid] = sum;
  }
  __syncthreads();  // part and rg are consumed
  part[tid] = lp_acc - log(prod);
  __syncthreads();
  if (tid == 0) {
    double sum = 0.0;
    for (int t = 0; t < NT; ++t) sum += part[t];
    my_scratch[0] = sum;
    my_scratch[1] = 0.0;
    my_scratch[2] = 0.0;
    my_scratch[3] = 0.0;
  }
  __syncthreads();
  double* fin = reinterpret_cast<double*>(stage0);  // [cols]
  finish_reduction<STAN_CUDA_CATEGORICAL_LOGIT, NT>(p, tid, cols, part, fin, &s_flag, [] { __syncthreads(); });
}

// =============================================================================================
static inline size_t align_up_mma(size_t v, size_t a) { return (v + a - 1) / a * a; }

static size_t mma_smem_bytes(const Tiling& t, int stages, int ntl, size_t* stage_bytes_out) {
  const size_t stage = align_up_mma(t.tile_bytes() + size_t(t.R) * sizeof(int), 128);
  if (stage_bytes_out) *stage_bytes_out = stage;
  const int ldl = ntl == 1 ? 8 : 24;
  const size_t doubles = size_t(t.kpad / 4) * ntl * 32 + 16 + size_t(64) * ldl + size_t(t.R / 4) * ntl * 32 +
                         size_t(kMmaThreads) * 4;
  return size_t(stages) * stage + doubles * sizeof(double) + size_t(stages) * sizeof(uint64_t) + 16;
}

template <int NTL, int MPW, int TPR, int CPT>
static int launch_mma_inst(const GlmKernelParams& p, int grid, size_t smem, cudaStream_t stream) {
  auto kern = glm_categorical_mma_kernel<NTL, MPW, TPR, CPT>;
  static std::atomic<size_t> configured[64];
  int dev = 0;
  SC_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && configured[dev].load(std::memory_order_relaxed) < smem) {
    SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev].store(smem, std::memory_order_relaxed);
  }
  SC_CUDA(launch_maybe_pdl(kern, dim3(unsigned(grid)), dim3(kMmaThreads), smem, stream, p.pdl != 0, p));
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

int launch_glm_categorical_mma(const stan_cuda_data* d, const LaunchIO& io, Workspace* ws, cudaStream_t stream) {
  const Tiling& t = d->t;
  if (d->n <= 0 || t.tiles <= 0) return fail(STAN_CUDA_ERR_ARG, "launch_glm: empty data block");
  if (t.layout != kLayoutMma || (t.kpad & 7) || t.R != (t.kpad <= 64 ? 64 : 32))
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit DMMA kernel needs the fragment tile layout");
  const int ncls = io.ncls;
  if (ncls < 1 || ncls > kMaxClasses || size_t(ncls) * t.kpad > size_t(kMaxClassWeights))
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit supports at most 16 classes and classes * k_pad <= 4096 "
                                   "on the device");
  if (!io.d_params) return fail(STAN_CUDA_ERR_ARG, "launch_glm: no parameter block");
  const int ntl = ncls <= 8 ? 1 : 2;
  const int MT = t.kpad / 8;
  if (MT > 32) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit DMMA kernel supports k_pad <= 256");
  const int sms = sm_count(d->device);
  // the kernel is latency-bound (dependent FP64 chains, three phases between barriers): as many CTAs per SM as
  // hold two stages each, then as many stages as still fit (<= 4).  The 64-row instantiations (k_pad <= 64) are
  // compiled for 80 registers, i.e. up to 3 CTAs per SM (K=32: 86 vs 95 us); the 32-row ones keep 128 registers
  // (their 32-64 KiB tiles allow 2 or 1 CTAs anyway, and the register cap cost them 4-9 %)
  int ctas_per_sm = t.R == 64 ? 3 : 2;
  size_t budget = 0;
  for (; ctas_per_sm >= 1; --ctas_per_sm) {
    budget = ctas_per_sm == 1 ? size_t(227) * 1024 : (size_t(228) * 1024) / ctas_per_sm - 1024 - 64;
    if (mma_smem_bytes(t, 2, ntl, nullptr) <= budget) break;
  }
  if (ctas_per_sm < 1) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit tile does not fit shared memory");
  if (tuning().ctas_per_sm > 0 && tuning().ctas_per_sm < ctas_per_sm) {
    ctas_per_sm = tuning().ctas_per_sm;
    budget = ctas_per_sm == 1 ? size_t(227) * 1024 : (size_t(228) * 1024) / ctas_per_sm - 1024 - 64;
  }
  int stages = 4;
  while (stages > 2 && mma_smem_bytes(t, stages, ntl, nullptr) > budget) --stages;
  size_t stage_bytes = 0;
  const size_t smem = mma_smem_bytes(t, stages, ntl, &stage_bytes);
  if (smem > size_t(227) * 1024) return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit tile does not fit shared memory");
  const size_t cols = 4 + size_t(ncls) * (t.kpad + 1);
  // the final reductions reuse the stage ring: [row slices][kpad][8 * ntl] partials, then `cols` doubles
  const size_t rs_n = MT >= 8 ? 1 : size_t(8 / MT);
  if (size_t(stages) * stage_bytes < std::max(cols, rs_n * t.kpad * 8 * ntl) * sizeof(double))
    return fail(STAN_CUDA_ERR_ARG, "launch_glm: categorical_logit result block does not fit the stage ring");

  int64_t grid64 = int64_t(sms) * ctas_per_sm;
  if (grid64 > t.tiles) grid64 = t.tiles;
  const bool pdl = pdl_enabled() && !io.shared_gpu;
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
  p.pdl = pdl ? 1 : 0;
  p.ncls = ncls;
  if (t.R == 64) {  // k_pad <= 64: one m-tile per warp, four threads per row in the softmax phase
    const int cpt = (ncls + 3) / 4;
    switch (cpt) {
      case 1: return launch_mma_inst<1, 1, 4, 1>(p, grid, smem, stream);
      case 2: return launch_mma_inst<1, 1, 4, 2>(p, grid, smem, stream);
      case 3: return launch_mma_inst<2, 1, 4, 3>(p, grid, smem, stream);
      default: return launch_mma_inst<2, 1, 4, 4>(p, grid, smem, stream);
    }
  }
  const bool wide = MT > 16;  // m-tiles per warp: 2 up to k_pad 128, 4 up to 256
  if (ntl == 1) return wide ? launch_mma_inst<1, 4, 8, 1>(p, grid, smem, stream) : launch_mma_inst<1, 2, 8, 1>(p, grid, smem, stream);
  return wide ? launch_mma_inst<2, 4, 8, 2>(p, grid, smem, stream) : launch_mma_inst<2, 2, 8, 2>(p, grid, smem, stream);
}

}  // namespace stancuda
