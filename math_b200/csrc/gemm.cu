// gemm.cu -- dense FP64 matrices on the device and their product.
//
// Replaces, for the `multiply` hooks of the reference (stan/math/prim/mat/fun/multiply.hpp:71-79,
// stan/math/rev/mat/fun/multiply.hpp:88-99, 111-123, 274-283, 296-303, 444-453, 466-473), what its OpenCL backend
// provides there: matrix_cl<double> (stan/math/opencl/matrix_cl.hpp:32-454), to_matrix_cl / from_matrix_cl
// (stan/math/opencl/copy.hpp:37-93), transpose (stan/math/opencl/transpose.hpp) and operator* / multiply
// (stan/math/opencl/multiply.hpp, kernels/matrix_multiply.hpp:18-130).
//
// Storage: column-major like Eigen and matrix_cl, leading dimension = rows rounded up to 16 doubles (128 B) so that
// every column starts on a 128-byte line and the kernels can use 16-byte loads; the pad rows are zero.
//
// Product: C = op(A) * op(B) in FP64 on the tensor cores' FP64 path (mma.sync m8n8k4 f64, SASS DMMA): a CTA of 256
// threads owns a BM x BN tile of C, walks K in steps of BK = 8 through two shared-memory stages, and each warp
// accumulates its part of the tile as 8 x 8 DMMA fragments in registers (dgemm_kernel below).  Either operand may be
// transposed: the shared-memory layout of its tile follows its storage order, no transposed copy is made.  The
// roofline is the FP64 rate (64 FMA / clk / SM), not HBM.
#include <algorithm>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "context.cuh"

struct stan_cuda_matrix {
  int device = -1;
  int64_t rows = 0, cols = 0, ld = 0;
  double* d = nullptr;
};

namespace stancuda {

int scope_device_ordinal();  // api.cu: device of the calling thread's shard scope, or the first device

namespace {

constexpr int kGemmThreads = 256;
constexpr int kBK = 8;

inline int64_t padded_ld(int64_t rows) { return rows <= 0 ? 16 : (rows + 15) / 16 * 16; }

// ---- C = alpha * op(A) * op(B) + beta * C, column-major ------------------------------------------------------
// D(8x8) += A(8x4) * B(4x8), FP64 (SASS DMMA.8x8x4 -- every f64 mma shape lowers to it on sm_100a).  Fragments (PTX
// ISA, mma.m8n8k4 .f64): a = A[lane / 4][lane % 4], b = B[lane % 4][lane / 4], {d0, d1} = D[lane / 4][2 * (lane % 4) + {0, 1}].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// Why DMMA although its peak equals the DFMA peak on this chip: operand delivery.  A register-tiled DFMA kernel
// with 8 x 8 accumulators per thread reads 128 B of shared memory per thread and K step for 64 FMA -- with four
// resident warps per SM that is exactly the 128 B / clk the shared-memory pipe moves, and the first version of this
// kernel sat at 57 % of the FP64 pipe with the LSU saturated (profiles/r02_ncu_dgemm_simt_4096.txt).  A warp that
// owns a 64 x 32 tile as 8 x 4 DMMA fragments reads 12 doubles per thread for 32 DMMAs (256 FMA per thread): a
// fifth of the traffic per flop.
//
// CTA: 8 warps as WARPS_M x (8 / WARPS_M); shared stages (two, BK = 8 deep), one layout per storage order of the
// operand so that no transposed copy is ever made:
//   "long" layout  T[k][x], row stride BX + 4 doubles -- for an operand whose M (resp. N) index is contiguous in
//                  memory (A not transposed, B transposed): 16-byte stores; the four k of a fragment land 8 banks
//                  apart -> conflict-free LDS.64
//   "deep" layout  T[x][k], row stride 12 doubles (8 used) -- for an operand whose K index is contiguous (A
//                  transposed, B not transposed): 16-byte stores; the rows of a fragment land 24 banks apart ->
//                  conflict-free LDS.64
// The global loads of stage t + 1 are issued before the DMMAs of stage t and stored after them: one barrier per
// stage.  Every element of C is one chain of fused multiply-adds over k ascending, whichever tile owns it.
// LOWER_ONLY skips the tiles strictly above the diagonal (the symmetric rank-k updates of the Cholesky routines).
template <int BX, bool DEEP>
struct OperandTile {
  static constexpr int kDoubles = DEEP ? BX * 12 : kBK * (BX + 4);
  static constexpr int kPerThread = BX * kBK / kGemmThreads;  // 2 or 4 contiguous doubles
  static_assert(kPerThread == 2 || kPerThread == 4, "one or two 16-byte loads per thread");
  // which (x, k) a thread loads: `kPerThread` consecutive elements along the contiguous index
  __device__ static int x_of(int tid) { return DEEP ? tid / (kBK / kPerThread) : (tid % (BX / kPerThread)) * kPerThread; }
  __device__ static int k_of(int tid) { return DEEP ? (tid % (kBK / kPerThread)) * kPerThread : tid / (BX / kPerThread); }
  __device__ static int offset(int x, int k) { return DEEP ? x * 12 + k : k * (BX + 4) + x; }
  // Per-thread loader of this operand's share of one stage: element (x, k) of op(P) is P[x + k * ld] when x is
  // contiguous and P[k + x * ld] when k is.  No test along x is ever needed: an element with x >= X only feeds rows
  // (columns) of C that the epilogue does not store, so such a thread just reads a valid address (x clamped to 0; a
  // group of 2 or 4 that straddles X stays inside the leading dimension, which is a multiple of 4).  Along k only
  // the last, partial stage tests, element by element.  The pointer advances by a constant per stage.
  struct Loader {
    const double* p;   // this thread's first element at k0 = 0
    int64_t step;      // pointer increment per stage
    int kt;            // this thread's k offset inside a stage
    __device__ Loader(const double* P, int64_t ld, int64_t x0, int64_t X, int tid) {
      int64_t x = x0 + x_of(tid);
      if (x >= X) x = 0;
      kt = k_of(tid);
      if (DEEP) {
        p = P + x * ld + kt;
        step = kBK;
      } else {
        p = P + int64_t(kt) * ld + x;
        step = int64_t(kBK) * ld;
      }
    }
    // k_left = K - k0 of the stage being loaded (> 0)
    __device__ void load(int64_t stage, int64_t k_left, double* r) const {
      const double* src = p + stage * step;
      if (k_left >= kBK) {
#pragma unroll
        for (int q = 0; q < kPerThread; q += 2) {
          const double2 v = __ldg(reinterpret_cast<const double2*>(src + q));
          r[q] = v.x;
          r[q + 1] = v.y;
        }
      } else if (DEEP) {
#pragma unroll
        for (int q = 0; q < kPerThread; ++q) r[q] = kt + q < k_left ? __ldg(src + q) : 0.0;
      } else {
#pragma unroll
        for (int q = 0; q < kPerThread; ++q) r[q] = kt < k_left ? __ldg(src + q) : 0.0;
      }
    }
  };
  __device__ static void store(double* T, int tid, const double* r) {
    double* t = T + offset(x_of(tid), k_of(tid));
#pragma unroll
    for (int q = 0; q < kPerThread; q += 2) *reinterpret_cast<double2*>(t + q) = make_double2(r[q], r[q + 1]);
  }
};

template <int BM, int BN, int WARPS_M, int MIN_CTAS, bool TA, bool TB, bool LOWER_ONLY>
__global__ void __launch_bounds__(kGemmThreads, MIN_CTAS)
dgemm_kernel(int64_t M, int64_t N, int64_t K, double alpha, const double* __restrict__ A, int64_t lda, const double* __restrict__ B,
             int64_t ldb, double beta, double* __restrict__ C, int64_t ldc, int64_t k_chunk, int64_t c_split_stride) {
  // split-K (gridDim.z > 1): slice z covers k in [z * k_chunk, min(K, (z + 1) * k_chunk)) and writes its partial
  // product (alpha = 1, beta = 0 from the launcher) into its own copy of C, c_split_stride doubles further on
  if (gridDim.z > 1) {
    const int64_t k_begin = int64_t(blockIdx.z) * k_chunk;
    A += TA ? k_begin : k_begin * lda;
    B += TB ? k_begin * ldb : k_begin;
    C += int64_t(blockIdx.z) * c_split_stride;
    K = min(K - k_begin, k_chunk);
  }
  constexpr int WARPS_N = kGemmThreads / 32 / WARPS_M;
  constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;  // warp tile
  constexpr int TM = WM / 8, TN = WN / 8;              // DMMA fragments per warp tile
  typedef OperandTile<BM, TA> TileA;                   // A transposed: K contiguous -> deep
  typedef OperandTile<BN, !TB> TileB;                  // B not transposed: K contiguous -> deep
  const int64_t m0 = int64_t(blockIdx.x) * BM, n0 = int64_t(blockIdx.y) * BN;
  if (LOWER_ONLY && n0 >= m0 + BM) return;             // every element of the tile has n > m
  __shared__ __align__(16) double As[2][TileA::kDoubles];
  __shared__ __align__(16) double Bs[2][TileB::kDoubles];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm0 = (warp % WARPS_M) * WM, wn0 = (warp / WARPS_M) * WN;
  const int fr = lane >> 2, fc = lane & 3;

  double acc[TM][TN][2];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  double ra[TileA::kPerThread], rb[TileB::kPerThread];
  const int64_t steps = (K + kBK - 1) / kBK;
  const typename TileA::Loader load_a(A, lda, m0, M, tid);
  const typename TileB::Loader load_b(B, ldb, n0, N, tid);
  load_a.load(0, K, ra);
  load_b.load(0, K, rb);
  TileA::store(As[0], tid, ra);
  TileB::store(Bs[0], tid, rb);
  __syncthreads();
  for (int64_t t = 0; t < steps; ++t) {
    const int buf = int(t & 1);
    if (t + 1 < steps) {
      load_a.load(t + 1, K - (t + 1) * kBK, ra);
      load_b.load(t + 1, K - (t + 1) * kBK, rb);
    }
#pragma unroll
    for (int ks = 0; ks < kBK; ks += 4) {
      double a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[buf][TileA::offset(wm0 + i * 8 + fr, ks + fc)];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[buf][TileB::offset(wn0 + j * 8 + fr, ks + fc)];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    if (t + 1 < steps) {
      TileA::store(As[buf ^ 1], tid, ra);
      TileB::store(Bs[buf ^ 1], tid, rb);
    }
    __syncthreads();
  }

  // epilogue: fragment (i, j) holds C[m0 + wm0 + i * 8 + fr][n0 + wn0 + j * 8 + 2 * fc + {0, 1}]
  const bool plain = alpha == 1.0 && beta == 0.0;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int64_t m = m0 + wm0 + i * 8 + fr;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int64_t n = n0 + wn0 + j * 8 + 2 * fc + e;
        if (n >= N || (LOWER_ONLY && n > m)) continue;
        double* c = C + n * ldc + m;
        if (plain) *c = acc[i][j][e];
        else *c = beta == 0.0 ? alpha * acc[i][j][e] : fma(alpha, acc[i][j][e], beta * *c);
      }
    }
  }
}

template <int BM, int BN, int WARPS_M, int MIN_CTAS, bool TA, bool TB, bool LOWER_ONLY>
int launch_dgemm(int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda, const double* B, int64_t ldb,
                 double beta, double* C, int64_t ldc, cudaStream_t stream, int splits = 1, int64_t k_chunk = 0,
                 int64_t c_split_stride = 0) {
  const int64_t gx = (M + BM - 1) / BM, gy = (N + BN - 1) / BN;
  if (gy > 65535) return fail(STAN_CUDA_ERR_ARG, "dgemm: more than 65535 column tiles");
  dgemm_kernel<BM, BN, WARPS_M, MIN_CTAS, TA, TB, LOWER_ONLY>
      <<<dim3(unsigned(gx), unsigned(gy), unsigned(splits)), kGemmThreads, 0, stream>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc,
                                                                                        k_chunk, c_split_stride);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

// C = alpha * (sum of the split-K partials, in slice order) + beta * C
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const double* __restrict__ part, int splits, int64_t stride, int64_t M,
                                                            int64_t N, int64_t ldp, double alpha, double beta, double* __restrict__ C,
                                                            int64_t ldc) {
  const int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (idx >= M * N) return;
  const int64_t m = idx % M, n = idx / M;
  double s = 0.0;
  for (int z = 0; z < splits; ++z) s += part[z * stride + n * ldp + m];
  double* c = C + n * ldc + m;
  *c = beta == 0.0 ? alpha * s : fma(alpha, s, beta * *c);
}

// Tile variants.  0: 128 x 128, warps 2 x 4 (64 x 32 each), one CTA per SM.  1: 128 x 64, warps 4 x 2 (32 x 32 each),
// two CTAs per SM: while one CTA waits at its stage barrier the other keeps the DMMA pipe busy.  2: 64 x 64 for
// products with fewer tiles than SMs.  STAN_CUDA_GEMM_TILE=0|1|2 forces one (for measurements); every variant gives
// the same bits (same FMA chain per element).
int gemm_tile_override() {
  static const int v = [] {
    const char* e = getenv("STAN_CUDA_GEMM_TILE");
    return e ? atoi(e) : -1;
  }();
  return v;
}

template <bool TA, bool TB, bool LOWER_ONLY>
int dgemm_variant(int device, int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda, const double* B,
                  int64_t ldb, double beta, double* C, int64_t ldc, cudaStream_t stream) {
  int variant = gemm_tile_override();
  const int sms = sm_count(device);
  if (variant < 0 || variant > 2) {
    const int64_t tiles = ((M + 127) / 128) * ((N + 63) / 64);
    variant = tiles >= 2 * int64_t(sms) * (LOWER_ONLY ? 2 : 1) ? 1 : 2;
  }
  if (variant == 0) return launch_dgemm<128, 128, 2, 1, TA, TB, LOWER_ONLY>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, stream);
  if (variant == 1) return launch_dgemm<128, 64, 4, 2, TA, TB, LOWER_ONLY>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, stream);
  // Few tiles and a long K (the A' B products of the Cholesky reverse sweep: a 64 x j result summed over thousands
  // of rows): split K over blockIdx.z so that the machine is filled, partial products into a workspace, then a
  // reduction in slice order -- the result does not depend on which SM ran which slice.
  const int64_t tiles64 = ((M + 63) / 64) * ((N + 63) / 64);
  if (!LOWER_ONLY && K >= 1024 && tiles64 * 2 <= sms) {
    int splits = int(std::min<int64_t>(std::min<int64_t>(32, K / 256), (2 * int64_t(sms) + tiles64 - 1) / tiles64));
    if (splits > 1) {
      const int64_t k_chunk = ((K + splits - 1) / splits + kBK - 1) / kBK * kBK;
      splits = int((K + k_chunk - 1) / k_chunk);
      const int64_t ldp = (M + 3) / 4 * 4, stride = ldp * N;
      double* part = nullptr;
      SC_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&part), size_t(stride) * splits * sizeof(double), stream));
      int rc = launch_dgemm<64, 64, 2, 2, TA, TB, false>(M, N, K, 1.0, A, lda, B, ldb, 0.0, part, ldp, stream, splits, k_chunk, stride);
      if (rc >= 0) {
        splitk_reduce_kernel<<<unsigned((M * N + 255) / 256), 256, 0, stream>>>(part, splits, stride, M, N, ldp, alpha, beta, C, ldc);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = fail_cuda(e, "splitk_reduce_kernel", __FILE__, __LINE__);
        counters().launches.fetch_add(1, std::memory_order_relaxed);
      }
      cudaFreeAsync(part, stream);
      return rc;
    }
  }
  return launch_dgemm<64, 64, 2, 2, TA, TB, LOWER_ONLY>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}

}  // namespace

// C (M x N) = alpha * op(A) * op(B) + beta * C on raw device pointers (sub-blocks of larger matrices included): every
// operand pointer 32-byte aligned, every operand leading dimension a multiple of 4 (the loads of a tile edge may read
// up to three elements past the last row of an operand, inside its leading dimension).  With lower_only only the elements with n <= m are
// computed and written (M == N for the callers: symmetric updates).
int dgemm(int device, bool trans_a, bool trans_b, bool lower_only, int64_t M, int64_t N, int64_t K, double alpha, const double* A,
          int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc, cudaStream_t stream) {
  if (M <= 0 || N <= 0) return STAN_CUDA_OK;
  if (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 31) || (lda & 3) || (ldb & 3))
    return fail(STAN_CUDA_ERR_ARG, "dgemm: operands must be 32-byte aligned with leading dimensions that are multiples of 4");
#define STAN_CUDA_DGEMM_CASE(TA, TB)                                                                                      \
  if (trans_a == TA && trans_b == TB)                                                                                     \
    return lower_only ? dgemm_variant<TA, TB, true>(device, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, stream)          \
                      : dgemm_variant<TA, TB, false>(device, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, stream);
  STAN_CUDA_DGEMM_CASE(false, false)
  STAN_CUDA_DGEMM_CASE(false, true)
  STAN_CUDA_DGEMM_CASE(true, false)
  STAN_CUDA_DGEMM_CASE(true, true)
#undef STAN_CUDA_DGEMM_CASE
  return STAN_CUDA_OK;
}

namespace {

// ---- host <-> device copies of a rows x cols column-major block ---------------------------------------------
// Copies from / to PAGEABLE host memory crawl on this platform: 0.75 GB/s measured for a 134 MB matrix, 1-D or 2-D
// alike (profiles/r02_gemm_timing_*.jsonl), against > 20 GB/s from page-locked memory.  Eigen's and the autodiff
// arena's memory is pageable, so every block larger than a few pages goes through two page-locked staging buffers
// of the calling thread: the CPU packs (unpacks) one chunk while the DMA engine moves the other.
// STAN_CUDA_DENSE_COPY=direct restores the plain cudaMemcpy path (for measuring the difference).
// The CPU side of a staged copy (pack into / unpack from page-locked memory) runs at the ~12 GB/s of one core, a
// quarter of what the DMA engine moves; a few helper threads bring the pair towards the PCIe rate.  A tiny persistent
// pool: the caller posts the slices of one chunk, works on the first slice itself and waits for the others.
// STAN_CUDA_COPY_THREADS (default 8 on hosts with 16 or more hardware threads, else 4; 1 = no helpers).
class CopyPool {
 public:
  static CopyPool& instance() {
    static CopyPool* pool = new CopyPool();  // leaked on purpose: helper threads must not be joined at static destruction
    return *pool;
  }
  int parts() const { return helpers_ + 1; }
  // fn(begin, end) over [0, n) in parts() contiguous slices; the caller runs the first slice and waits for the rest
  template <typename F>
  void for_slices(size_t n, const F& fn) {
    const size_t p = size_t(parts());
    if (p == 1 || n < p) {
      fn(size_t(0), n);
      return;
    }
    const size_t slice = (n + p - 1) / p;
    std::function<void(size_t, size_t)> f = fn;
    std::unique_lock<std::mutex> call(call_mu_);  // one parallel copy at a time (callers on other threads queue here)
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &f;
      tasks_.clear();
      for (size_t q = 1; q < p; ++q) {
        const size_t b = std::min(n, slice * q), e = std::min(n, slice * (q + 1));
        if (e > b) tasks_.push_back(std::make_pair(b, e));
      }
      pending_ = int(tasks_.size());
      next_ = 0;
    }
    cv_.notify_all();
    fn(size_t(0), std::min(n, slice));
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }

 private:
  CopyPool() {
    const char* e = getenv("STAN_CUDA_COPY_THREADS");
    const unsigned hw = std::thread::hardware_concurrency();
    int n = e ? atoi(e) : (hw >= 16 ? 8 : 4);  // measured: 134 MB up at 12.9 / 24.6 / 32.6 GB/s with 1 / 4 / 8 threads
    if (hw > 0 && n > int(hw)) n = int(hw);
    helpers_ = n > 1 ? n - 1 : 0;
    for (int q = 0; q < helpers_; ++q) std::thread([this] { work(); }).detach();
  }
  void work() {
    for (;;) {
      std::pair<size_t, size_t> t;
      const std::function<void(size_t, size_t)>* f;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return next_ < tasks_.size(); });
        t = tasks_[next_++];
        f = fn_;
      }
      (*f)(t.first, t.second);
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (--pending_ == 0) done_cv_.notify_all();
      }
    }
  }
  int helpers_ = 0;
  std::mutex call_mu_, mu_;
  std::condition_variable cv_, done_cv_;
  std::vector<std::pair<size_t, size_t>> tasks_;
  const std::function<void(size_t, size_t)>* fn_ = nullptr;
  size_t next_ = 0;
  int pending_ = 0;
};

struct CopyStage {
  double* buf[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};
  size_t cap = 0;  // doubles per buffer
  ~CopyStage() {
    for (int q = 0; q < 2; ++q) {
      if (buf[q]) cudaFreeHost(buf[q]);
      if (done[q]) cudaEventDestroy(done[q]);
    }
  }
};
constexpr size_t kStageDoubles = size_t(1) << 20;  // 8 MiB per buffer
constexpr int64_t kDirectCopyDoubles = 8192;       // blocks up to 64 KiB: one plain copy

int copy_stage(CopyStage** out) {
  static thread_local CopyStage st;
  if (!st.buf[0]) {
    for (int q = 0; q < 2; ++q) {
      SC_CUDA(cudaMallocHost(reinterpret_cast<void**>(&st.buf[q]), kStageDoubles * sizeof(double)));
      SC_CUDA(cudaEventCreateWithFlags(&st.done[q], cudaEventDisableTiming));
    }
    st.cap = kStageDoubles;
  }
  *out = &st;
  return STAN_CUDA_OK;
}

bool direct_copies() {
  static const bool direct = [] {
    const char* e = getenv("STAN_CUDA_DENSE_COPY");
    return e && std::string(e) == "direct";
  }();
  return direct;
}

// One chunk = `nc` pieces of `len` doubles, `host_stride` apart on the host and `dev_stride` apart on the device.
// Walks the block in chunks that fit a staging buffer; `to_device` selects the direction.  Device -> host returns
// with the data in place (the stream is synchronised).  max_chunk_cols > 0 caps the columns of a chunk; `gate`, when
// given, is called in front of the DMA of every chunk with the index one past the chunk's last column (a caller whose
// columns become final one group after the other makes the stream wait for the group's event there).
// `tri` (square blocks only): 1 = only the lower triangle travels, 2 = only the upper -- per chunk of columns the rows
// from the chunk's first diagonal element down (lower) or down to its last (upper); towards the host the rest of the
// columns is filled with zeros, on the device it is left as it is.
int staged_copy(bool to_device, double* dev, int64_t ld_dev, double* host, int64_t ld_host, int64_t rows, int64_t cols,
                cudaStream_t stream, int64_t max_chunk_cols = 0, const std::function<int(int64_t)>* gate = nullptr,
                int tri = 0) {
  CopyStage* st;
  SC_TRY(copy_stage(&st));
  const int64_t cap = int64_t(st->cap);
  const int64_t piece = std::min(rows, cap);                         // doubles of one column handled at a time
  const int64_t pieces_per_col = (rows + piece - 1) / piece;
  int64_t cols_per_chunk = pieces_per_col == 1 ? std::max<int64_t>(1, cap / rows) : 1;
  if (max_chunk_cols > 0 && cols_per_chunk > max_chunk_cols) cols_per_chunk = max_chunk_cols;
  static const bool tri_enabled = [] {
    const char* e = getenv("STAN_CUDA_TRI_COPY");  // 0: triangular operands travel whole (A/B)
    return !e || atoi(e) != 0;
  }();
  if (pieces_per_col != 1 || rows != cols || !tri_enabled) tri = 0;
  struct Chunk { int64_t j0, nc, r0, len; int q; };
  auto dma = [&](const Chunk& c) -> int {
    if (gate) SC_TRY((*gate)(c.j0 + c.nc));
    double* d = dev + c.j0 * ld_dev + c.r0;
    const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    if (to_device) {
      SC_CUDA(cudaMemcpy2DAsync(d, size_t(ld_dev) * sizeof(double), st->buf[c.q], size_t(c.len) * sizeof(double),
                                size_t(c.len) * sizeof(double), size_t(c.nc), kind, stream));
    } else {
      SC_CUDA(cudaMemcpy2DAsync(st->buf[c.q], size_t(c.len) * sizeof(double), d, size_t(ld_dev) * sizeof(double),
                                size_t(c.len) * sizeof(double), size_t(c.nc), kind, stream));
    }
    SC_CUDA(cudaEventRecord(st->done[c.q], stream));
    return STAN_CUDA_OK;
  };
  auto cpu_side = [&](const Chunk& c) -> int {  // pack before the DMA (to device) or unpack after it (to host)
    SC_CUDA(cudaEventSynchronize(st->done[c.q]));
    double* stage = st->buf[c.q];
    // host bytes this chunk touches (a triangle on its way to the host also fills the rest of its columns with zeros)
    const size_t touched = size_t(c.nc) * size_t(tri && !to_device ? rows : c.len) * sizeof(double);
    const bool big = touched >= (size_t(1) << 20);
    if (c.nc == 1 && !tri) {  // one long piece: byte ranges (4 KiB granules) over the copy threads
      char* h = reinterpret_cast<char*>(host + c.j0 * ld_host + c.r0);
      char* b = reinterpret_cast<char*>(stage);
      const size_t bytes = size_t(c.len) * sizeof(double), granules = (bytes + 4095) / 4096;
      auto piece = [&](size_t g0, size_t g1) {
        const size_t o = g0 * 4096, e = std::min(bytes, g1 * 4096);
        if (to_device) memcpy(b + o, h + o, e - o);
        else memcpy(h + o, b + o, e - o);
      };
      if (big) CopyPool::instance().for_slices(granules, piece);
      else piece(0, granules);
    } else {          // whole columns: column ranges over the copy threads
      auto cols_fn = [&](size_t j0, size_t j1) {
        for (size_t j = j0; j < j1; ++j) {
          double* col = host + (c.j0 + int64_t(j)) * ld_host;
          double* h = col + c.r0;
          double* b = stage + int64_t(j) * c.len;
          if (to_device) {
            memcpy(b, h, size_t(c.len) * sizeof(double));
          } else {
            memcpy(h, b, size_t(c.len) * sizeof(double));
            if (tri) {  // the part of the column that did not travel
              if (c.r0 > 0) memset(col, 0, size_t(c.r0) * sizeof(double));
              if (c.r0 + c.len < rows) memset(h + c.len, 0, size_t(rows - c.r0 - c.len) * sizeof(double));
            }
          }
        }
      };
      if (big) CopyPool::instance().for_slices(size_t(c.nc), cols_fn);
      else cols_fn(0, size_t(c.nc));
    }
    return STAN_CUDA_OK;
  };
  bool have_pending = false;
  Chunk pending{};
  int q = 0;
  for (int64_t j0 = 0; j0 < cols; j0 += cols_per_chunk)
    for (int64_t r0 = 0; r0 < rows; r0 += piece) {
      Chunk c{j0, std::min(cols_per_chunk, cols - j0), r0, std::min(piece, rows - r0), q};
      if (tri == 1) {         // rows from the first column's diagonal element (even start: 16-byte aligned pieces)
        c.r0 = j0 & ~int64_t(1);
        c.len = rows - c.r0;
      } else if (tri == 2) {  // rows down to the last column's diagonal element
        c.len = std::min(rows, j0 + c.nc);
      }
      q ^= 1;
      if (to_device) {
        SC_TRY(cpu_side(c));  // waits for the DMA that last read this buffer, then packs
        SC_TRY(dma(c));
      } else {
        SC_TRY(dma(c));       // buffer c.q was unpacked one iteration ago
        if (have_pending) SC_TRY(cpu_side(pending));
        pending = c;
        have_pending = true;
      }
    }
  if (!to_device) {
    if (have_pending) SC_TRY(cpu_side(pending));
    SC_CUDA(cudaStreamSynchronize(stream));
  }
  return STAN_CUDA_OK;
}

int copy_to_device(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t rows, int64_t cols, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return STAN_CUDA_OK;
  const bool contiguous = cols == 1 || (ld_dev == rows && ld_host == rows);
  if (rows * cols <= kDirectCopyDoubles || direct_copies()) {
    if (contiguous) {
      SC_CUDA(cudaMemcpyAsync(dev, host, size_t(rows) * size_t(cols) * sizeof(double), cudaMemcpyHostToDevice, stream));
    } else {
      SC_CUDA(cudaMemcpy2DAsync(dev, size_t(ld_dev) * sizeof(double), host, size_t(ld_host) * sizeof(double),
                                size_t(rows) * sizeof(double), size_t(cols), cudaMemcpyHostToDevice, stream));
    }
    return STAN_CUDA_OK;
  }
  if (contiguous) return staged_copy(true, dev, rows * cols, const_cast<double*>(host), rows * cols, rows * cols, 1, stream);
  return staged_copy(true, dev, ld_dev, const_cast<double*>(host), ld_host, rows, cols, stream);
}

// returns with the data in `host` (synchronises the stream)
int copy_to_host(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) {
    SC_CUDA(cudaStreamSynchronize(stream));
    return STAN_CUDA_OK;
  }
  const bool contiguous = cols == 1 || (ld_dev == rows && ld_host == rows);
  if (rows * cols <= kDirectCopyDoubles || direct_copies()) {
    if (contiguous) {
      SC_CUDA(cudaMemcpyAsync(host, dev, size_t(rows) * size_t(cols) * sizeof(double), cudaMemcpyDeviceToHost, stream));
    } else {
      SC_CUDA(cudaMemcpy2DAsync(host, size_t(ld_host) * sizeof(double), dev, size_t(ld_dev) * sizeof(double),
                                size_t(rows) * sizeof(double), size_t(cols), cudaMemcpyDeviceToHost, stream));
    }
    SC_CUDA(cudaStreamSynchronize(stream));
    return STAN_CUDA_OK;
  }
  double* d = const_cast<double*>(dev);
  if (contiguous) return staged_copy(false, d, rows * cols, host, rows * cols, rows * cols, 1, stream);
  return staged_copy(false, d, ld_dev, host, ld_host, rows, cols, stream);
}

// device -> host in column chunks, each behind its gate (see staged_copy); returns with the data in `host`
int copy_to_host_gated(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols,
                       cudaStream_t stream, int64_t chunk_cols, const std::function<int(int64_t)>& gate, int tri) {
  if (rows <= 0 || cols <= 0 || direct_copies() || rows * cols <= kDirectCopyDoubles) {
    SC_TRY(gate(cols));
    return copy_to_host(host, ld_host, dev, ld_dev, rows, cols, stream);
  }
  return staged_copy(false, const_cast<double*>(dev), ld_dev, host, ld_host, rows, cols, stream, chunk_cols, &gate, tri);
}

// Square n x n blocks of which one triangle matters (tri = 1: lower, 2: upper): only that triangle crosses the bus.
// Towards the device the other triangle of `dev` keeps its contents (the callers clear the buffer first); towards the
// host it is filled with zeros.  Small blocks travel whole (the device copy of the other triangle is then the caller's).
int copy_to_device_tri(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t n, cudaStream_t stream, int tri) {
  if (n <= 0) return STAN_CUDA_OK;
  if (n * n <= kDirectCopyDoubles || direct_copies()) return copy_to_device(dev, ld_dev, host, ld_host, n, n, stream);
  return staged_copy(true, dev, ld_dev, const_cast<double*>(host), ld_host, n, n, stream, 0, nullptr, tri);
}
int copy_to_host_tri(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t n, cudaStream_t stream, int tri) {
  if (n <= 0 || n * n <= kDirectCopyDoubles || direct_copies()) return copy_to_host(host, ld_host, dev, ld_dev, n, n, stream);
  return staged_copy(false, const_cast<double*>(dev), ld_dev, host, ld_host, n, n, stream, 0, nullptr, tri);
}

int use_device_of(const stan_cuda_matrix* m, ThreadDev** td) {
  SC_TRY(thread_dev(m->device, td));
  SC_CUDA(cudaSetDevice(m->device));
  return STAN_CUDA_OK;
}

// the pool keeps freed blocks: the matrices of one log-density evaluation are the matrices of the next
int tune_pool_once(int device) {
  static std::mutex mu;
  static std::vector<int> done;
  std::lock_guard<std::mutex> lk(mu);
  if (std::find(done.begin(), done.end(), device) != done.end()) return STAN_CUDA_OK;
  cudaMemPool_t pool;
  SC_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
  unsigned long long keep = 16ull << 30;
  SC_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  done.push_back(device);
  return STAN_CUDA_OK;
}

int matrix_create_on(int device, int64_t rows, int64_t cols, bool zero, stan_cuda_matrix** out) {
  if (rows < 0 || cols < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_create: negative size");
  ThreadDev* td;
  SC_TRY(thread_dev(device, &td));
  SC_CUDA(cudaSetDevice(device));
  SC_TRY(tune_pool_once(device));
  stan_cuda_matrix* m = new stan_cuda_matrix();
  m->device = device;
  m->rows = rows;
  m->cols = cols;
  m->ld = padded_ld(rows);
  const size_t bytes = size_t(m->ld) * size_t(std::max<int64_t>(cols, 1)) * sizeof(double);
  cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&m->d), bytes, td->stream);
  if (e == cudaSuccess && zero) e = cudaMemsetAsync(m->d, 0, bytes, td->stream);
  if (e != cudaSuccess) {
    delete m;
    return fail_cuda(e, "cudaMallocAsync / cudaMemsetAsync", __FILE__, __LINE__);
  }
  *out = m;
  return STAN_CUDA_OK;
}

}  // namespace

int dense_copy_to_device(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t rows, int64_t cols, cudaStream_t stream) {
  return copy_to_device(dev, ld_dev, host, ld_host, rows, cols, stream);
}
int dense_copy_to_host(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols, cudaStream_t stream) {
  return copy_to_host(host, ld_host, dev, ld_dev, rows, cols, stream);
}
int dense_copy_to_host_gated(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols,
                             cudaStream_t stream, int64_t chunk_cols, const std::function<int(int64_t)>& gate, int tri) {
  return copy_to_host_gated(host, ld_host, dev, ld_dev, rows, cols, stream, chunk_cols, gate, tri);
}
int dense_copy_to_device_tri(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t n, cudaStream_t stream, int tri) {
  return copy_to_device_tri(dev, ld_dev, host, ld_host, n, stream, tri);
}
int dense_copy_to_host_tri(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t n, cudaStream_t stream, int tri) {
  return copy_to_host_tri(host, ld_host, dev, ld_dev, n, stream, tri);
}

int matrix_create(int64_t rows, int64_t cols, stan_cuda_matrix** out) {
  if (!out) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_create: null out");
  SC_TRY(ensure_init());
  stan_cuda_matrix* m = nullptr;
  SC_TRY(matrix_create_on(scope_device_ordinal(), rows, cols, true, &m));
  ThreadDev* td;
  SC_TRY(thread_dev(m->device, &td));
  SC_CUDA(cudaStreamSynchronize(td->stream));  // handles may travel between host threads: nothing is left in flight
  *out = m;
  return STAN_CUDA_OK;
}

int matrix_free(stan_cuda_matrix* m) {
  if (!m) return STAN_CUDA_OK;
  ThreadDev* td;
  int rc = thread_dev(m->device, &td);
  if (rc >= 0) {
    cudaSetDevice(m->device);
    cudaFreeAsync(m->d, td->stream);
  }
  delete m;
  return rc < 0 ? rc : STAN_CUDA_OK;
}

int matrix_upload(stan_cuda_matrix* m, const double* host, int64_t ld) {
  if (!m || (!host && m->rows * m->cols > 0)) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_upload: null argument");
  if (ld < m->rows) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_upload: leading dimension below the row count");
  if (m->rows == 0 || m->cols == 0) return STAN_CUDA_OK;
  NvtxRange range("stan_cuda_matrix_upload");
  ThreadDev* td;
  SC_TRY(use_device_of(m, &td));
  SC_TRY(copy_to_device(m->d, m->ld, host, ld, m->rows, m->cols, td->stream));
  SC_CUDA(cudaStreamSynchronize(td->stream));
  counters().bytes_uploaded.fetch_add(uint64_t(m->rows) * uint64_t(m->cols) * sizeof(double));
  return STAN_CUDA_OK;
}

int matrix_download(const stan_cuda_matrix* m, double* host, int64_t ld) {
  if (!m || (!host && m->rows * m->cols > 0)) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_download: null argument");
  if (ld < m->rows) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_download: leading dimension below the row count");
  if (m->rows == 0 || m->cols == 0) return STAN_CUDA_OK;
  NvtxRange range("stan_cuda_matrix_download");
  ThreadDev* td;
  SC_TRY(use_device_of(m, &td));
  SC_TRY(copy_to_host(host, ld, m->d, m->ld, m->rows, m->cols, td->stream));
  return STAN_CUDA_OK;
}

// c = op(a) * op(b); c must have the shape of the product.  Asynchronous on the calling thread's stream when
// `sync` is false (used by the host-level entry, which synchronises once at the end).
int matrix_multiply(const stan_cuda_matrix* a, int trans_a, const stan_cuda_matrix* b, int trans_b, stan_cuda_matrix* c, bool sync) {
  if (!a || !b || !c) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: null matrix");
  if (a->device != b->device || a->device != c->device)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: operands on different devices");
  const int64_t M = trans_a ? a->cols : a->rows, K = trans_a ? a->rows : a->cols;
  const int64_t Kb = trans_b ? b->cols : b->rows, N = trans_b ? b->rows : b->cols;
  if (K != Kb) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: inner dimensions differ");
  if (c->rows != M || c->cols != N) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: result has the wrong shape");
  if (c == a || c == b) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: the result aliases an operand");
  NvtxRange range("stan_cuda_matrix_multiply");
  ThreadDev* td;
  SC_TRY(use_device_of(c, &td));
  if (M == 0 || N == 0) return STAN_CUDA_OK;
  if (K == 0) {  // empty sum: the zero matrix (Eigen's product of an M x 0 and a 0 x N matrix)
    SC_CUDA(cudaMemsetAsync(c->d, 0, size_t(c->ld) * size_t(c->cols) * sizeof(double), td->stream));
    if (sync) SC_CUDA(cudaStreamSynchronize(td->stream));
    return STAN_CUDA_OK;
  }
  int rc = dgemm(c->device, trans_a != 0, trans_b != 0, false, M, N, K, 1.0, a->d, a->ld, b->d, b->ld, 0.0, c->d, c->ld, td->stream);
  if (rc >= 0 && sync) SC_CUDA(cudaStreamSynchronize(td->stream));
  return rc;
}

// host matrices in, host matrix out: upload, multiply, download -- one synchronisation
int multiply_host(const stan_cuda_multiply_args* g) {
  if (!g) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_multiply: null args");
  if (g->m < 0 || g->n < 0 || g->k < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_multiply: negative size");
  if (g->m == 0 || g->n == 0) return STAN_CUDA_OK;
  const int64_t a_rows = g->trans_a ? g->k : g->m, a_cols = g->trans_a ? g->m : g->k;
  const int64_t b_rows = g->trans_b ? g->n : g->k, b_cols = g->trans_b ? g->k : g->n;
  if (!g->c || g->ldc < g->m) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_multiply: bad result matrix");
  if (g->k > 0 && (!g->a || !g->b || g->lda < a_rows || g->ldb < b_rows))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_multiply: bad operand");
  NvtxRange range("stan_cuda_multiply");
  SC_TRY(ensure_init());
  const int dev = scope_device_ordinal();
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  stan_cuda_matrix *a = nullptr, *b = nullptr, *c = nullptr;
  int rc = matrix_create_on(dev, a_rows, a_cols, true, &a);
  if (rc >= 0) rc = matrix_create_on(dev, b_rows, b_cols, true, &b);
  if (rc >= 0) rc = matrix_create_on(dev, g->m, g->n, false, &c);
  auto copy_in = [&](stan_cuda_matrix* m, const double* host, int64_t ld) -> int {
    if (m->rows == 0 || m->cols == 0) return STAN_CUDA_OK;
    SC_TRY(copy_to_device(m->d, m->ld, host, ld, m->rows, m->cols, td->stream));
    counters().bytes_uploaded.fetch_add(uint64_t(m->rows) * uint64_t(m->cols) * sizeof(double));
    return STAN_CUDA_OK;
  };
  if (rc >= 0) rc = copy_in(a, g->a, g->lda);
  if (rc >= 0) rc = copy_in(b, g->b, g->ldb);
  if (rc >= 0) rc = matrix_multiply(a, g->trans_a, b, g->trans_b, c, false);
  if (rc >= 0) rc = copy_to_host(g->c, g->ldc, c->d, c->ld, g->m, g->n, td->stream);
  if (a) matrix_free(a);
  if (b) matrix_free(b);
  if (c) matrix_free(c);
  if (rc >= 0) counters().evals.fetch_add(1, std::memory_order_relaxed);
  return rc;
}

}  // namespace stancuda

extern "C" {
int stan_cuda_matrix_create(int64_t rows, int64_t cols, stan_cuda_matrix** out) { return stancuda::matrix_create(rows, cols, out); }
void stan_cuda_matrix_free(stan_cuda_matrix* m) { stancuda::matrix_free(m); }
int64_t stan_cuda_matrix_rows(const stan_cuda_matrix* m) { return m ? m->rows : 0; }
int64_t stan_cuda_matrix_cols(const stan_cuda_matrix* m) { return m ? m->cols : 0; }
int stan_cuda_matrix_upload(stan_cuda_matrix* m, const double* host, int64_t ld) { return stancuda::matrix_upload(m, host, ld); }
int stan_cuda_matrix_download(const stan_cuda_matrix* m, double* host, int64_t ld) { return stancuda::matrix_download(m, host, ld); }
int stan_cuda_matrix_multiply(const stan_cuda_matrix* a, int trans_a, const stan_cuda_matrix* b, int trans_b, stan_cuda_matrix* c) {
  return stancuda::matrix_multiply(a, trans_a, b, trans_b, c, true);
}
int stan_cuda_multiply(const stan_cuda_multiply_args* a) { return stancuda::multiply_host(a); }
}
