// cholesky.cu -- Cholesky factorisation and its reverse-mode adjoint on the device.
//
// Replaces, for the cholesky_decompose hooks of the reference (stan/math/prim/mat/fun/cholesky_decompose.hpp:34-62,
// stan/math/rev/mat/fun/cholesky_decompose.hpp:255-383, 396-440), what its OpenCL backend provides there:
// opencl::cholesky_decompose (stan/math/opencl/cholesky_decompose.hpp:36-107: blocked, diagonal block factorised by
// one work-group, panel by tri_inverse + multiply, trailing matrix by multiply_transpose) and the blocked reverse
// sweep of cholesky_opencl::chain() (rev/mat/fun/cholesky_decompose.hpp:329-382; same algebra as the CPU
// cholesky_block::chain(), :120-167 -- Murray 2016, "Differentiation of the Cholesky decomposition").
//
// Forward, lower triangular, blocks of NB = 64 columns; A is consumed, L is written to a second matrix:
//   1. potrf_inv_kernel   one CTA: L_D = chol(D) for the 64 x 64 diagonal block in shared memory, and W = L_D^-1
//   2. dgemm (N, T)       panel     L[j+64:, j:j+64] = A[j+64:, j:j+64] * W'
//   3. dgemm (N, T), lower-only     A[j+64:, j+64:] -= panel * panel'
// Reverse, blocks from the bottom up (the block boundaries are multiples of 64 from the top: any blocking is valid
// and this one keeps every sub-block 16-byte aligned):
//   W_b = D_b^-1 for all blocks (tri_inv_blocks_kernel, one launch);  Cb = Cb W;  Bb -= Cb R;  Db -= Cb' C;  Db = sym(D, W, Db) (symbolic_rev_kernel);
//   Rb -= Cb' B + Db_sym R;  diag(Db) *= 1/2.
// The products are the DMMA kernel of gemm.cu on sub-blocks; the three small kernels run one CTA each and are the
// serial spine of both sweeps (n / 64 steps).
#include <algorithm>
#include <functional>
#include <vector>

#include "context.cuh"

namespace stancuda {

int scope_device_ordinal();  // api.cu

namespace {

constexpr int kNB = 64;          // block size of both sweeps
constexpr int kLD = kNB + 1;     // row stride of the 64 x 64 blocks in shared memory (bank spread)
constexpr int kThreads = 256;

__device__ __forceinline__ double& at(double* m, int r, int c) { return m[r * kLD + c]; }

// W = inverse of the lower-triangular jb x jb matrix Lm (both in shared memory, W zero above the diagonal), in two
// levels.  (1) The four 16 x 16 diagonal blocks by forward substitution, all at once: column c of its block solves
// L x = e_c, four lanes of one warp share a column and split the dot product of a step (two shuffles, fixed order).
// (2) The blocks below the diagonal by substitution at block level, W[r][c] = -W[r][r] * sum_k L[r][k] W[k][c] for
// k = c .. r - 1, one thread per element of a 16 x 16 block.  15 dependent steps + 6 small block products instead of
// the 63 dependent steps of a plain column sweep, which took 44 us -- half of the serial spine of both sweeps.
__device__ void tri_inverse_shared(const double* Lm, double* W, int jb, int tid) {
  __shared__ double Tm[16][17];
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) at(W, idx / kNB, idx % kNB) = 0.0;
  __syncthreads();
  {
    const int c = tid >> 2, part = tid & 3;  // 64 columns x 4 lanes; a warp holds 8 columns of one diagonal block
    const int base = c & ~15;
    const bool live = c < jb;
    if (live && part == 0) at(W, c, c) = 1.0 / Lm[c * kLD + c];
    __syncwarp();
    for (int ii = 1; ii < 16; ++ii) {
      const int i = base + ii;
      const bool mine = live && i > c && i < jb;
      double s = 0.0;
      if (mine) {
        for (int k = c + part; k < i; k += 4) s = fma(Lm[i * kLD + k], at(W, k, c), s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (mine && part == 0) at(W, i, c) = -s / Lm[i * kLD + i];
      __syncwarp();
    }
  }
  __syncthreads();
  const int tr = tid & 15, tc = tid >> 4;
  for (int d = 1; d < 4; ++d) {
    for (int rb = d; rb < 4; ++rb) {
      const int cb = rb - d;
      double t = 0.0;
      for (int kb = cb; kb < rb; ++kb) {
#pragma unroll 4
        for (int q = 0; q < 16; ++q) t = fma(Lm[(rb * 16 + tr) * kLD + kb * 16 + q], at(W, kb * 16 + q, cb * 16 + tc), t);
      }
      Tm[tr][tc] = t;
      __syncthreads();
      double w = 0.0;
#pragma unroll 4
      for (int q = 0; q < 16; ++q) w = fma(at(W, rb * 16 + tr, rb * 16 + q), Tm[q][tc], w);
      at(W, rb * 16 + tr, cb * 16 + tc) = -w;
      __syncthreads();
    }
  }
}

// load the lower triangle of a jb x jb block (zero elsewhere, zero padding up to 64 x 64)
__device__ void load_lower(double* dst, const double* src, int64_t ld, int jb, int tid) {
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) {
    const int r = idx % kNB, c = idx / kNB;
    at(dst, r, c) = (r < jb && c <= r) ? src[int64_t(c) * ld + r] : 0.0;
  }
}

// step 1 of the forward sweep.  info: 0, or 1 + the index of the first pivot that is not positive (NaN included)
__global__ void __launch_bounds__(kThreads) potrf_inv_kernel(const double* __restrict__ A, int64_t lda, int jb, int64_t j0,
                                                            double* __restrict__ L, int64_t ldl, double* __restrict__ W,
                                                            int* __restrict__ info) {
  extern __shared__ double sm[];
  double* D = sm;
  double* Wm = sm + kNB * kLD;
  __shared__ int failed;
  const int tid = threadIdx.x;
  if (*info != 0) return;  // an earlier block failed: nothing below it is meaningful
  if (tid == 0) failed = 0;
  load_lower(D, A, lda, jb, tid);
  // Left-looking column Cholesky (the form of Eigen's unblocked LLT): column j is the original column minus the
  // products with the columns already done, then scaled by the pivot's square root.  Four lanes share a row and
  // split that dot product -- no read-modify-write chain through shared memory, two barriers per column.  (The first
  // version updated the whole trailing block after every column: 1.4 us per column, the longest kernel of the sweep.)
  __shared__ double pivot_s;
  const int r = tid >> 2, part = tid & 3;
  for (int j = 0; j < jb; ++j) {
    __syncthreads();  // column j - 1 is in place
    double sum = 0.0;
    if (r >= j && r < jb) {
      for (int k = part; k < j; k += 4) sum = fma(at(D, r, k), at(D, j, k), sum);
    }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    const double v = (r >= j && r < jb) ? at(D, r, j) - sum : 0.0;
    if (r == j && part == 0) pivot_s = v;
    __syncthreads();
    const double pivot = pivot_s;
    if (!(pivot > 0.0)) {  // uniform: every thread reads the same value
      if (tid == 0) {
        *info = int(j0) + j + 1;
        failed = 1;
      }
      break;
    }
    const double ljj = sqrt(pivot);
    if (part == 0 && r >= j && r < jb) at(D, r, j) = (r == j) ? ljj : v / ljj;
  }
  __syncthreads();
  if (failed) return;
  tri_inverse_shared(D, Wm, jb, tid);
  for (int idx = tid; idx < jb * jb; idx += kThreads) {
    const int r = idx % jb, c = idx / jb;
    if (r >= c) L[int64_t(c) * ldl + r] = at(D, r, c);
  }
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) W[idx] = at(Wm, idx % kNB, idx / kNB);  // column-major, ld = 64
}

// W_b = D_b^-1 for every 64 x 64 diagonal block D_b of a triangular n x n matrix, one CTA per block (the blocks are
// independent: one launch instead of one per step of a sweep).  lower: D_b is the lower triangle of the block.
// Otherwise the block is upper triangular, it is loaded transposed, and W_b = (U_b')^-1 = (U_b^-1)': the callers
// flip the transpose flag of the product that applies it.
__global__ void __launch_bounds__(kThreads) tri_inv_blocks_kernel(const double* __restrict__ A, int64_t ld, int64_t n, int lower,
                                                                 double* __restrict__ Wall) {
  extern __shared__ double sm[];
  double* D = sm;
  double* Wm = sm + kNB * kLD;
  const int tid = threadIdx.x;
  const int64_t j = int64_t(blockIdx.x) * kNB;
  const int jb = int(min(int64_t(kNB), n - j));
  const double* Dg = A + j + j * ld;
  if (lower) {
    load_lower(D, Dg, ld, jb, tid);
  } else {
    for (int idx = tid; idx < kNB * kNB; idx += kThreads) {
      const int r = idx / kNB, c = idx % kNB;  // element (r, c) of the transposed block = U(c, r), kept for r >= c
      at(D, r, c) = (r < jb && c <= r) ? Dg[int64_t(r) * ld + c] : 0.0;
    }
  }
  __syncthreads();
  tri_inverse_shared(D, Wm, jb, tid);
  double* W = Wall + int64_t(blockIdx.x) * kNB * kNB;
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) W[idx] = at(Wm, idx % kNB, idx / kNB);
}

// out = X' * Y or X * Y for 64 x 64 blocks in shared memory; thread (tx, ty) of a 16 x 16 grid owns a 4 x 4 block
template <bool TRANS_X>
__device__ void block_product(const double* X, const double* Y, double* out, int tid) {
  const int r0 = (tid % 16) * 4, c0 = (tid / 16) * 4;
  double acc[4][4] = {};
  for (int k = 0; k < kNB; ++k) {
    double x[4], y[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) x[i] = TRANS_X ? X[k * kLD + r0 + i] : X[(r0 + i) * kLD + k];
#pragma unroll
    for (int j = 0; j < 4; ++j) y[j] = Y[k * kLD + c0 + j];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fma(x[i], y[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) out[(r0 + i) * kLD + c0 + j] = acc[i][j];
}

// symbolic_rev (rev/mat/fun/cholesky_decompose.hpp:102-112 / :301-307): with P = tril(Db),
//   Phi = symmetrised lower triangle of D' P;   Db_sym = D^-T Phi D^-1 = W' Phi W.
// Writes Db_sym (full, column-major, ld = 64) for the update of R and, into the adjoint matrix, its lower triangle with
// the diagonal halved (:160-161).
__global__ void __launch_bounds__(kThreads) symbolic_rev_kernel(const double* __restrict__ Dg, int64_t ldd, const double* __restrict__ Wg,
                                                               double* __restrict__ Db, int64_t ldb, int jb, double* __restrict__ Dsym) {
  extern __shared__ double sm[];
  double* D = sm;
  double* W = sm + kNB * kLD;
  double* P = sm + 2 * kNB * kLD;
  double* T = sm + 3 * kNB * kLD;
  const int tid = threadIdx.x;
  load_lower(D, Dg, ldd, jb, tid);
  load_lower(P, Db, ldb, jb, tid);
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) at(W, idx % kNB, idx / kNB) = Wg[idx];
  __syncthreads();
  block_product<true>(D, P, T, tid);  // T = D' P
  __syncthreads();
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) {
    const int r = idx / kNB, c = idx % kNB;
    at(P, r, c) = r >= c ? at(T, r, c) : at(T, c, r);  // Phi
  }
  __syncthreads();
  block_product<false>(P, W, T, tid);  // T = Phi W
  __syncthreads();
  block_product<true>(W, T, P, tid);   // P = W' T
  __syncthreads();
  for (int idx = tid; idx < kNB * kNB; idx += kThreads) {
    const int r = idx % kNB, c = idx / kNB;
    Dsym[idx] = at(P, r, c);
    if (r < jb && c < jb && r >= c) Db[int64_t(c) * ldb + r] = (r == c ? 0.5 : 1.0) * at(P, r, c);
  }
}

// zero the strict upper triangle of an n x n matrix
__global__ void zero_strict_upper_kernel(double* __restrict__ L, int64_t ld, int64_t n) {
  const int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, c = blockIdx.y * int64_t(64);
  if (r >= n) return;
  for (int64_t cc = c; cc < min(n, c + 64); ++cc)
    if (r < cc) L[cc * ld + r] = 0.0;
}

// check_symmetric of the reference (stan/math/prim/mat/err/check_symmetric.hpp:39-46) on the uploaded matrix:
// flag[1] = 1 when some pair fails  |A(m, n) - A(n, m)| <= tol  (NaN fails, as there).  32 x 32 tiles, the mirrored
// tile read through shared memory so that both reads are coalesced; only tiles on or below the diagonal work.
__global__ void __launch_bounds__(256) symmetry_kernel(const double* __restrict__ A, int64_t ld, int64_t n, double tol,
                                                       int* __restrict__ flag) {
  if (blockIdx.y > blockIdx.x) return;
  __shared__ double mirror[32][33];
  const int64_t r0 = int64_t(blockIdx.x) * 32, c0 = int64_t(blockIdx.y) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {  // mirror[j][tx] = A(c0 + tx, r0 + j): the tile across the diagonal
    const int64_t r = c0 + tx, c = r0 + j;
    mirror[j][tx] = (r < n && c < n) ? A[c * ld + r] : 0.0;
  }
  __syncthreads();
  bool bad = false;
  for (int j = ty; j < 32; j += 8) {  // A(r0 + tx, c0 + j) against A(c0 + j, r0 + tx) = mirror[tx][j]
    const int64_t r = r0 + tx, c = c0 + j;
    if (r < n && c < n && r > c) bad = bad || !(fabs(A[c * ld + r] - mirror[tx][j]) <= tol);
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag + 1, 1);
}

struct DeviceBuf {
  double* p = nullptr;
  cudaStream_t stream = nullptr;
  ~DeviceBuf() {
    if (p) cudaFreeAsync(p, stream);
  }
  int alloc(size_t doubles, cudaStream_t s) {
    stream = s;
    SC_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&p), std::max<size_t>(doubles, 2) * sizeof(double), s));
    return STAN_CUDA_OK;
  }
};

int64_t padded(int64_t n) { return n <= 0 ? 16 : (n + 15) / 16 * 16; }

int set_smem_limits() {  // per device (the attribute belongs to the function on the current device): cheap, every call
  SC_CUDA(cudaFuncSetAttribute(symbolic_rev_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(4 * kNB * kLD * sizeof(double))));
  SC_CUDA(cudaFuncSetAttribute(potrf_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(2 * kNB * kLD * sizeof(double))));
  SC_CUDA(cudaFuncSetAttribute(tri_inv_blocks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(2 * kNB * kLD * sizeof(double))));
  return STAN_CUDA_OK;
}

}  // namespace

// host <-> device block copies through page-locked staging (gemm.cu)
int dense_copy_to_device(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t rows, int64_t cols, cudaStream_t stream);
int dense_copy_to_host(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols, cudaStream_t stream);
int dense_copy_to_host_gated(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols,
                             cudaStream_t stream, int64_t chunk_cols, const std::function<int(int64_t)>& gate, int tri);
// one triangle of a square block (tri = 1: lower, 2: upper); see gemm.cu
int dense_copy_to_device_tri(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t n, cudaStream_t stream, int tri);
int dense_copy_to_host_tri(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t n, cudaStream_t stream, int tri);

namespace {
// A second stream and one event per block column for the calling thread: the factor goes back to the host block column
// group by group WHILE the sweep works on the columns to the right (a block column of L is final as soon as its panel
// product is done).  Unpacking into the caller's fresh pages is the longest part of a host-to-host factorisation
// (3.8 of 7.7 ms at n = 2048); the sweep (2.1 ms) hides behind it.
struct DownloadOverlap {
  int device = -1;
  cudaStream_t copy = nullptr;
  std::vector<cudaEvent_t> done;  // done[b]: block column b of L is final
  int prepare(int dev, size_t nblk) {
    if (device != dev) {
      release();
      SC_CUDA(cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking));
      device = dev;
    }
    while (done.size() < nblk) {
      cudaEvent_t e;
      SC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      done.push_back(e);
    }
    return STAN_CUDA_OK;
  }
  void release() {
    for (cudaEvent_t e : done) cudaEventDestroy(e);
    done.clear();
    if (copy) cudaStreamDestroy(copy);
    copy = nullptr;
    device = -1;
  }
};
bool download_overlap_enabled() {
  static const bool on = [] {
    const char* e = getenv("STAN_CUDA_CHOLESKY_OVERLAP");  // 0: factor first, then one download (A/B)
    return !e || atoi(e) != 0;
  }();
  return on;
}
}  // namespace

// A (n x n, its lower triangle is read) is consumed; L receives the factor (zeros above the diagonal).  *info_host: 0
// or 1 + index of the first non-positive pivot.
// column_done, when given: one event per block column, recorded when that block column of L is final.
int cholesky_device(int dev, double* A, double* L, int64_t ld, int64_t n, int* info_dev, double* W, cudaStream_t stream,
                    const cudaEvent_t* column_done = nullptr) {
  SC_TRY(set_smem_limits());
  const size_t smem2 = 2 * kNB * kLD * sizeof(double);
  for (int64_t j = 0; j < n; j += kNB) {
    const int jb = int(std::min<int64_t>(kNB, n - j));
    const int64_t m = n - j - jb;
    potrf_inv_kernel<<<1, kThreads, smem2, stream>>>(A + j + j * ld, ld, jb, j, L + j + j * ld, ld, W, info_dev);
    SC_CUDA(cudaGetLastError());
    counters().launches.fetch_add(1, std::memory_order_relaxed);
    if (m > 0) {
      // panel = A_panel * W'  (m x jb) ; trailing -= panel * panel' (lower)
      SC_TRY(dgemm(dev, false, true, false, m, jb, jb, 1.0, A + (j + jb) + j * ld, ld, W, kNB, 0.0, L + (j + jb) + j * ld, ld, stream));
      if (column_done) SC_CUDA(cudaEventRecord(column_done[j / kNB], stream));
      SC_TRY(dgemm(dev, false, true, true, m, m, jb, -1.0, L + (j + jb) + j * ld, ld, L + (j + jb) + j * ld, ld, 1.0,
                   A + (j + jb) + (j + jb) * ld, ld, stream));
    } else if (column_done) {
      SC_CUDA(cudaEventRecord(column_done[j / kNB], stream));
    }
  }
  return STAN_CUDA_OK;
}

int cholesky_host(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl, double symmetry_tolerance) {
  NvtxRange range("stan_cuda_cholesky_decompose");
  if (n < 0 || (n > 0 && (!a || !l || lda < n || ldl < n))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_cholesky_decompose: bad argument");
  if (n == 0) return STAN_CUDA_OK;
  SC_TRY(ensure_init());
  const int dev = scope_device_ordinal();
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  const int64_t ld = padded(n);
  DeviceBuf A, L, W, info;
  SC_TRY(A.alloc(size_t(ld) * n, td->stream));
  SC_TRY(L.alloc(size_t(ld) * n, td->stream));
  SC_TRY(W.alloc(kNB * kNB, td->stream));
  SC_TRY(info.alloc(2, td->stream));
  SC_CUDA(cudaMemsetAsync(L.p, 0, size_t(ld) * n * sizeof(double), td->stream));
  SC_CUDA(cudaMemsetAsync(info.p, 0, 2 * sizeof(double), td->stream));
  SC_TRY(dense_copy_to_device(A.p, ld, a, lda, n, n, td->stream));
  counters().bytes_uploaded.fetch_add(uint64_t(n) * uint64_t(n) * sizeof(double));
  if (symmetry_tolerance >= 0.0 && n > 1) {
    const unsigned tiles = unsigned((n + 31) / 32);
    symmetry_kernel<<<dim3(tiles, tiles), 256, 0, td->stream>>>(A.p, ld, n, symmetry_tolerance, reinterpret_cast<int*>(info.p));
    SC_CUDA(cudaGetLastError());
    counters().launches.fetch_add(1, std::memory_order_relaxed);
  }
  SC_TRY(td->reserve_host(8));
  const int64_t nblk = (n + kNB - 1) / kNB;
  // chunks of about 2 MiB, whole block columns: the first one leaves while the sweep is a few steps in
  const int64_t chunk_cols = std::max<int64_t>(kNB, (int64_t(1) << 18) / n / kNB * kNB);
  if (download_overlap_enabled() && nblk > 1 && chunk_cols < n) {
    static thread_local DownloadOverlap ov;
    SC_TRY(ov.prepare(dev, size_t(nblk)));
    SC_TRY(cholesky_device(dev, A.p, L.p, ld, n, reinterpret_cast<int*>(info.p), W.p, td->stream, ov.done.data()));
    SC_CUDA(cudaMemcpyAsync(td->h_pinned, info.p, sizeof(double), cudaMemcpyDeviceToHost, td->stream));
    const std::function<int(int64_t)> gate = [&](int64_t col_end) -> int {
      SC_CUDA(cudaStreamWaitEvent(ov.copy, ov.done[size_t((col_end - 1) / kNB)], 0));
      return STAN_CUDA_OK;
    };
    SC_TRY(dense_copy_to_host_gated(l, ldl, L.p, ld, n, n, ov.copy, chunk_cols, gate, 1));  // lower triangle; synchronises the copy stream
    SC_CUDA(cudaStreamSynchronize(td->stream));  // the flags; the buffers are released on this stream
  } else {
    SC_TRY(cholesky_device(dev, A.p, L.p, ld, n, reinterpret_cast<int*>(info.p), W.p, td->stream));
    SC_CUDA(cudaMemcpyAsync(td->h_pinned, info.p, sizeof(double), cudaMemcpyDeviceToHost, td->stream));
    SC_TRY(dense_copy_to_host_tri(l, ldl, L.p, ld, n, td->stream, 1));  // synchronises
  }
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  const int* info_host = reinterpret_cast<const int*>(td->h_pinned);  // [0]: pivot, [1]: symmetry
  if (info_host[1] != 0) return STAN_CUDA_NOT_SYMMETRIC;  // the reference checks symmetry first (:60-63)
  return info_host[0] != 0 ? STAN_CUDA_NOT_POS_DEF : STAN_CUDA_OK;
}

// a_adj (lower triangle, n x n) = adjoint of A given L and the adjoint of L (lower triangles are read)
int cholesky_adjoint_host(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj, double* a_adj, int64_t lda) {
  NvtxRange range("stan_cuda_cholesky_adjoint");
  if (n < 0 || (n > 0 && (!l || !l_adj || !a_adj || ldl < n || ldadj < n || lda < n)))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_cholesky_adjoint: bad argument");
  if (n == 0) return STAN_CUDA_OK;
  SC_TRY(ensure_init());
  SC_TRY(set_smem_limits());
  const int dev = scope_device_ordinal();
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  cudaStream_t s = td->stream;
  const int64_t ld = padded(n);
  DeviceBuf L, Lb, W, Dsym, Tmp;
  SC_TRY(L.alloc(size_t(ld) * n, s));
  SC_TRY(Lb.alloc(size_t(ld) * n, s));
  const int64_t nblk = (n + kNB - 1) / kNB;
  SC_TRY(W.alloc(size_t(nblk) * kNB * kNB, s));
  SC_TRY(Dsym.alloc(kNB * kNB, s));
  SC_TRY(Tmp.alloc(size_t(ld) * kNB, s));
  // both operands are lower triangular: only the triangles travel, into cleared buffers
  SC_CUDA(cudaMemsetAsync(L.p, 0, size_t(ld) * n * sizeof(double), s));
  SC_CUDA(cudaMemsetAsync(Lb.p, 0, size_t(ld) * n * sizeof(double), s));
  SC_TRY(dense_copy_to_device_tri(L.p, ld, l, ldl, n, s, 1));
  SC_TRY(dense_copy_to_device_tri(Lb.p, ld, l_adj, ldadj, n, s, 1));
  counters().bytes_uploaded.fetch_add(uint64_t(n) * uint64_t(n + 1) * sizeof(double));
  const size_t smem2 = 2 * kNB * kLD * sizeof(double), smem4 = 4 * kNB * kLD * sizeof(double);
  tri_inv_blocks_kernel<<<unsigned(nblk), kThreads, smem2, s>>>(L.p, ld, n, 1, W.p);  // D_b^-1 for every diagonal block
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  for (int64_t b = nblk - 1; b >= 0; --b) {
    const int64_t j = b * kNB, k = std::min<int64_t>(n, j + kNB);
    const int kj = int(k - j);
    const int64_t mk = n - k;
    double *D = L.p + j + j * ld, *R = L.p + j, *B = L.p + k, *C = L.p + k + j * ld;
    double *Db = Lb.p + j + j * ld, *Rb = Lb.p + j, *Bb = Lb.p + k, *Cb = Lb.p + k + j * ld;
    const double* Wb = W.p + b * kNB * kNB;
    if (mk > 0) {
      // Cb = Cb D^-1                                                         (:147-151 / :358)
      SC_TRY(dgemm(dev, false, false, false, mk, kj, kj, 1.0, Cb, ld, Wb, kNB, 0.0, Tmp.p, ld, s));
      SC_CUDA(cudaMemcpy2DAsync(Cb, size_t(ld) * sizeof(double), Tmp.p, size_t(ld) * sizeof(double), size_t(mk) * sizeof(double),
                                size_t(kj), cudaMemcpyDeviceToDevice, s));
      if (j > 0) SC_TRY(dgemm(dev, false, false, false, mk, j, kj, -1.0, Cb, ld, R, ld, 1.0, Bb, ld, s));  // Bb -= Cb R   (:152 / :359)
      SC_TRY(dgemm(dev, true, false, false, kj, kj, mk, -1.0, Cb, ld, C, ld, 1.0, Db, ld, s));            // Db -= Cb' C  (:153 / :360)
    }
    symbolic_rev_kernel<<<1, kThreads, smem4, s>>>(D, ld, Wb, Db, ld, kj, Dsym.p);                         // :155 / :362
    SC_CUDA(cudaGetLastError());
    counters().launches.fetch_add(1, std::memory_order_relaxed);
    if (j > 0) {
      if (mk > 0) SC_TRY(dgemm(dev, true, false, false, kj, j, mk, -1.0, Cb, ld, B, ld, 1.0, Rb, ld, s));  // Rb -= Cb' B   (:156 / :364)
      SC_TRY(dgemm(dev, false, false, false, kj, j, kj, -1.0, Dsym.p, kNB, R, ld, 1.0, Rb, ld, s));        // Rb -= Db_sym R (:157 / :364)
    }
  }
  // the strict upper triangle holds scratch of the symmetric updates: the operands are the lower triangle of A only
  zero_strict_upper_kernel<<<dim3(unsigned((n + 255) / 256), unsigned((n + 63) / 64)), 256, 0, s>>>(Lb.p, ld, n);
  SC_CUDA(cudaGetLastError());
  SC_TRY(dense_copy_to_host_tri(a_adj, lda, Lb.p, ld, n, s, 1));
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

// x (n x m) = op(A)^-1 b for a triangular A (lower != 0: the lower triangle is read, else the upper), op = transpose
// when trans != 0.  Blocked substitution: all diagonal blocks are inverted in one launch, then for each block row i
// (top down for lower / no transpose and upper / transpose, bottom up otherwise)
//   X_i = op(A_ii)^-1 X_i      a 64 x m product with the block inverse
//   X_r -= op(A)[r, i] X_i     for the rows r still to be solved: one product with inner dimension 64
// The reference's OpenCL path forms the whole inverse (tri_inverse) and multiplies (prim/mat/fun/mdivide_left_tri.hpp
// :63-72); substitution does a third of the work and keeps the error at that of the 64 x 64 blocks.
int tri_solve_host(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b, int64_t ldb, double* x,
                   int64_t ldx) {
  NvtxRange range("stan_cuda_tri_solve");
  if (n < 0 || m < 0 || (n > 0 && (!a || lda < n)) || (n > 0 && m > 0 && (!b || !x || ldb < n || ldx < n)))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_tri_solve: bad argument");
  if (n == 0 || m == 0) return STAN_CUDA_OK;
  SC_TRY(ensure_init());
  SC_TRY(set_smem_limits());
  const int dev = scope_device_ordinal();
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  cudaStream_t s = td->stream;
  const int64_t ld = padded(n), nblk = (n + kNB - 1) / kNB;
  DeviceBuf A, X, W, Tmp;
  SC_TRY(A.alloc(size_t(ld) * n, s));
  SC_TRY(X.alloc(size_t(ld) * m, s));
  SC_TRY(W.alloc(size_t(nblk) * kNB * kNB, s));
  SC_TRY(Tmp.alloc(size_t(kNB) * m, s));
  SC_CUDA(cudaMemsetAsync(A.p, 0, size_t(ld) * n * sizeof(double), s));
  SC_TRY(dense_copy_to_device_tri(A.p, ld, a, lda, n, s, lower ? 1 : 2));  // the triangle that is read
  SC_TRY(dense_copy_to_device(X.p, ld, b, ldb, n, m, s));
  counters().bytes_uploaded.fetch_add((uint64_t(n) * uint64_t(n + 1) / 2 + uint64_t(n) * uint64_t(m)) * sizeof(double));
  tri_inv_blocks_kernel<<<unsigned(nblk), kThreads, 2 * kNB * kLD * sizeof(double), s>>>(A.p, ld, n, lower, W.p);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  const bool top_down = (lower != 0) != (trans != 0);
  const bool trans_w = lower ? trans != 0 : trans == 0;  // W_b of an upper block is the transpose of its inverse
  for (int64_t step = 0; step < nblk; ++step) {
    const int64_t bi = top_down ? step : nblk - 1 - step;
    const int64_t i0 = bi * kNB, ib = std::min<int64_t>(kNB, n - i0);
    double* Xi = X.p + i0;
    SC_TRY(dgemm(dev, trans_w, false, false, ib, m, ib, 1.0, W.p + bi * kNB * kNB, kNB, Xi, ld, 0.0, Tmp.p, kNB, s));
    SC_CUDA(cudaMemcpy2DAsync(Xi, size_t(ld) * sizeof(double), Tmp.p, size_t(kNB) * sizeof(double), size_t(ib) * sizeof(double),
                              size_t(m), cudaMemcpyDeviceToDevice, s));
    // rows still to be solved and the block of op(A) that couples them to block row i
    const int64_t r0 = top_down ? i0 + ib : 0, nr = top_down ? n - i0 - ib : i0;
    if (nr > 0) {
      // not transposed: A[r, i] (rows r0.., columns i0..);  transposed: A[i, r]' (rows i0.., columns r0..)
      const double* Ari = trans ? A.p + i0 + r0 * ld : A.p + r0 + i0 * ld;
      SC_TRY(dgemm(dev, trans != 0, false, false, nr, m, ib, -1.0, Ari, ld, Tmp.p, kNB, 1.0, X.p + r0, ld, s));
    }
  }
  SC_TRY(dense_copy_to_host(x, ldx, X.p, ld, n, m, s));
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda

extern "C" {
int stan_cuda_cholesky_decompose(int64_t n, const double* a, int64_t lda, double* l, int64_t ldl, double symmetry_tolerance) {
  return stancuda::cholesky_host(n, a, lda, l, ldl, symmetry_tolerance);
}
int stan_cuda_cholesky_adjoint(int64_t n, const double* l, int64_t ldl, const double* l_adj, int64_t ldadj, double* a_adj, int64_t lda) {
  return stancuda::cholesky_adjoint_host(n, l, ldl, l_adj, ldadj, a_adj, lda);
}
int stan_cuda_tri_solve(int lower, int trans, int64_t n, int64_t m, const double* a, int64_t lda, const double* b, int64_t ldb, double* x,
                        int64_t ldx) {
  return stancuda::tri_solve_host(lower, trans, n, m, a, lda, b, ldb, x, ldx);
}
}
