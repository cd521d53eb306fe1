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
// Product: C = op(A) * op(B) in FP64 on the DFMA pipe.  On B200 the FP64 tensor rate equals the vector rate
// (~64 DFMA / clk / SM), so this is a classic register-tiled kernel rather than a DMMA one: a CTA of 256 threads owns
// a BM x BN tile of C, walks K in steps of BK = 8 through two shared-memory stages (the global loads of step t + 1
// are issued before the arithmetic of step t and stored after it: one barrier per step), and each thread
// accumulates a TM x TN block in registers.  op(A) must be M-contiguous and op(B) K-contiguous (plain column-major
// A and B); a transposed operand is first transposed on the device (2 x its bytes of HBM traffic against 2 M N K
// flops: noise for any shape worth sending here).  The roofline is the FP64 issue rate, not HBM.
#include <algorithm>

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

// ---- transpose: out(c, r) = in(r, c); 32 x 32 tiles through shared memory ---------------------------------
__global__ void __launch_bounds__(256) transpose_kernel(const double* __restrict__ in, int64_t ld_in, int64_t rows, int64_t cols,
                                                        double* __restrict__ out, int64_t ld_out) {
  __shared__ double tile[32][33];
  const int64_t r0 = int64_t(blockIdx.x) * 32, c0 = int64_t(blockIdx.y) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int64_t r = r0 + tx, c = c0 + ty + j;
    tile[ty + j][tx] = (r < rows && c < cols) ? in[c * ld_in + r] : 0.0;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int64_t c = c0 + tx, r = r0 + ty + j;  // element (r, c) of `in` goes to out(c, r)
    if (r < rows && c < cols) out[r * ld_out + c] = tile[tx][ty + j];
  }
}

// ---- C = A * B, column-major, A: M x K (ld lda), B: K x N (ld ldb), C: M x N (ld ldc) ----------------------
// Thread layout: tx = tid % (BM / TMH) along M, ty = tid / (BM / TMH) along N, where a thread owns rows
// {tx * 4 .. tx * 4 + 3} (+ BM / 2 for the second half when TM == 8) and the columns likewise.
template <int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__(kGemmThreads, TM == 8 ? 1 : 2)
dgemm_nn_kernel(int64_t M, int64_t N, int64_t K, const double* __restrict__ A, int64_t lda, const double* __restrict__ B,
                int64_t ldb, double* __restrict__ C, int64_t ldc) {
  static_assert((BM / TM) * (BN / TN) == kGemmThreads, "one thread per TM x TN block");
  static_assert(TM % 4 == 0 && TN % 4 == 0, "blocks of four");
  constexpr int LDB_S = BN + 2;                       // +2: the transposing stores of the B tile hit distinct banks
  constexpr int LA = BM * kBK / kGemmThreads;         // doubles of the A tile per thread (contiguous along M)
  constexpr int LB = BN * kBK / kGemmThreads;         // doubles of the B tile per thread (contiguous along K)
  static_assert(LA == 2 || LA == 4, "A tile: one or two 16-byte loads per thread");
  static_assert(LB == 2 || LB == 4, "B tile: one or two 16-byte loads per thread");
  __shared__ __align__(16) double As[2][kBK][BM];
  __shared__ __align__(16) double Bs[2][kBK][LDB_S];

  const int tid = threadIdx.x;
  const int64_t m0 = int64_t(blockIdx.x) * BM, n0 = int64_t(blockIdx.y) * BN;
  // global -> shared assignment
  const int a_k = tid / (BM / LA), a_m = (tid % (BM / LA)) * LA;  // A(m0 + a_m .. + LA - 1, k0 + a_k)
  const int b_n = tid / (kBK / LB), b_k = (tid % (kBK / LB)) * LB;  // B(k0 + b_k .. + LB - 1, n0 + b_n)
  const bool a_row_ok = m0 + a_m < M;  // rows beyond M inside the leading dimension are zero; beyond it, nothing to read
  const bool b_col_ok = n0 + b_n < N;
  const double* a_ptr = A + (m0 + a_m);
  const double* b_ptr = B + (n0 + b_n) * ldb;
  // compute assignment
  constexpr int TXN = BM / TM;  // threads along M
  const int tx = tid % TXN, ty = tid / TXN;

  double acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.0;

  double ra[LA], rb[LB];
  auto load_global = [&](int64_t k0) {
    const int64_t ka = k0 + a_k;
    if (a_row_ok && ka < K) {
      // lda is a multiple of 16 and a_m of LA: 16-byte aligned; rows in [M, lda) are zero by construction
      const double2* p = reinterpret_cast<const double2*>(a_ptr + ka * lda);
#pragma unroll
      for (int q = 0; q < LA / 2; ++q) {
        const double2 v = (m0 + a_m + 2 * q < lda) ? __ldg(p + q) : make_double2(0.0, 0.0);
        ra[2 * q] = v.x;
        ra[2 * q + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int q = 0; q < LA; ++q) ra[q] = 0.0;
    }
    const int64_t kb = k0 + b_k;
    if (b_col_ok && kb < K) {
      // rows of B in [K, ldb) are zero by construction; kb is a multiple of LB
      const double2* p = reinterpret_cast<const double2*>(b_ptr + kb);
#pragma unroll
      for (int q = 0; q < LB / 2; ++q) {
        const double2 v = (kb + 2 * q < ldb) ? __ldg(p + q) : make_double2(0.0, 0.0);
        rb[2 * q] = v.x;
        rb[2 * q + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int q = 0; q < LB; ++q) rb[q] = 0.0;
    }
  };
  auto store_shared = [&](int buf) {
#pragma unroll
    for (int q = 0; q < LA / 2; ++q)
      *reinterpret_cast<double2*>(&As[buf][a_k][a_m + 2 * q]) = make_double2(ra[2 * q], ra[2 * q + 1]);
#pragma unroll
    for (int q = 0; q < LB; ++q) Bs[buf][b_k + q][b_n] = rb[q];
  };

  const int64_t steps = (K + kBK - 1) / kBK;
  load_global(0);
  store_shared(0);
  __syncthreads();
  for (int64_t t = 0; t < steps; ++t) {
    const int buf = int(t & 1);
    if (t + 1 < steps) load_global((t + 1) * kBK);
#pragma unroll
    for (int kk = 0; kk < kBK; ++kk) {
      double a[TM], b[TN];
#pragma unroll
      for (int h = 0; h < TM / 4; ++h) {
        const double2 v0 = *reinterpret_cast<const double2*>(&As[buf][kk][h * (BM / 2) + tx * 4]);
        const double2 v1 = *reinterpret_cast<const double2*>(&As[buf][kk][h * (BM / 2) + tx * 4 + 2]);
        a[h * 4] = v0.x; a[h * 4 + 1] = v0.y; a[h * 4 + 2] = v1.x; a[h * 4 + 3] = v1.y;
      }
#pragma unroll
      for (int h = 0; h < TN / 4; ++h) {
        const double2 v0 = *reinterpret_cast<const double2*>(&Bs[buf][kk][h * (BN / 2) + ty * 4]);
        const double2 v1 = *reinterpret_cast<const double2*>(&Bs[buf][kk][h * (BN / 2) + ty * 4 + 2]);
        b[h * 4] = v0.x; b[h * 4 + 1] = v0.y; b[h * 4 + 2] = v1.x; b[h * 4 + 3] = v1.y;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    if (t + 1 < steps) store_shared(buf ^ 1);
    __syncthreads();
  }

  // epilogue: four consecutive rows per (half, column) -> two 16-byte stores
#pragma unroll
  for (int hj = 0; hj < TN / 4; ++hj)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t n = n0 + hj * (BN / 2) + ty * 4 + j;
      if (n >= N) continue;
#pragma unroll
      for (int hi = 0; hi < TM / 4; ++hi) {
        const int64_t m = m0 + hi * (BM / 2) + tx * 4;
        double* c = C + n * ldc + m;
        const int r = hi * 4, cj = hj * 4 + j;  // rows r .. r + 3 of column cj of the thread's block
        if (m + 3 < M) {
          *reinterpret_cast<double2*>(c) = make_double2(acc[r][cj], acc[r + 1][cj]);
          *reinterpret_cast<double2*>(c + 2) = make_double2(acc[r + 2][cj], acc[r + 3][cj]);
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (m + i < M) c[i] = acc[r + i][cj];
        }
      }
    }
}

template <int BM, int BN, int TM, int TN>
int launch_dgemm(int64_t M, int64_t N, int64_t K, const double* A, int64_t lda, const double* B, int64_t ldb, double* C,
                 int64_t ldc, cudaStream_t stream) {
  const int64_t gx = (M + BM - 1) / BM, gy = (N + BN - 1) / BN;
  if (gy > 65535) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: more than 65535 column tiles");
  dgemm_nn_kernel<BM, BN, TM, TN><<<dim3(unsigned(gx), unsigned(gy)), kGemmThreads, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc);
  SC_CUDA(cudaGetLastError());
  counters().launches.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

int dgemm_nn(int device, int64_t M, int64_t N, int64_t K, const double* A, int64_t lda, const double* B, int64_t ldb, double* C,
             int64_t ldc, cudaStream_t stream) {
  // 128 x 128 tiles when they fill the machine at least once, 64 x 64 below that
  const int64_t big_tiles = ((M + 127) / 128) * ((N + 127) / 128);
  if (big_tiles >= sm_count(device)) return launch_dgemm<128, 128, 8, 8>(M, N, K, A, lda, B, ldb, C, ldc, stream);
  return launch_dgemm<64, 64, 4, 4>(M, N, K, A, lda, B, ldb, C, ldc, stream);
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
  SC_CUDA(cudaMemcpy2DAsync(m->d, size_t(m->ld) * sizeof(double), host, size_t(ld) * sizeof(double), size_t(m->rows) * sizeof(double),
                            size_t(m->cols), cudaMemcpyHostToDevice, td->stream));
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
  SC_CUDA(cudaMemcpy2DAsync(host, size_t(ld) * sizeof(double), m->d, size_t(m->ld) * sizeof(double), size_t(m->rows) * sizeof(double),
                            size_t(m->cols), cudaMemcpyDeviceToHost, td->stream));
  SC_CUDA(cudaStreamSynchronize(td->stream));
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
  stan_cuda_matrix* at = nullptr;
  stan_cuda_matrix* bt = nullptr;
  int rc = STAN_CUDA_OK;
  auto transposed = [&](const stan_cuda_matrix* src, stan_cuda_matrix** dst) -> int {
    SC_TRY(matrix_create_on(src->device, src->cols, src->rows, true, dst));
    const dim3 grid(unsigned((src->rows + 31) / 32), unsigned((src->cols + 31) / 32));
    if (grid.y > 65535) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_matrix_multiply: transpose of more than 2097120 columns");
    transpose_kernel<<<grid, 256, 0, td->stream>>>(src->d, src->ld, src->rows, src->cols, (*dst)->d, (*dst)->ld);
    SC_CUDA(cudaGetLastError());
    counters().launches.fetch_add(1, std::memory_order_relaxed);
    return STAN_CUDA_OK;
  };
  if (trans_a) rc = transposed(a, &at);
  if (rc >= 0 && trans_b) rc = transposed(b, &bt);
  if (rc >= 0) {
    const stan_cuda_matrix* aa = at ? at : a;
    const stan_cuda_matrix* bb = bt ? bt : b;
    rc = dgemm_nn(c->device, M, N, K, aa->d, aa->ld, bb->d, bb->ld, c->d, c->ld, td->stream);
  }
  if (at) matrix_free(at);  // stream-ordered: freed after the product that reads it
  if (bt) matrix_free(bt);
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
    SC_CUDA(cudaMemcpy2DAsync(m->d, size_t(m->ld) * sizeof(double), host, size_t(ld) * sizeof(double), size_t(m->rows) * sizeof(double),
                              size_t(m->cols), cudaMemcpyHostToDevice, td->stream));
    counters().bytes_uploaded.fetch_add(uint64_t(m->rows) * uint64_t(m->cols) * sizeof(double));
    return STAN_CUDA_OK;
  };
  if (rc >= 0) rc = copy_in(a, g->a, g->lda);
  if (rc >= 0) rc = copy_in(b, g->b, g->ldb);
  if (rc >= 0) rc = matrix_multiply(a, g->trans_a, b, g->trans_b, c, false);
  if (rc >= 0) {
    cudaError_t e = cudaMemcpy2DAsync(g->c, size_t(g->ldc) * sizeof(double), c->d, size_t(c->ld) * sizeof(double),
                                      size_t(g->m) * sizeof(double), size_t(g->n), cudaMemcpyDeviceToHost, td->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(td->stream);
    if (e != cudaSuccess) rc = fail_cuda(e, "download of the product", __FILE__, __LINE__);
  }
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
