// gp.cu -- squared-exponential covariance matrix on the device.
//
// Replaces, for the gp_exp_quad_cov hooks of the reference (stan/math/prim/mat/fun/gp_exp_quad_cov.hpp:344-655, six
// double specialisations under #ifdef STAN_OPENCL), the OpenCL kernels gp_exp_quad_cov / gp_exp_quad_cov_cross
// (stan/math/opencl/kernels/gp_exp_quad_cov.hpp):  cov(i, j) = sigma_sq * exp(factor * |x1_i - x2_j|^2)
// (internal::gp_exp_quad_cov, :40-170: factor = -0.5 / l^2 for a scalar length scale, -0.5 with the points already
// divided by a vector of length scales; the diagonal of the one-sample form is sigma_sq exactly, :53, :88).
// One thread per element, points stored dimension-major so that a warp reads consecutive x1 and one x2; the kernel
// is bound by its n1 * n2 * 8 bytes of output (and the call by their way home over PCIe).
#include <algorithm>

#include "context.cuh"

namespace stancuda {

int scope_device_ordinal();  // api.cu
int dense_copy_to_host(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols, cudaStream_t stream);

namespace {

__global__ void __launch_bounds__(256) gp_exp_quad_cov_kernel(const double* __restrict__ x1, const double* __restrict__ x2, int64_t n1,
                                                              int64_t n2, int dim, double sigma_sq, double factor, int symmetric,
                                                              double* __restrict__ cov, int64_t ldc) {
  const int64_t i = blockIdx.x * int64_t(32) + (threadIdx.x & 31);
  const int64_t j0 = (blockIdx.y * int64_t(8) + (threadIdx.x >> 5)) * 4;
  if (i >= n1) return;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int64_t j = j0 + q;
    if (j >= n2) return;
    double v;
    if (symmetric && i == j) {
      v = sigma_sq;
    } else {
      double d2 = 0.0;
      for (int d = 0; d < dim; ++d) {
        const double diff = x1[int64_t(d) * n1 + i] - x2[int64_t(d) * n2 + j];
        d2 = fma(diff, diff, d2);
      }
      v = sigma_sq * exp(d2 * factor);
    }
    cov[j * ldc + i] = v;
  }
}

}  // namespace

int gp_exp_quad_cov_host(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2, double sigma_sq, double factor,
                         double* cov, int64_t ldc) {
  NvtxRange range("stan_cuda_gp_exp_quad_cov");
  if (n1 < 0 || n2 < 0 || dim < 1 || dim > (1 << 20) || (n1 > 0 && !x1) || (n1 > 0 && n2 > 0 && (!cov || ldc < n1)))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_gp_exp_quad_cov: bad argument");
  const int symmetric = x2 == nullptr;
  if (symmetric) {
    x2 = x1;
    n2 = n1;
  }
  if (n1 == 0 || n2 == 0) return STAN_CUDA_OK;
  SC_TRY(ensure_init());
  const int dev = scope_device_ordinal();
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  cudaStream_t s = td->stream;
  const int64_t ld = (n1 + 15) / 16 * 16;
  double *d_x1 = nullptr, *d_x2 = nullptr, *d_cov = nullptr;
  SC_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&d_x1), size_t(n1) * dim * sizeof(double), s));
  cudaError_t e = cudaMemcpyAsync(d_x1, x1, size_t(n1) * dim * sizeof(double), cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess && !symmetric) {
    e = cudaMallocAsync(reinterpret_cast<void**>(&d_x2), size_t(n2) * dim * sizeof(double), s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_x2, x2, size_t(n2) * dim * sizeof(double), cudaMemcpyHostToDevice, s);
  }
  if (e == cudaSuccess) e = cudaMallocAsync(reinterpret_cast<void**>(&d_cov), size_t(ld) * n2 * sizeof(double), s);
  int rc = STAN_CUDA_OK;
  if (e == cudaSuccess) {
    const dim3 grid(unsigned((n1 + 31) / 32), unsigned((n2 + 31) / 32));
    if (grid.y > 65535) {
      rc = fail(STAN_CUDA_ERR_ARG, "stan_cuda_gp_exp_quad_cov: more than 2097120 columns");
    } else {
      gp_exp_quad_cov_kernel<<<grid, 256, 0, s>>>(d_x1, symmetric ? d_x1 : d_x2, n1, n2, int(dim), sigma_sq, factor, symmetric, d_cov, ld);
      e = cudaGetLastError();
      counters().launches.fetch_add(1, std::memory_order_relaxed);
      counters().bytes_uploaded.fetch_add(uint64_t(n1 + (symmetric ? 0 : n2)) * uint64_t(dim) * sizeof(double));
    }
  }
  if (e != cudaSuccess) rc = fail_cuda(e, "stan_cuda_gp_exp_quad_cov", __FILE__, __LINE__);
  if (rc >= 0) rc = dense_copy_to_host(cov, ldc, d_cov, ld, n1, n2, s);
  if (d_x1) cudaFreeAsync(d_x1, s);
  if (d_x2) cudaFreeAsync(d_x2, s);
  if (d_cov) cudaFreeAsync(d_cov, s);
  if (rc >= 0) counters().evals.fetch_add(1, std::memory_order_relaxed);
  return rc;
}

}  // namespace stancuda

extern "C" int stan_cuda_gp_exp_quad_cov(int64_t n1, int64_t n2, int64_t dim, const double* x1, const double* x2, double sigma_sq,
                                         double factor, double* cov, int64_t ldc) {
  return stancuda::gp_exp_quad_cov_host(n1, n2, dim, x1, x2, sigma_sq, factor, cov, ldc);
}
