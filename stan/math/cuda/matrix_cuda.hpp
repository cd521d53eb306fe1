#ifndef STAN_MATH_CUDA_MATRIX_CUDA_HPP
#define STAN_MATH_CUDA_MATRIX_CUDA_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/c_api.h>
#include <stan/math/cuda/err.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/scal/err/check_size_match.hpp>
#include <cstdint>
#include <utility>

namespace stan {
namespace math {
namespace cuda {

/**
 * A dense FP64 matrix in device memory: the counterpart of the OpenCL backend's matrix_cl<double>
 * (stan/math/opencl/matrix_cl.hpp:32-454) for the `multiply` hooks.  Column-major like Eigen.
 * Construction from an Eigen matrix uploads (matrix_cl(A), to_matrix_cl: stan/math/opencl/copy.hpp:37-58),
 * from_matrix_cuda downloads (from_matrix_cl, :69-93), operator* / multiply is the product and transpose(m) a view
 * of the transposed matrix (stan/math/opencl/multiply.hpp, transpose.hpp).  Move-only: the device block is owned.
 */
class matrix_cuda {
 public:
  matrix_cuda() : h_(nullptr) {}
  matrix_cuda(int64_t rows, int64_t cols) : h_(nullptr) {
    check_cuda_rc("matrix_cuda", stan_cuda_matrix_create(rows, cols, &h_));
  }
  /** rows x cols column-major doubles at `data` */
  matrix_cuda(const double* data, int64_t rows, int64_t cols) : h_(nullptr) {
    check_cuda_rc("matrix_cuda", stan_cuda_matrix_create(rows, cols, &h_));
    const int rc = stan_cuda_matrix_upload(h_, data, rows > 0 ? rows : 1);
    if (rc < 0) {
      stan_cuda_matrix_free(h_);
      h_ = nullptr;
      check_cuda_rc("matrix_cuda", rc);
    }
  }
  template <int R, int C>
  explicit matrix_cuda(const Eigen::Matrix<double, R, C>& A) : matrix_cuda(A.data(), A.rows(), A.cols()) {}
  matrix_cuda(const matrix_cuda&) = delete;
  matrix_cuda& operator=(const matrix_cuda&) = delete;
  matrix_cuda(matrix_cuda&& o) noexcept : h_(o.h_) { o.h_ = nullptr; }
  matrix_cuda& operator=(matrix_cuda&& o) noexcept {
    if (this != &o) {
      stan_cuda_matrix_free(h_);
      h_ = o.h_;
      o.h_ = nullptr;
    }
    return *this;
  }
  ~matrix_cuda() { stan_cuda_matrix_free(h_); }

  int64_t rows() const { return stan_cuda_matrix_rows(h_); }
  int64_t cols() const { return stan_cuda_matrix_cols(h_); }
  int64_t size() const { return rows() * cols(); }
  stan_cuda_matrix* handle() const { return h_; }
  /** copies the matrix into rows() x cols() column-major doubles at `out` */
  void download(double* out) const {
    check_cuda_rc("from_matrix_cuda", stan_cuda_matrix_download(h_, out, rows() > 0 ? rows() : 1));
  }

 private:
  stan_cuda_matrix* h_;
};

/** view of a device matrix as its transpose (no data moves until it is multiplied) */
struct transposed_matrix_cuda {
  const matrix_cuda& m;
};
inline transposed_matrix_cuda transpose(const matrix_cuda& m) { return transposed_matrix_cuda{m}; }

namespace detail {
inline matrix_cuda multiply_impl(const matrix_cuda& a, bool ta, const matrix_cuda& b, bool tb) {
  const int64_t m = ta ? a.cols() : a.rows(), ka = ta ? a.rows() : a.cols();
  const int64_t kb = tb ? b.cols() : b.rows(), n = tb ? b.rows() : b.cols();
  check_size_match("multiply (CUDA)", "Columns of ", "A", ka, "Rows of ", "B", kb);  // opencl/multiply.hpp: same check
  matrix_cuda c(m, n);
  check_cuda_rc("multiply (CUDA)", stan_cuda_matrix_multiply(a.handle(), ta ? 1 : 0, b.handle(), tb ? 1 : 0, c.handle()));
  return c;
}
}  // namespace detail

inline matrix_cuda multiply(const matrix_cuda& a, const matrix_cuda& b) { return detail::multiply_impl(a, false, b, false); }
inline matrix_cuda multiply(const transposed_matrix_cuda& a, const matrix_cuda& b) {
  return detail::multiply_impl(a.m, true, b, false);
}
inline matrix_cuda multiply(const matrix_cuda& a, const transposed_matrix_cuda& b) {
  return detail::multiply_impl(a, false, b.m, true);
}
inline matrix_cuda multiply(const transposed_matrix_cuda& a, const transposed_matrix_cuda& b) {
  return detail::multiply_impl(a.m, true, b.m, true);
}
inline matrix_cuda operator*(const matrix_cuda& a, const matrix_cuda& b) { return multiply(a, b); }
inline matrix_cuda operator*(const transposed_matrix_cuda& a, const matrix_cuda& b) { return multiply(a, b); }
inline matrix_cuda operator*(const matrix_cuda& a, const transposed_matrix_cuda& b) { return multiply(a, b); }
inline matrix_cuda operator*(const transposed_matrix_cuda& a, const transposed_matrix_cuda& b) { return multiply(a, b); }

/** from_matrix_cl (stan/math/opencl/copy.hpp:69-93) */
template <int R = Eigen::Dynamic, int C = Eigen::Dynamic>
inline Eigen::Matrix<double, R, C> from_matrix_cuda(const matrix_cuda& src) {
  Eigen::Matrix<double, R, C> dst(src.rows(), src.cols());
  if (src.size() > 0) {
    src.download(dst.data());
  }
  return dst;
}

/** to_matrix_cl (stan/math/opencl/copy.hpp:37-58) */
template <int R, int C>
inline matrix_cuda to_matrix_cuda(const Eigen::Matrix<double, R, C>& src) {
  return matrix_cuda(src);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
