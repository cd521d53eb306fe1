// oracle/ref_shim_dense.cpp -- TEST INFRASTRUCTURE, third translation unit of _ref/libstanref.so.
//
// The UNMODIFIED reference's dense `multiply` (stan/math/prim/mat/fun/multiply.hpp:66-83: Eigen's product, and
// stan/math/rev/mat/fun/multiply.hpp:35-150, 225-328, 395-501, 619-633: multiply_mat_vari) behind C entry points, so
// the parity tests can check the device product and its reverse-mode adjoints against the reference itself.
// Compiled WITHOUT STAN_OPENCL / STAN_CUDA: the reference's CPU path.  Nothing is copied.
#include <stan/math/rev/mat.hpp>

#include <chrono>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace {
using stan::math::var;
typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> matrix_d;
typedef Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic> matrix_v;

template <typename F>
int guarded(F&& f, char* err, int errlen) {
  int rc = 0;
  try {
    f();
  } catch (const std::exception& e) {
    if (err && errlen > 0) {
      std::strncpy(err, e.what(), size_t(errlen) - 1);
      err[errlen - 1] = 0;
    }
    rc = dynamic_cast<const std::domain_error*>(&e) ? 1 : dynamic_cast<const std::invalid_argument*>(&e) ? 2 : 3;
  }
  stan::math::recover_memory();
  return rc;
}

matrix_v to_var(const matrix_d& m) {
  matrix_v v(m.rows(), m.cols());
  for (Eigen::Index i = 0; i < m.size(); ++i) v(i) = var(m(i));
  return v;
}
}  // namespace

extern "C" {

// C (m x n) = A (m x k) * B (k x n), dense column-major doubles
int stanref_multiply(long m, long n, long k, const double* a, const double* b, double* c, double* seconds, char* err,
                     int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, m, k), B = Eigen::Map<const matrix_d>(b, k, n);
        auto t0 = std::chrono::steady_clock::now();
        matrix_d C = stan::math::multiply(A, B);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        Eigen::Map<matrix_d>(c, m, n) = C;
      },
      err, errlen);
}

// Reverse mode: C = multiply(A, B) with A and / or B matrices of var (var_mask bit 0: A, bit 1: B), then the
// gradient of f = sum_ij W_ij C_ij: dA = W B', dB = A' W.  seconds[0]: forward, seconds[1]: grad().
int stanref_multiply_rev(long m, long n, long k, const double* a, const double* b, const double* w, int var_mask, double* c,
                         double* da, double* db, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, m, k), B = Eigen::Map<const matrix_d>(b, k, n);
        Eigen::Map<const matrix_d> W(w, m, n);
        matrix_v Av, Bv, Cv;
        if (var_mask & 1) Av = to_var(A);
        if (var_mask & 2) Bv = to_var(B);
        auto t0 = std::chrono::steady_clock::now();
        if ((var_mask & 3) == 3) Cv = stan::math::multiply(Av, Bv);
        else if (var_mask & 1) Cv = stan::math::multiply(Av, B);
        else if (var_mask & 2) Cv = stan::math::multiply(A, Bv);
        else throw std::invalid_argument("stanref_multiply_rev: no variable operand");
        auto t1 = std::chrono::steady_clock::now();
        var f = 0;
        {
          std::vector<var> ops(Cv.data(), Cv.data() + Cv.size());
          std::vector<double> grads(w, w + m * n);
          double val = 0;
          for (long i = 0; i < m * n; ++i) val += w[i] * Cv(i).val();
          f = stan::math::precomputed_gradients(val, ops, grads);
        }
        auto t2 = std::chrono::steady_clock::now();
        f.grad();
        auto t3 = std::chrono::steady_clock::now();
        if (seconds) {
          seconds[0] = std::chrono::duration<double>(t1 - t0).count();
          seconds[1] = std::chrono::duration<double>(t3 - t2).count();
        }
        for (long i = 0; i < m * n; ++i) c[i] = Cv(i).val();
        if ((var_mask & 1) && da)
          for (long i = 0; i < m * k; ++i) da[i] = Av(i).adj();
        if ((var_mask & 2) && db)
          for (long i = 0; i < k * n; ++i) db[i] = Bv(i).adj();
      },
      err, errlen);
}

// L = cholesky_decompose(A) for a double matrix (prim/mat/fun/cholesky_decompose.hpp:34-62); rc 1 when the
// reference throws std::domain_error (not symmetric / not positive definite)
int stanref_cholesky(long n, const double* a, double* l, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, n, n);
        auto t0 = std::chrono::steady_clock::now();
        matrix_d L = stan::math::cholesky_decompose(A);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        Eigen::Map<matrix_d>(l, n, n) = L;
      },
      err, errlen);
}

// Reverse mode: L = cholesky_decompose(A_v) (rev/mat/fun/cholesky_decompose.hpp:396-440: cholesky_scalar up to
// n = 35, cholesky_block above), f = sum_ij W_ij L_ij over the lower triangle, a_adj = df/dA for the lower triangle
// of A (the operands of the vari; zeros above).  seconds[0]: forward, seconds[1]: grad().
int stanref_cholesky_rev(long n, const double* a, const double* w, double* l, double* a_adj, double* seconds, char* err,
                         int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, n, n);
        matrix_v Av = to_var(A);
        auto t0 = std::chrono::steady_clock::now();
        matrix_v Lv = stan::math::cholesky_decompose(Av);
        auto t1 = std::chrono::steady_clock::now();
        std::vector<var> ops;
        std::vector<double> grads;
        double val = 0;
        for (long j = 0; j < n; ++j)
          for (long i = j; i < n; ++i) {
            ops.push_back(Lv(i, j));
            grads.push_back(w[j * n + i]);
            val += w[j * n + i] * Lv(i, j).val();
          }
        var f = stan::math::precomputed_gradients(val, ops, grads);
        auto t2 = std::chrono::steady_clock::now();
        f.grad();
        auto t3 = std::chrono::steady_clock::now();
        if (seconds) {
          seconds[0] = std::chrono::duration<double>(t1 - t0).count();
          seconds[1] = std::chrono::duration<double>(t3 - t2).count();
        }
        for (long j = 0; j < n; ++j)
          for (long i = 0; i < n; ++i) {
            l[j * n + i] = i >= j ? Lv(i, j).val() : 0.0;
            a_adj[j * n + i] = i >= j ? Av(i, j).adj() : 0.0;
          }
      },
      err, errlen);
}

// X = mdivide_left_tri<Lower | Upper>(A, B) for double matrices (prim/mat/fun/mdivide_left_tri.hpp:58-77)
int stanref_mdivide_left_tri(int lower, long n, long m, const double* a, const double* b, double* x, double* seconds, char* err,
                             int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, n, n), B = Eigen::Map<const matrix_d>(b, n, m);
        auto t0 = std::chrono::steady_clock::now();
        matrix_d X = lower ? stan::math::mdivide_left_tri<Eigen::Lower>(A, B) : stan::math::mdivide_left_tri<Eigen::Upper>(A, B);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        Eigen::Map<matrix_d>(x, n, m) = X;
      },
      err, errlen);
}

// Reverse mode (rev/mat/fun/mdivide_left_tri.hpp: the vv / dv / vd nodes): X = mdivide_left_tri(A, B) with A and / or B
// matrices of var (var_mask bit 0: A, bit 1: B), f = sum(W * X): da (the triangle that is an operand, zeros
// elsewhere) and db.  seconds[0]: forward, seconds[1]: grad().
int stanref_mdivide_left_tri_rev(int lower, int var_mask, long n, long m, const double* a, const double* b, const double* w,
                                 double* x, double* da, double* db, double* seconds, char* err, int errlen) {
  return guarded(
      [&]() {
        matrix_d A = Eigen::Map<const matrix_d>(a, n, n), B = Eigen::Map<const matrix_d>(b, n, m);
        matrix_v Av, Bv, Xv;
        if (var_mask & 1) Av = to_var(A);
        if (var_mask & 2) Bv = to_var(B);
        auto t0 = std::chrono::steady_clock::now();
#define STANREF_TRI(UPLO)                                                                          \
  ((var_mask & 3) == 3 ? stan::math::mdivide_left_tri<UPLO>(Av, Bv)                                \
                       : (var_mask & 1) ? stan::math::mdivide_left_tri<UPLO>(Av, B) : stan::math::mdivide_left_tri<UPLO>(A, Bv))
        if (!(var_mask & 3)) throw std::invalid_argument("stanref_mdivide_left_tri_rev: no variable operand");
        if (lower) Xv = STANREF_TRI(Eigen::Lower); else Xv = STANREF_TRI(Eigen::Upper);
#undef STANREF_TRI
        auto t1 = std::chrono::steady_clock::now();
        std::vector<var> ops(Xv.data(), Xv.data() + Xv.size());
        std::vector<double> grads(w, w + n * m);
        double val = 0;
        for (long i = 0; i < n * m; ++i) val += w[i] * Xv(i).val();
        var f = stan::math::precomputed_gradients(val, ops, grads);
        auto t2 = std::chrono::steady_clock::now();
        f.grad();
        auto t3 = std::chrono::steady_clock::now();
        if (seconds) {
          seconds[0] = std::chrono::duration<double>(t1 - t0).count();
          seconds[1] = std::chrono::duration<double>(t3 - t2).count();
        }
        for (long i = 0; i < n * m; ++i) x[i] = Xv(i).val();
        if ((var_mask & 1) && da)
          for (long i = 0; i < n * n; ++i) da[i] = Av(i).adj();
        if ((var_mask & 2) && db)
          for (long i = 0; i < n * m; ++i) db[i] = Bv(i).adj();
      },
      err, errlen);
}

}  // extern "C"
