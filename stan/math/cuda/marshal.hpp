#ifndef STAN_MATH_CUDA_MARSHAL_HPP
#define STAN_MATH_CUDA_MARSHAL_HPP
#ifdef STAN_CUDA

// Compile-time routing and marshalling between the reference's argument kinds and the C ABI.
// C++14 only (the reference builds with -std=c++1y): overloads and tag types, no if constexpr.

#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/meta.hpp>
#include <stan/math/prim/scal/fun/value_of_rec.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/rev/meta.hpp>
#ifdef STAN_THREADS
#include <tbb/blocked_range.h>
#include <tbb/parallel_for.h>
#include <tbb/task_arena.h>
#endif
#include <cstdlib>
#include <memory>
#include <type_traits>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/**
 * body(begin, end) over [0, n): in parallel over the TBB workers for the N-element operand passes of the
 * elementwise densities, serially below 65536 elements or without STAN_THREADS.
 *
 * The parallel loop runs ISOLATED (tbb::this_task_arena::isolate).  These passes are nested parallelism whenever the
 * density is evaluated inside a map_rect job: a worker that waits in a plain nested parallel_for may be handed
 * ANOTHER map_rect job, which would evaluate on the same thread -- overwriting this thread's staging buffer
 * (stan_cuda_host_staging is per thread) and leaving the shard scope of the outer job closed.  Isolation restricts
 * the waiting thread to tasks of this loop (test: map_rect_dispatch_mock_test.cpp, nested_parallel_gather_...).
 */
/** How many workers share one of these passes (default 2, STAN_CUDA_HOST_PARTS).  Measured on the B200 hosts at
 * N = 1e6 (profiles/r02_host_thread_sweep.txt): 2-3 workers are the optimum -- C1 6.2-6.4 ms against 6.8 ms serial,
 * 6.9-7.3 ms with 4-6 and 9.3 ms with all 16 cores.  The passes stream 8-24 MB that ONE thread then walks again (the D2H copy into the
 * partials, grad() over the tape, the caller reading the adjoints): spread over every core's private cache the
 * data has to travel back line by line (device call 0.75 -> 1.56 ms, grad() 2.9 -> 5.2 ms with 16 workers). */
inline size_t host_pass_parts() {
  static const size_t parts = [] {
    const char* v = std::getenv("STAN_CUDA_HOST_PARTS");
    const long p = v ? std::atol(v) : 2;
    return static_cast<size_t>(p < 1 ? 1 : (p > 64 ? 64 : p));
  }();
  return parts;
}

template <typename Body>
inline void for_each_chunk(size_t n, const Body& body) {
#ifdef STAN_THREADS
  const size_t parts = host_pass_parts();
  if (n >= (size_t(1) << 16) && parts > 1) {
    const size_t per = (n + parts - 1) / parts;
    tbb::this_task_arena::isolate([&] {
      tbb::parallel_for(size_t(0), parts, [&](size_t q) {
        const size_t begin = q * per, end = begin + per < n ? begin + per : n;
        if (begin < end) {
          body(begin, end);
        }
      });
    });
    return;
  }
#endif
  body(size_t(0), n);
}

// ---- which instantiations take the GPU path ---------------------------------------
template <typename T>
struct is_double_or_var
    : std::integral_constant<bool, std::is_same<T, double>::value || std::is_same<T, var>::value> {};

/** x: an Eigen matrix of doubles (data).  var / fvar matrices stay on the reference body. */
template <typename T>
struct is_data_matrix : std::false_type {};
template <int R, int C, int O, int MR, int MC>
struct is_data_matrix<Eigen::Matrix<double, R, C, O, MR, MC>> : std::true_type {};

/**
 * Zero-copy view of column-major double data that already lives somewhere else, e.g. the x_r
 * array a map_rect functor receives:  data_matrix_view x(x_r.data(), rows, cols);
 * The reference's GLMs only accept an owning Eigen::Matrix (value_of_rec has no Map overload,
 * stan/math/prim/mat/fun/value_of_rec.hpp:22-49), so a functor has to copy x_r into a matrix on
 * every call; the device path never reads x on the host after the first call, so it also takes a
 * view (an extension: calls with a view exist only under STAN_CUDA).
 */
typedef Eigen::Map<const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>> data_matrix_view;
template <>
struct is_data_matrix<data_matrix_view> : std::true_type {};

/** x as a parameter: a column-major Eigen matrix of var.  Its values are gathered into a
 * temporary on every call and travel to the device with it (stan_cuda_glm_args.x_per_call); the N x K partials
 * d * beta^T come back into the x edge of the tape node. */
template <typename T>
struct is_var_matrix : std::false_type {};
template <int O, int MR, int MC>
struct is_var_matrix<Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic, O, MR, MC>>
    : std::integral_constant<bool, (O & Eigen::RowMajor) == 0> {};
template <typename T>
struct is_design_matrix : std::integral_constant<bool, is_data_matrix<T>::value || is_var_matrix<T>::value> {};

/** What the x argument contributes to the tape node: nothing (declared on a double, so that views of data work
 * too) for data, the matrix itself for a var matrix. */
template <typename T_x>
struct x_operand {
  typedef double type;
  typedef double scalar;
  static double get(const T_x& /*x*/) { return 0.0; }
};
template <int O, int MR, int MC>
struct x_operand<Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic, O, MR, MC>> {
  typedef Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic, O, MR, MC> type;
  typedef var scalar;
  static const type& get(const type& x) { return x; }
};

/** scalar, std::vector or Eigen vector whose scalar type satisfies Pred */
template <typename T, template <typename> class Pred>
struct is_scalar_or_vector_of : Pred<T> {};
template <typename T, template <typename> class Pred>
struct is_scalar_or_vector_of<std::vector<T>, Pred> : Pred<T> {};
template <typename T, int R, int C, int O, int MR, int MC, template <typename> class Pred>
struct is_scalar_or_vector_of<Eigen::Matrix<T, R, C, O, MR, MC>, Pred>
    : std::integral_constant<bool, Pred<T>::value && (R == 1 || C == 1)> {};

template <typename T>
struct is_int_scalar : std::is_same<T, int> {};
template <typename T>
struct is_double_int_or_var
    : std::integral_constant<bool, is_double_or_var<T>::value || std::is_same<T, int>::value> {};

/** parameters (alpha, beta, sigma, mu): double / int / var, scalar or vector */
template <typename T>
struct is_param_arg : is_scalar_or_vector_of<T, is_double_int_or_var> {};
/** integer outcomes (bernoulli, poisson): int scalar / std::vector<int> / Eigen int vector */
template <typename T>
struct is_int_outcome : is_scalar_or_vector_of<T, is_int_scalar> {};
/** real outcomes (normal): double or var, scalar or vector */
template <typename T>
struct is_real_outcome : is_scalar_or_vector_of<T, is_double_or_var> {};

template <typename T_y, typename T_x, typename T_alpha, typename T_beta>
struct int_glm_eligible
    : std::integral_constant<bool, is_int_outcome<T_y>::value && is_design_matrix<T_x>::value
                                       && is_param_arg<T_alpha>::value && is_param_arg<T_beta>::value> {};

template <typename T_y, typename T_x, typename T_alpha, typename T_beta, typename T_scale>
struct normal_glm_eligible
    : std::integral_constant<bool, is_real_outcome<T_y>::value && is_design_matrix<T_x>::value
                                       && is_param_arg<T_alpha>::value && is_param_arg<T_beta>::value
                                       && is_param_arg<T_scale>::value> {};

// ---- element access without a copy ----------------------------------------------------------------
/**
 * i-th scalar of a scalar / std::vector / Eigen vector argument, by reference.  The reference's
 * scalar_seq_view holds the container BY VALUE (prim/scal/meta/scalar_seq_view.hpp:36: `C c_;`), i.e. it copies
 * an N-element var vector (8 MB at N = 1e6, plus the page faults of a fresh allocation) before the first element
 * is read; the device path reads each element exactly once, so it indexes the caller's container directly.
 */
template <typename T>
struct seq_access {
  static const T& at(const T& v, size_t /*i*/) { return v; }
};
template <typename S>
struct seq_access<std::vector<S>> {
  static const S& at(const std::vector<S>& v, size_t i) { return v[i]; }
};
template <typename S, int R, int C, int O, int MR, int MC>
struct seq_access<Eigen::Matrix<S, R, C, O, MR, MC>> {
  static const S& at(const Eigen::Matrix<S, R, C, O, MR, MC>& v, size_t i) { return v.data()[i]; }
};

// ---- values of an argument as contiguous doubles ------------------------------------------------
/**
 * View of the values of a scalar / std::vector / Eigen vector argument as `size()` contiguous
 * doubles.  Double containers are borrowed (no copy, stable address -> usable as a cache
 * identity); everything else is copied with value_of_rec.
 */
template <typename T>
class values {
 public:
  explicit values(const T& v) : owned_(new double[length(v) > 0 ? length(v) : 1]), size_(length(v)) {
    double* dst = owned_.get();  // uninitialised storage: every element is written by the gather below
    for_each_chunk(size_, [&](size_t begin, size_t end) {  // N-element intercepts / scales: in parallel
      for (size_t i = begin; i != end; ++i) {
        dst[i] = value_of_rec(seq_access<T>::at(v, i));
      }
    });
  }
  const double* data() const { return owned_.get(); }
  size_t size() const { return size_; }
  bool borrowed() const { return false; }

 private:
  std::unique_ptr<double[]> owned_;
  size_t size_;
};

template <>
class values<std::vector<double>> {
 public:
  explicit values(const std::vector<double>& v) : ptr_(v.data()), size_(v.size()) {}
  const double* data() const { return ptr_; }
  size_t size() const { return size_; }
  bool borrowed() const { return true; }

 private:
  const double* ptr_;
  size_t size_;
};

template <int R, int C, int O, int MR, int MC>
class values<Eigen::Matrix<double, R, C, O, MR, MC>> {
 public:
  explicit values(const Eigen::Matrix<double, R, C, O, MR, MC>& v) : ptr_(v.data()), size_(v.size()) {}
  const double* data() const { return ptr_; }
  size_t size() const { return size_; }
  bool borrowed() const { return true; }

 private:
  const double* ptr_;
  size_t size_;
};

/** Outcome vector of exactly n doubles: a scalar outcome is broadcast (the reference's
 * as_array_or_scalar does the same implicitly). */
template <typename T>
class real_outcome {
 public:
  real_outcome(const T& y, size_t n) : vals_(y) {
    if (vals_.size() == n) {
      ptr_ = vals_.data();
    } else {
      expanded_.assign(n, vals_.size() ? vals_.data()[0] : 0.0);
      ptr_ = expanded_.data();
    }
  }
  const double* data() const { return ptr_; }
  /** true when the doubles live in a temporary of this call (a var outcome vector, a broadcast scalar): their
   * address is no identity and their content may change from call to call, so the device cache is keyed on x
   * alone and y travels with every call (stan_cuda_glm_args.y_per_call) */
  bool per_call() const { return !vals_.borrowed() || ptr_ != vals_.data(); }

 private:
  values<T> vals_;
  std::vector<double> expanded_;
  const double* ptr_;
};

/** Integer outcome vector of exactly n int32: borrowed when the caller holds ints already. */
template <typename T>
class int_outcome;

template <>
class int_outcome<int> {
 public:
  int_outcome(const int& y, size_t n) : owned_(n, y) {}
  const int* data() const { return owned_.data(); }
  bool per_call() const { return true; }  // a temporary: see real_outcome::per_call

 private:
  std::vector<int> owned_;
};

template <>
class int_outcome<std::vector<int>> {
 public:
  int_outcome(const std::vector<int>& y, size_t /*n*/) : ptr_(y.data()) {}
  const int* data() const { return ptr_; }
  bool per_call() const { return false; }

 private:
  const int* ptr_;
};

template <int R, int C, int O, int MR, int MC>
class int_outcome<Eigen::Matrix<int, R, C, O, MR, MC>> {
 public:
  int_outcome(const Eigen::Matrix<int, R, C, O, MR, MC>& y, size_t /*n*/) : ptr_(y.data()) {}
  const int* data() const { return ptr_; }
  bool per_call() const { return false; }

 private:
  const int* ptr_;
};

// ---- where the partials of an operands_and_partials edge live --------------------------------
// Constant operands have an empty edge (prim/scal/meta/operands_and_partials.hpp:35-51): nullptr.
template <typename ViewElt, typename Op>
inline double* partials_data(internal::ops_partials_edge<ViewElt, Op>& /*edge*/) {
  return nullptr;
}
inline double* partials_data(internal::ops_partials_edge<double, var>& edge) { return &edge.partial_; }
inline double* partials_data(internal::ops_partials_edge<double, std::vector<var>>& edge) {
  return edge.partials_.data();
}
template <int R, int C>
inline double* partials_data(internal::ops_partials_edge<double, Eigen::Matrix<var, R, C>>& edge) {
  return edge.partials_.data();
}

/** The design matrix as doubles: data is borrowed, a var matrix is gathered (in parallel) into a temporary. */
template <typename T_x>
class design_values {
 public:
  explicit design_values(const T_x& x) : x_(x) {}
  const T_x& matrix() const { return x_; }
  bool per_call() const { return false; }

 private:
  const T_x& x_;
};
template <int O, int MR, int MC>
class design_values<Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic, O, MR, MC>> {
 public:
  typedef Eigen::Matrix<var, Eigen::Dynamic, Eigen::Dynamic, O, MR, MC> T_x;
  explicit design_values(const T_x& x) : vals_(x.rows(), x.cols()) {
    const var* src = x.data();  // column-major (is_var_matrix): same order as the values and as the x edge
    double* dst = vals_.data();
    for_each_chunk(static_cast<size_t>(x.size()), [&](size_t begin, size_t end) {
      for (size_t i = begin; i != end; ++i) {
        dst[i] = src[i].vi_->val_;
      }
    });
  }
  const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& matrix() const { return vals_; }
  bool per_call() const { return true; }

 private:
  Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> vals_;
};

/** Matrix layout as the C ABI wants it. */
template <int R, int C, int O, int MR, int MC>
inline void matrix_layout(const Eigen::Matrix<double, R, C, O, MR, MC>& x, const double** data, int64_t* ld,
                          int32_t* row_major) {
  *data = x.data();
  *row_major = (O & Eigen::RowMajor) ? 1 : 0;
  *ld = *row_major ? x.cols() : x.rows();
  if (*ld < 1) {
    *ld = 1;
  }
}

inline void matrix_layout(const data_matrix_view& x, const double** data, int64_t* ld, int32_t* row_major) {
  *data = x.data();
  *row_major = 0;
  *ld = x.rows() < 1 ? 1 : x.rows();
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
