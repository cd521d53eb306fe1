#ifndef STAN_MATH_CUDA_REV_MULTIPLY_HPP
#define STAN_MATH_CUDA_REV_MULTIPLY_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/multiply.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/rev/mat/fun/typedefs.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <type_traits>

namespace stan {
namespace math {
namespace cuda {

/**
 * The operands of one product, resident on the device between the forward sweep and chain().  The reference's
 * OpenCL branch uploads A and B again in chain() (rev/mat/fun/multiply.hpp:114-116); here they stay where the
 * forward product left them.  A chainable_alloc: recover_memory() / recover_memory_nested() destroys it together
 * with the tape that refers to it (rev/core/chainable_alloc.hpp, rev/core/recover_memory.hpp).
 */
struct resident_operands : public chainable_alloc {
  matrix_cuda A, B;
  virtual ~resident_operands() {}
};

namespace detail {
template <typename T>
struct is_var_scalar : std::is_same<T, var> {};

/** values of a matrix of double (borrowed) or var (gathered) as column-major doubles */
template <int R, int C>
inline const double* values_of(const Eigen::Matrix<double, R, C>& m, Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& /*tmp*/) {
  return m.data();
}
template <int R, int C>
inline const double* values_of(const Eigen::Matrix<var, R, C>& m, Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic>& tmp) {
  tmp.resize(m.rows(), m.cols());
  const var* src = m.data();
  double* dst = tmp.data();
  for (Eigen::Index i = 0; i < m.size(); ++i) {
    dst[i] = src[i].vi_->val_;
  }
  return tmp.data();
}
template <int R, int C>
inline void varis_of(const Eigen::Matrix<double, R, C>& /*m*/, vari** /*dst*/) {}
template <int R, int C>
inline void varis_of(const Eigen::Matrix<var, R, C>& m, vari** dst) {
  const var* src = m.data();
  for (Eigen::Index i = 0; i < m.size(); ++i) {
    dst[i] = src[i].vi_;
  }
}
}  // namespace detail

/**
 * multiply_mat_vari (stan/math/rev/mat/fun/multiply.hpp:35-150 var x var, :225-328 double x var, :395-501
 * var x double) with its three products on the device: AB in the constructor (:88-99), adjAB * B' and A' * adjAB in
 * chain() (:111-123).  Same tape layout: one node on the chaining stack that owns the element varis of AB
 * (`new vari(x, false)`, :95-96).  The host copies Ad_ / Bd_ of the reference are not kept: the device copies
 * serve chain().
 */
template <typename Ta, int Ra, int Ca, typename Tb, int Cb>
class multiply_mat_vari_cuda : public vari {
 public:
  int A_rows_;
  int A_cols_;
  int B_cols_;
  vari** variRefA_;
  vari** variRefB_;
  vari** variRefAB_;
  resident_operands* dev_;

  multiply_mat_vari_cuda(const Eigen::Matrix<Ta, Ra, Ca>& A, const Eigen::Matrix<Tb, Ca, Cb>& B)
      : vari(0.0),
        A_rows_(A.rows()),
        A_cols_(A.cols()),
        B_cols_(B.cols()),
        variRefA_(detail::is_var_scalar<Ta>::value ? ChainableStack::instance_->memalloc_.alloc_array<vari*>(A.size()) : nullptr),
        variRefB_(detail::is_var_scalar<Tb>::value ? ChainableStack::instance_->memalloc_.alloc_array<vari*>(B.size()) : nullptr),
        variRefAB_(ChainableStack::instance_->memalloc_.alloc_array<vari*>(A.rows() * B.cols())),
        dev_(new resident_operands()) {
    detail::varis_of(A, variRefA_);
    detail::varis_of(B, variRefB_);
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> tmp;
    dev_->A = matrix_cuda(detail::values_of(A, tmp), A.rows(), A.cols());
    dev_->B = matrix_cuda(detail::values_of(B, tmp), B.rows(), B.cols());
    matrix_cuda AB = dev_->A * dev_->B;
    tmp.resize(A_rows_, B_cols_);
    AB.download(tmp.data());
    const double* v = tmp.data();
    const Eigen::Index n = tmp.size();
    for (Eigen::Index i = 0; i < n; ++i) {
      variRefAB_[i] = new vari(v[i], false);
    }
  }

  virtual void chain() {
    const Eigen::Index n_ab = static_cast<Eigen::Index>(A_rows_) * B_cols_;
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> adjAB(A_rows_, B_cols_);
    double* g = adjAB.data();
    for (Eigen::Index i = 0; i < n_ab; ++i) {
      g[i] = variRefAB_[i]->adj_;
    }
    matrix_cuda adjAB_dev(adjAB);
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> tmp;
    if (detail::is_var_scalar<Ta>::value) {  // adjA += adjAB * B'  (:117, :128-129)
      matrix_cuda adjA = adjAB_dev * transpose(dev_->B);
      tmp.resize(A_rows_, A_cols_);
      adjA.download(tmp.data());
      const double* t = tmp.data();
      const Eigen::Index n = tmp.size();
      for (Eigen::Index i = 0; i < n; ++i) {
        variRefA_[i]->adj_ += t[i];
      }
    }
    if (detail::is_var_scalar<Tb>::value) {  // adjB += A' * adjAB  (:118, :130-131)
      matrix_cuda adjB = transpose(dev_->A) * adjAB_dev;
      tmp.resize(A_cols_, B_cols_);
      adjB.download(tmp.data());
      const double* t = tmp.data();
      const Eigen::Index n = tmp.size();
      for (Eigen::Index i = 0; i < n; ++i) {
        variRefB_[i]->adj_ += t[i];
      }
    }
  }
};

/** GPU body of multiply(A, B) with at least one matrix of var (stan/math/rev/mat/fun/multiply.hpp:619-633) */
template <typename Ta, int Ra, int Ca, typename Tb, int Cb>
inline Eigen::Matrix<var, Ra, Cb> multiply_device_rev(const Eigen::Matrix<Ta, Ra, Ca>& A, const Eigen::Matrix<Tb, Ca, Cb>& B) {
  multiply_mat_vari_cuda<Ta, Ra, Ca, Tb, Cb>* baseVari = new multiply_mat_vari_cuda<Ta, Ra, Ca, Tb, Cb>(A, B);
  Eigen::Matrix<var, Ra, Cb> AB_v(A.rows(), B.cols());
  var* out = AB_v.data();
  const Eigen::Index n = AB_v.size();
  for (Eigen::Index i = 0; i < n; ++i) {
    out[i].vi_ = baseVari->variRefAB_[i];
  }
  return AB_v;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
