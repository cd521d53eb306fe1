#ifndef STAN_MATH_CUDA_REV_MDIVIDE_LEFT_TRI_HPP
#define STAN_MATH_CUDA_REV_MDIVIDE_LEFT_TRI_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/mdivide_left_tri.hpp>
#include <stan/math/cuda/multiply.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <type_traits>

namespace stan {
namespace math {
namespace cuda {

/**
 * The three tape nodes of mdivide_left_tri (stan/math/rev/mat/fun/mdivide_left_tri.hpp: _vv_ :18-130, _dv_ :132-203,
 * _vd_ :205-290) as one class with the solves and the product on the device: C = A^-1 B in the constructor (:69-82),
 * adjB = A^-T adjC and adjA = -adjB C' in chain() (:91-105).  Storage as in the reference: the values of A and C in
 * the arena, the varis of the triangle of A column by column (:45-62), of B and of C.
 */
template <int TriView, typename Ta, int R1, int C1, typename Tb, int R2, int C2>
class mdivide_left_tri_vari_cuda : public vari {
 public:
  static constexpr bool a_is_var = std::is_same<Ta, var>::value;
  static constexpr bool b_is_var = std::is_same<Tb, var>::value;
  int M_;  // A.rows() = A.cols() = B.rows()
  int N_;  // B.cols()
  double* A_;
  double* C_;
  vari** variRefA_;
  vari** variRefB_;
  vari** variRefC_;

  mdivide_left_tri_vari_cuda(const Eigen::Matrix<Ta, R1, C1>& A, const Eigen::Matrix<Tb, R2, C2>& B)
      : vari(0.0),
        M_(A.rows()),
        N_(B.cols()),
        A_(ChainableStack::instance_->memalloc_.alloc_array<double>(A.rows() * A.cols())),
        C_(ChainableStack::instance_->memalloc_.alloc_array<double>(B.rows() * B.cols())),
        variRefA_(a_is_var ? ChainableStack::instance_->memalloc_.alloc_array<vari*>(A.rows() * (A.rows() + 1) / 2) : nullptr),
        variRefB_(b_is_var ? ChainableStack::instance_->memalloc_.alloc_array<vari*>(B.rows() * B.cols()) : nullptr),
        variRefC_(ChainableStack::instance_->memalloc_.alloc_array<vari*>(B.rows() * B.cols())) {
    gather_a(A, std::integral_constant<bool, a_is_var>());
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> Bd(M_, N_);
    gather_b(B, Bd.data(), std::integral_constant<bool, b_is_var>());
    tri_solve_on_device(TriView == Eigen::Lower, false, M_, N_, A_, Bd.data(), C_);
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * N_;
    for (Eigen::Index i = 0; i < n; ++i) {
      variRefC_[i] = new vari(C_[i], false);
    }
  }

  virtual void chain() {
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * N_;
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> adjC(M_, N_), adjB(M_, N_);
    for (Eigen::Index i = 0; i < n; ++i) {
      adjC.data()[i] = variRefC_[i]->adj_;
    }
    tri_solve_on_device(TriView == Eigen::Lower, true, M_, N_, A_, adjC.data(), adjB.data());  // adjB = A^-T adjC
    if (a_is_var) {
      Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> adjA(M_, M_);
      if (multiply_worth_transfer(M_, N_, M_)) {  // adjA = -adjB C'
        multiply_on_device(M_, M_, N_, adjB.data(), M_, false, C_, M_, true, adjA.data(), M_);
        adjA = -adjA;
      } else {
        adjA.noalias() = -adjB * Eigen::Map<Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> >(C_, M_, N_).transpose();
      }
      size_t pos = 0;
      if (TriView == Eigen::Lower) {
        for (Eigen::Index j = 0; j < M_; ++j) {
          for (Eigen::Index i = j; i < M_; ++i) {
            variRefA_[pos++]->adj_ += adjA(i, j);
          }
        }
      } else {
        for (Eigen::Index j = 0; j < M_; ++j) {
          for (Eigen::Index i = 0; i < j + 1; ++i) {
            variRefA_[pos++]->adj_ += adjA(i, j);
          }
        }
      }
    }
    if (b_is_var) {
      for (Eigen::Index i = 0; i < n; ++i) {
        variRefB_[i]->adj_ += adjB.data()[i];
      }
    }
  }

 private:
  void gather_a(const Eigen::Matrix<Ta, R1, C1>& A, std::true_type) {
    size_t pos = 0;
    if (TriView == Eigen::Lower) {
      for (Eigen::Index j = 0; j < M_; ++j) {
        for (Eigen::Index i = j; i < M_; ++i) {
          variRefA_[pos++] = A(i, j).vi_;
        }
      }
    } else {
      for (Eigen::Index j = 0; j < M_; ++j) {
        for (Eigen::Index i = 0; i < j + 1; ++i) {
          variRefA_[pos++] = A(i, j).vi_;
        }
      }
    }
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * M_;
    for (Eigen::Index i = 0; i < n; ++i) {
      A_[i] = A.data()[i].vi_->val_;
    }
  }
  void gather_a(const Eigen::Matrix<Ta, R1, C1>& A, std::false_type) {
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * M_;
    for (Eigen::Index i = 0; i < n; ++i) {
      A_[i] = A.data()[i];
    }
  }
  void gather_b(const Eigen::Matrix<Tb, R2, C2>& B, double* dst, std::true_type) {
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * N_;
    for (Eigen::Index i = 0; i < n; ++i) {
      variRefB_[i] = B.data()[i].vi_;
      dst[i] = B.data()[i].vi_->val_;
    }
  }
  void gather_b(const Eigen::Matrix<Tb, R2, C2>& B, double* dst, std::false_type) {
    const Eigen::Index n = static_cast<Eigen::Index>(M_) * N_;
    for (Eigen::Index i = 0; i < n; ++i) {
      dst[i] = B.data()[i];
    }
  }
};

template <int TriView, typename Ta, int R1, int C1, typename Tb, int R2, int C2>
inline Eigen::Matrix<var, R1, C2> mdivide_left_tri_device_rev(const Eigen::Matrix<Ta, R1, C1>& A,
                                                              const Eigen::Matrix<Tb, R2, C2>& b) {
  mdivide_left_tri_vari_cuda<TriView, Ta, R1, C1, Tb, R2, C2>* baseVari
      = new mdivide_left_tri_vari_cuda<TriView, Ta, R1, C1, Tb, R2, C2>(A, b);
  Eigen::Matrix<var, R1, C2> res(b.rows(), b.cols());
  const Eigen::Index n = res.size();
  for (Eigen::Index i = 0; i < n; ++i) {
    res.data()[i].vi_ = baseVari->variRefC_[i];
  }
  return res;
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
