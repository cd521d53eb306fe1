#ifndef STAN_MATH_CUDA_REV_CHOLESKY_DECOMPOSE_HPP
#define STAN_MATH_CUDA_REV_CHOLESKY_DECOMPOSE_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/cholesky_decompose.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>

namespace stan {
namespace math {
namespace cuda {

/**
 * The tape node of L = cholesky_decompose(A) whose reverse sweep runs on the device: the counterpart of the
 * reference's cholesky_opencl (stan/math/rev/mat/fun/cholesky_decompose.hpp:255-383).  Same storage as
 * cholesky_block / cholesky_opencl -- the varis of the lower triangles of A and L, packed column by column
 * (:273-292) -- so the reference's set_lower_tri_coeff_ref builds the returned matrix from it.  chain() gathers L
 * and its adjoints into dense matrices, calls stan_cuda_cholesky_adjoint (the blocked algorithm of :329-382 with the
 * products on the DMMA kernel) and scatters the result into the adjoints of A (:378-381).
 */
class cholesky_cuda : public vari {
 public:
  int M_;
  vari** vari_ref_A_;
  vari** vari_ref_L_;

  cholesky_cuda(const Eigen::Matrix<var, -1, -1>& A, const Eigen::Matrix<double, -1, -1>& L_A)
      : vari(0.0),
        M_(A.rows()),
        vari_ref_A_(ChainableStack::instance_->memalloc_.alloc_array<vari*>(A.rows() * (A.rows() + 1) / 2)),
        vari_ref_L_(ChainableStack::instance_->memalloc_.alloc_array<vari*>(A.rows() * (A.rows() + 1) / 2)) {
    size_t pos = 0;
    for (Eigen::Index j = 0; j < M_; ++j) {
      for (Eigen::Index i = j; i < M_; ++i) {
        vari_ref_A_[pos] = A.coeffRef(i, j).vi_;
        vari_ref_L_[pos] = new vari(L_A.coeffRef(i, j), false);
        ++pos;
      }
    }
  }

  virtual void chain() {
    Eigen::MatrixXd L = Eigen::MatrixXd::Zero(M_, M_), L_adj = Eigen::MatrixXd::Zero(M_, M_), A_adj(M_, M_);
    size_t pos = 0;
    for (Eigen::Index j = 0; j < M_; ++j) {
      for (Eigen::Index i = j; i < M_; ++i) {
        L_adj.coeffRef(i, j) = vari_ref_L_[pos]->adj_;
        L.coeffRef(i, j) = vari_ref_L_[pos]->val_;
        ++pos;
      }
    }
    check_cuda_rc("cholesky_decompose", stan_cuda_cholesky_adjoint(M_, L.data(), M_, L_adj.data(), M_, A_adj.data(), M_));
    pos = 0;
    for (Eigen::Index j = 0; j < M_; ++j) {
      for (Eigen::Index i = j; i < M_; ++i) {
        vari_ref_A_[pos++]->adj_ += A_adj.coeffRef(i, j);
      }
    }
  }
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
