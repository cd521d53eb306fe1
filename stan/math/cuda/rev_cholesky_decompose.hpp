#ifndef STAN_MATH_CUDA_REV_CHOLESKY_DECOMPOSE_HPP
#define STAN_MATH_CUDA_REV_CHOLESKY_DECOMPOSE_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/cholesky_decompose.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <vector>

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
    // three n x n work matrices, kept per thread across evaluations: a fresh Eigen matrix of this size is 32 MB of
    // page faults at n = 2048, every evaluation, and only the lower triangles are read or written on either side
    const size_t nn = static_cast<size_t>(M_) * M_;
    static thread_local std::vector<double> work;
    if (work.size() < 3 * nn) {
      work.assign(3 * nn, 0.0);
    }
    double* L = work.data();
    double* L_adj = L + nn;
    double* A_adj = L_adj + nn;
    size_t pos = 0;
    for (Eigen::Index j = 0; j < M_; ++j) {
      for (Eigen::Index i = j; i < M_; ++i) {
        L_adj[j * M_ + i] = vari_ref_L_[pos]->adj_;
        L[j * M_ + i] = vari_ref_L_[pos]->val_;
        ++pos;
      }
    }
    check_cuda_rc("cholesky_decompose", stan_cuda_cholesky_adjoint(M_, L, M_, L_adj, M_, A_adj, M_));
    pos = 0;
    for (Eigen::Index j = 0; j < M_; ++j) {
      for (Eigen::Index i = j; i < M_; ++i) {
        vari_ref_A_[pos++]->adj_ += A_adj[j * M_ + i];
      }
    }
  }
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
