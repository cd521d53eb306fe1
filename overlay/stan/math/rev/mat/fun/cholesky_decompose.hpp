// Overlay of stan/math/rev/mat/fun/cholesky_decompose.hpp (see the prim overlay of the same name).
// The reference's reverse-mode function is kept as cholesky_decompose_reference_cpu; with -DSTAN_CUDA a matrix of var
// of at least cholesky_size_worth_transfer rows gets its factor from the device and a cholesky_cuda node
// (stan/math/cuda/rev_cholesky_decompose.hpp) whose reverse sweep runs there too -- the two places where the
// reference switches to its OpenCL backend (:402 and :427-431).
#ifndef STAN_MATH_CUDA_OVERLAY_REV_CHOLESKY_DECOMPOSE_HPP
#define STAN_MATH_CUDA_OVERLAY_REV_CHOLESKY_DECOMPOSE_HPP

// everything the reference header includes, first (see overlay/stan/math/rev/mat/fun/multiply.hpp)
#include <stan/math/rev/meta.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/mat/fun/typedefs.hpp>
#include <stan/math/prim/mat/fun/cholesky_decompose.hpp>
#include <stan/math/rev/scal/fun/value_of_rec.hpp>
#include <stan/math/rev/scal/fun/value_of.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/prim/mat/fun/value_of_rec.hpp>
#include <stan/math/prim/mat/err/check_pos_definite.hpp>
#include <stan/math/prim/mat/err/check_square.hpp>
#include <stan/math/prim/mat/err/check_symmetric.hpp>
#include <algorithm>
#include <vector>

#define cholesky_decompose cholesky_decompose_reference_cpu
#include_next <stan/math/rev/mat/fun/cholesky_decompose.hpp>
#undef cholesky_decompose

#ifdef STAN_CUDA
#include <stan/math/cuda/rev_cholesky_decompose.hpp>
#endif

namespace stan {
namespace math {

inline Eigen::Matrix<var, -1, -1> cholesky_decompose(const Eigen::Matrix<var, -1, -1>& A) {
#ifdef STAN_CUDA
  if (A.rows() >= cuda::cholesky_size_worth_transfer() && A.rows() == A.cols()) {
    Eigen::Matrix<double, -1, -1> A_val(value_of_rec(A));
    Eigen::Matrix<double, -1, -1> L_A;
    if (cuda::cholesky_on_device(A_val, L_A) == STAN_CUDA_OK) {  // check_symmetric (:404) runs on the device copy
      cuda::cholesky_cuda* baseVari = new cuda::cholesky_cuda(A, L_A);
      Eigen::Matrix<var, -1, -1> L(A.rows(), A.cols());
      internal::set_lower_tri_coeff_ref(L, baseVari->vari_ref_L_);  // :428-429
      return L;
    }
    // not symmetric or not positive definite: the reference's own checks, in its order, throw its exception
  }
#endif
  return cholesky_decompose_reference_cpu(A);
}

}  // namespace math
}  // namespace stan
#endif
