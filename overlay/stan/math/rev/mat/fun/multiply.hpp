// Overlay of stan/math/rev/mat/fun/multiply.hpp (see overlay/stan/math/prim/mat/fun/multiply.hpp).
// The reference's reverse-mode products keep their overloads under the name multiply_reference_cpu; the public
// multiply() of the prim overlay finds them by argument-dependent lookup.  With -DSTAN_CUDA a matrix x matrix
// product with at least one matrix of var whose dimensions pass the transfer rule is built from
// multiply_mat_vari_cuda (stan/math/cuda/rev_multiply.hpp) instead of multiply_mat_vari: the places where the
// reference itself switches to its OpenCL backend (:88-99, :111-123, :274-303, :444-473).
#ifndef STAN_MATH_CUDA_OVERLAY_REV_MULTIPLY_HPP
#define STAN_MATH_CUDA_OVERLAY_REV_MULTIPLY_HPP

// everything the reference header includes, first: a header pulled in for the first time while the rename below is
// active would have its own `multiply` tokens renamed (or, for the prim overlay, would end the rename early)
#include <stan/math/rev/meta.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/rev/mat/fun/to_var.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/prim/mat/err/check_multiplicable.hpp>
#include <stan/math/prim/scal/err/check_not_nan.hpp>
#include <stan/math/prim/mat.hpp>
#include <stan/math/prim/meta.hpp>
#include <boost/math/tools/promotion.hpp>
#include <type_traits>

#define multiply multiply_reference_cpu
#include_next <stan/math/rev/mat/fun/multiply.hpp>
#undef multiply

#ifdef STAN_CUDA
#include <stan/math/cuda/rev_multiply.hpp>

namespace stan {
namespace math {
namespace cuda_dispatch {

/** matrix x matrix with at least one var scalar type, not the row x column dot product (:647-660) */
template <typename Ta, int Ra, int Ca, typename Tb, int Rb, int Cb>
struct multiply_route<Eigen::Matrix<Ta, Ra, Ca>, Eigen::Matrix<Tb, Rb, Cb>,
                      typename std::enable_if<(std::is_same<Ta, var>::value || std::is_same<Tb, var>::value)
                                              && (std::is_same<Ta, var>::value || std::is_same<Ta, double>::value)
                                              && (std::is_same<Tb, var>::value || std::is_same<Tb, double>::value)
                                              && !(Ra == 1 && Cb == 1) && (Ca == Rb)>::type> {
  static Eigen::Matrix<var, Ra, Cb> run(const Eigen::Matrix<Ta, Ra, Ca>& A, const Eigen::Matrix<Tb, Rb, Cb>& B) {
    if (A.cols() == B.rows() && cuda::multiply_worth_transfer(A.rows(), A.cols(), B.cols())) {
      check_not_nan("multiply", "A", A);  // :624-625
      check_not_nan("multiply", "B", B);
      return cuda::multiply_device_rev<Ta, Ra, Ca, Tb, Cb>(A, B);
    }
    return multiply_reference_cpu(A, B);  // also throws the reference's size error
  }
};

}  // namespace cuda_dispatch
}  // namespace math
}  // namespace stan
#endif
#endif
