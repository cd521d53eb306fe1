// Overlay of stan/math/rev/mat/fun/mdivide_left_tri.hpp (see the prim overlay of the same name).
// The reference's reverse-mode overloads are kept as mdivide_left_tri_reference_cpu (found by the public forwarding
// template through argument-dependent lookup); with -DSTAN_CUDA systems of at least tri_size_worth_transfer rows with
// a matrix of var on either side get a mdivide_left_tri_vari_cuda node (stan/math/cuda/rev_mdivide_left_tri.hpp)
// -- the places where the reference switches to its OpenCL backend (:69-82, :91-105, :145-159, :170-183).
#ifndef STAN_MATH_CUDA_OVERLAY_REV_MDIVIDE_LEFT_TRI_HPP
#define STAN_MATH_CUDA_OVERLAY_REV_MDIVIDE_LEFT_TRI_HPP

// everything the reference header includes, first (see overlay/stan/math/rev/mat/fun/multiply.hpp)
#include <stan/math/rev/meta.hpp>
#include <stan/math/prim/mat/fun/Eigen.hpp>
#include <stan/math/prim/mat/err/check_multiplicable.hpp>
#include <stan/math/prim/mat/err/check_square.hpp>
#include <stan/math/rev/core.hpp>
#include <stan/math/rev/mat/fun/typedefs.hpp>
#include <stan/math/prim/mat/fun/typedefs.hpp>
#include <stan/math/prim/mat/fun/mdivide_left_tri.hpp>

#define mdivide_left_tri mdivide_left_tri_reference_cpu
#include_next <stan/math/rev/mat/fun/mdivide_left_tri.hpp>
#undef mdivide_left_tri

#ifdef STAN_CUDA
#include <stan/math/cuda/rev_mdivide_left_tri.hpp>

namespace stan {
namespace math {
namespace cuda_dispatch {

template <int TriView, typename Ta, int R1, int C1, typename Tb, int R2, int C2>
struct mdivide_left_tri_route<TriView, Eigen::Matrix<Ta, R1, C1>, Eigen::Matrix<Tb, R2, C2>,
                              typename std::enable_if<(TriView == Eigen::Lower || TriView == Eigen::Upper)
                                                      && (std::is_same<Ta, var>::value || std::is_same<Tb, var>::value)
                                                      && (std::is_same<Ta, var>::value || std::is_same<Ta, double>::value)
                                                      && (std::is_same<Tb, var>::value || std::is_same<Tb, double>::value)>::type> {
  static Eigen::Matrix<var, R1, C2> run(const Eigen::Matrix<Ta, R1, C1>& A, const Eigen::Matrix<Tb, R2, C2>& b) {
    if (A.rows() == A.cols() && A.cols() == b.rows() && cuda::tri_solve_worth_transfer(A.rows(), b.cols())) {
      return cuda::mdivide_left_tri_device_rev<TriView, Ta, R1, C1, Tb, R2, C2>(A, b);
    }
    return mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A, b);
  }
};

}  // namespace cuda_dispatch
}  // namespace math
}  // namespace stan
#endif
#endif
