// Overlay of stan/math/prim/mat/fun/mdivide_right_tri.hpp (how overlays work: see
// overlay/stan/math/prim/mat/fun/multiply.hpp).  The reference's transparent GPU hook sits in the double overload
// (:63-72, #ifdef STAN_OPENCL: b * tri_inverse(A)); with -DSTAN_CUDA x = b A^-1 is solved on the B200 as
// x' = A^-T b' with the triangular solve of mdivide_left_tri (stan_cuda_tri_solve, transposed operand).
#ifndef STAN_MATH_CUDA_OVERLAY_PRIM_MDIVIDE_RIGHT_TRI_HPP
#define STAN_MATH_CUDA_OVERLAY_PRIM_MDIVIDE_RIGHT_TRI_HPP

#include <stan/math/prim/mat/fun/mdivide_left_tri.hpp>

#define mdivide_right_tri mdivide_right_tri_reference_cpu
#include_next <stan/math/prim/mat/fun/mdivide_right_tri.hpp>
#undef mdivide_right_tri

#ifdef STAN_CUDA
#include <stan/math/cuda/mdivide_left_tri.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {

template <Eigen::UpLoType TriView, typename TB, typename TA, typename Enable = void>
struct mdivide_right_tri_route {
  template <typename B_, typename A_>
  static auto run(const B_& b, const A_& A) -> decltype(mdivide_right_tri_reference_cpu<TriView>(b, A)) {
    return mdivide_right_tri_reference_cpu<TriView>(b, A);
  }
};

#ifdef STAN_CUDA
template <Eigen::UpLoType TriView, int R1, int C1, int R2, int C2>
struct mdivide_right_tri_route<TriView, Eigen::Matrix<double, R1, C1>, Eigen::Matrix<double, R2, C2>,
                               typename std::enable_if<TriView == Eigen::Lower || TriView == Eigen::Upper>::type> {
  static Eigen::Matrix<double, R1, C2> run(const Eigen::Matrix<double, R1, C1>& b, const Eigen::Matrix<double, R2, C2>& A) {
    if (A.rows() == A.cols() && b.cols() == A.rows() && cuda::tri_solve_worth_transfer(A.rows(), b.rows())) {
      const Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> bt = b.transpose();
      Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> xt(bt.rows(), bt.cols());
      cuda::tri_solve_on_device(TriView == Eigen::Lower, true, A.rows(), bt.cols(), A.data(), bt.data(), xt.data());
      return xt.transpose();
    }
    return mdivide_right_tri_reference_cpu<TriView>(b, A);  // also throws the size errors (:61-62)
  }
};
#endif
}  // namespace cuda_dispatch

template <Eigen::UpLoType TriView, typename TB, typename TA>
inline auto mdivide_right_tri(const TB& b, const TA& A) -> decltype(mdivide_right_tri_reference_cpu<TriView>(b, A)) {
  return cuda_dispatch::mdivide_right_tri_route<TriView, TB, TA>::run(b, A);
}

}  // namespace math
}  // namespace stan
#endif
