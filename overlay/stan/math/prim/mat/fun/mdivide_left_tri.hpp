// Overlay of stan/math/prim/mat/fun/mdivide_left_tri.hpp (how overlays work: see
// overlay/stan/math/prim/mat/fun/multiply.hpp).  The reference's own transparent GPU hook sits in the double
// overloads (:63-72 and :88-98, #ifdef STAN_OPENCL: tri_inverse + multiply above tri_inverse_size_worth_transfer
// rows); with -DSTAN_CUDA the same decision sends the solve to the B200 (blocked substitution).  Every overload of
// the reference header is kept, under the name mdivide_left_tri_reference_cpu; the public name is a forwarding
// template with the same explicit first template argument (callers write mdivide_left_tri<Eigen::Lower>(A, b)).
#ifndef STAN_MATH_CUDA_OVERLAY_PRIM_MDIVIDE_LEFT_TRI_HPP
#define STAN_MATH_CUDA_OVERLAY_PRIM_MDIVIDE_LEFT_TRI_HPP

#define mdivide_left_tri mdivide_left_tri_reference_cpu
#include_next <stan/math/prim/mat/fun/mdivide_left_tri.hpp>
#undef mdivide_left_tri

#ifdef STAN_CUDA
#include <stan/math/cuda/mdivide_left_tri.hpp>
#endif
#include <type_traits>

namespace stan {
namespace math {
namespace cuda_dispatch {

/** primary: the reference overload */
template <int TriView, typename TA, typename TB, typename Enable = void>
struct mdivide_left_tri_route {
  template <typename A_, typename B_>
  static auto run(const A_& A, const B_& b)
      -> decltype(mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A, b)) {
    return mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A, b);
  }
};

#ifdef STAN_CUDA
template <int TriView, int R1, int C1, int R2, int C2>
struct mdivide_left_tri_route<TriView, Eigen::Matrix<double, R1, C1>, Eigen::Matrix<double, R2, C2>,
                              typename std::enable_if<TriView == Eigen::Lower || TriView == Eigen::Upper>::type> {
  static Eigen::Matrix<double, R1, C2> run(const Eigen::Matrix<double, R1, C1>& A, const Eigen::Matrix<double, R2, C2>& b) {
    if (A.rows() == A.cols() && A.cols() == b.rows() && cuda::tri_solve_worth_transfer(A.rows(), b.cols())) {
      return cuda::mdivide_left_tri_device<TriView>(A, b);
    }
    return mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A, b);  // also throws the size errors (:61-62)
  }
};
#endif
}  // namespace cuda_dispatch

template <int TriView, typename TA, typename TB>
inline auto mdivide_left_tri(const TA& A, const TB& b)
    -> decltype(mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A, b)) {
  return cuda_dispatch::mdivide_left_tri_route<TriView, TA, TB>::run(A, b);
}

/** the inverse of a triangular matrix (:38-46, :80-100): solve against the identity */
template <int TriView, typename TA>
inline auto mdivide_left_tri(const TA& A) -> decltype(mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A)) {
  return mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A);
}

#ifdef STAN_CUDA
template <int TriView, int R1, int C1>
inline Eigen::Matrix<double, R1, C1> mdivide_left_tri(const Eigen::Matrix<double, R1, C1>& A) {
  if ((TriView == Eigen::Lower || TriView == Eigen::Upper) && A.rows() == A.cols()
      && cuda::tri_solve_worth_transfer(A.rows(), A.rows())) {
    Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> eye = Eigen::MatrixXd::Identity(A.rows(), A.rows());
    Eigen::Matrix<double, R1, C1> x(A.rows(), A.cols());
    cuda::tri_solve_on_device(TriView == Eigen::Lower, false, A.rows(), A.rows(), A.data(), eye.data(), x.data());
    return x;
  }
  return mdivide_left_tri_reference_cpu<static_cast<Eigen::UpLoType>(TriView)>(A);
}
#endif

}  // namespace math
}  // namespace stan
#endif
