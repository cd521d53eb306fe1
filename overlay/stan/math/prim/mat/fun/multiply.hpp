// Overlay of stan/math/prim/mat/fun/multiply.hpp (how overlays work: see
// overlay/stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp).  SURVEY 8f rank 4: the reference's own
// transparent GPU hook sits inside multiply(m1, m2) for double matrices (:71-79, #ifdef STAN_OPENCL); with
// -DSTAN_CUDA the same decision -- product of the three dimensions above a transfer threshold -- sends the product
// to the B200 instead.  Every overload of the reference header is kept, under the name multiply_reference_cpu; the
// public name is ONE forwarding template whose return type is whatever the reference overload returns, so every
// call that compiled before compiles now and resolves to the same reference overload unless a route is
// specialised for its argument types (here: two matrices of double; in the rev overlay: matrices with var).
#ifndef STAN_MATH_CUDA_OVERLAY_PRIM_MULTIPLY_HPP
#define STAN_MATH_CUDA_OVERLAY_PRIM_MULTIPLY_HPP

#define multiply multiply_reference_cpu
#include_next <stan/math/prim/mat/fun/multiply.hpp>
#undef multiply

#ifdef STAN_CUDA
#include <stan/math/cuda/multiply.hpp>
#endif
#include <type_traits>
#include <utility>

namespace stan {
namespace math {
namespace cuda_dispatch {

/** primary: the reference overload */
template <typename T1, typename T2, typename Enable = void>
struct multiply_route {
  template <typename A, typename B>
  static auto run(const A& a, const B& b) -> decltype(multiply_reference_cpu(a, b)) {
    return multiply_reference_cpu(a, b);
  }
};

#ifdef STAN_CUDA
/** matrix of double x matrix of double, not the row-vector x column-vector dot product (:99-104) */
template <int R1, int C1, int R2, int C2>
struct multiply_route<Eigen::Matrix<double, R1, C1>, Eigen::Matrix<double, R2, C2>,
                      typename std::enable_if<!(R1 == 1 && C2 == 1)>::type> {
  static Eigen::Matrix<double, R1, C2> run(const Eigen::Matrix<double, R1, C1>& m1, const Eigen::Matrix<double, R2, C2>& m2) {
    check_multiplicable("multiply", "m1", m1, "m2", m2);  // :69
    if (cuda::multiply_worth_transfer(m1.rows(), m1.cols(), m2.cols())) {
      return cuda::multiply_device(m1, m2);
    }
    return multiply_reference_cpu(m1, m2);
  }
};
#endif
}  // namespace cuda_dispatch

template <typename T1, typename T2>
inline auto multiply(const T1& a, const T2& b) -> decltype(multiply_reference_cpu(a, b)) {
  return cuda_dispatch::multiply_route<T1, T2>::run(a, b);
}

}  // namespace math
}  // namespace stan
#endif
