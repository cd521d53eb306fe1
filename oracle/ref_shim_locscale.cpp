// oracle/ref_shim_locscale.cpp -- TEST INFRASTRUCTURE, a translation unit of _ref/libstanref.so.
//
// The UNMODIFIED reference's normal_lpdf / cauchy_lpdf / lognormal_lpdf / gamma_lpdf / exponential_lpdf
// (stan/math/prim/scal/prob/normal_lpdf.hpp:38-113, cauchy_lpdf.hpp:35-113, lognormal_lpdf.hpp:19-133, gamma_lpdf.hpp:41-145,
// exponential_lpdf.hpp:42-95) behind one C entry point, for every mix of {double, var} x
// {scalar, vector} over the three operands, so that the parity tests can check each `propto` summand selection and
// each partial against the reference itself.  Compiled from the headers where they lie under /root/reference
// (oracle/Makefile); nothing is copied.
#include <stan/math/rev/mat.hpp>

#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

using stan::math::var;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_d;
typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_v;

struct Operand {
  const double* p;
  long len;
  int is_var;
  vector_d vd;
  vector_v vv;
  var sv;
  void build() {
    if (len != 1) {
      if (is_var) {
        vv.resize(len);
        for (long i = 0; i < len; ++i) vv[i] = var(p[i]);
      } else {
        vd = Eigen::Map<const vector_d>(p, len);
      }
    } else if (is_var) {
      sv = var(p[0]);
    }
  }
  int kind() const { return (len != 1 ? 2 : 0) + (is_var ? 1 : 0); }
  void adjoints(double* out) const {
    if (!out || !is_var) return;
    if (len != 1) {
      for (long i = 0; i < len; ++i) out[i] = vv[i].adj();
    } else {
      out[0] = sv.adj();
    }
  }
};

inline double value_of_result(double r) { return r; }
inline double value_of_result(const var& r) { return r.val(); }
inline void grad_of_result(double) {}
inline void grad_of_result(var& r) { r.grad(); }

template <int FAMILY, bool propto>
struct Call;
template <bool propto>
struct Call<0, propto> {
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& mu, const C& sigma) -> decltype(stan::math::normal_lpdf<propto>(y, mu, sigma)) {
    return stan::math::normal_lpdf<propto>(y, mu, sigma);
  }
};
template <bool propto>
struct Call<1, propto> {
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& mu, const C& sigma) -> decltype(stan::math::cauchy_lpdf<propto>(y, mu, sigma)) {
    return stan::math::cauchy_lpdf<propto>(y, mu, sigma);
  }
};
template <bool propto>
struct Call<2, propto> {
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& mu, const C& sigma) -> decltype(stan::math::lognormal_lpdf<propto>(y, mu, sigma)) {
    return stan::math::lognormal_lpdf<propto>(y, mu, sigma);
  }
};

template <bool propto>
struct Call<3, propto> {  // gamma_lpdf(y | alpha, beta): alpha in the mu slot, beta in the sigma slot
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& alpha, const C& beta) -> decltype(stan::math::gamma_lpdf<propto>(y, alpha, beta)) {
    return stan::math::gamma_lpdf<propto>(y, alpha, beta);
  }
};
template <bool propto>
struct Call<4, propto> {  // exponential_lpdf(y | beta): beta in the sigma slot, the mu slot is ignored
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B&, const C& beta) -> decltype(stan::math::exponential_lpdf<propto>(y, beta)) {
    return stan::math::exponential_lpdf<propto>(y, beta);
  }
};

template <bool propto>
struct Call<5, propto> {  // beta_lpdf(y | alpha, beta)
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& alpha, const C& beta) -> decltype(stan::math::beta_lpdf<propto>(y, alpha, beta)) {
    return stan::math::beta_lpdf<propto>(y, alpha, beta);
  }
};
template <bool propto>
struct Call<6, propto> {  // weibull_lpdf(y | alpha, sigma)
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& alpha, const C& sigma) -> decltype(stan::math::weibull_lpdf<propto>(y, alpha, sigma)) {
    return stan::math::weibull_lpdf<propto>(y, alpha, sigma);
  }
};

template <bool propto>
struct Call<7, propto> {  // inv_gamma_lpdf(y | alpha, beta)
  template <typename A, typename B, typename C>
  static auto go(const A& y, const B& alpha, const C& beta) -> decltype(stan::math::inv_gamma_lpdf<propto>(y, alpha, beta)) {
    return stan::math::inv_gamma_lpdf<propto>(y, alpha, beta);
  }
};

// peel one operand at a time into its static type, then call the reference
template <int FAMILY, bool propto, typename... Done>
struct Dispatch {
  template <typename... Rest>
  static double run(const Done&... done, const Operand& next, const Rest&... rest) {
    switch (next.kind()) {
      case 0: return Dispatch<FAMILY, propto, Done..., double>::run(done..., next.p[0], rest...);
      case 1: return Dispatch<FAMILY, propto, Done..., var>::run(done..., next.sv, rest...);
      case 2: return Dispatch<FAMILY, propto, Done..., vector_d>::run(done..., next.vd, rest...);
      default: return Dispatch<FAMILY, propto, Done..., vector_v>::run(done..., next.vv, rest...);
    }
  }
  static double run(const Done&... done) {
    auto r = Call<FAMILY, propto>::go(done...);
    grad_of_result(r);
    return value_of_result(r);
  }
};

template <int FAMILY>
double run_family(bool propto, const Operand& y, const Operand& mu, const Operand& sigma) {
  return propto ? Dispatch<FAMILY, true>::run(y, mu, sigma) : Dispatch<FAMILY, false>::run(y, mu, sigma);
}

}  // namespace

extern "C" {

// density 0: normal_lpdf, 1: cauchy_lpdf, 2: lognormal_lpdf, 3: gamma_lpdf(y | alpha = mu, beta = sigma),
// 4: exponential_lpdf(y | beta = sigma; pass mu as one double), 5: beta_lpdf(y | alpha = mu, beta = sigma),
// 6: weibull_lpdf(y | alpha = mu, sigma), 7: inv_gamma_lpdf(y | alpha = mu, beta = sigma).  var_mask bit 0: y, bit 1: mu, bit 2: sigma.  Lengths are 1
// (scalar) or n (vector).  Returns 0, or 1 (std::domain_error) / 2 (std::invalid_argument) / 3 (other) with the
// exception text in err.
int stanref_locscale_lpdf(int density, int propto, int var_mask, const double* y, long y_len, const double* mu, long mu_len,
                          const double* sigma, long sigma_len, double* lp, double* dy, double* dmu, double* dsigma,
                          double* seconds, char* err, int errlen) {
  int rc = 0;
  try {
    Operand oy{y, y_len, var_mask & 1}, om{mu, mu_len, (var_mask >> 1) & 1}, os{sigma, sigma_len, (var_mask >> 2) & 1};
    oy.build(); om.build(); os.build();
    auto t0 = std::chrono::steady_clock::now();
    const double r = density == 0   ? run_family<0>(propto != 0, oy, om, os)
                     : density == 1 ? run_family<1>(propto != 0, oy, om, os)
                     : density == 2 ? run_family<2>(propto != 0, oy, om, os)
                     : density == 3 ? run_family<3>(propto != 0, oy, om, os)
                     : density == 4 ? run_family<4>(propto != 0, oy, om, os)
                     : density == 5 ? run_family<5>(propto != 0, oy, om, os)
                     : density == 6 ? run_family<6>(propto != 0, oy, om, os) : run_family<7>(propto != 0, oy, om, os);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (lp) *lp = r;
    oy.adjoints(dy); om.adjoints(dmu); os.adjoints(dsigma);
  } catch (const std::exception& e) {  // same codes as ref_shim.cpp's guarded()
    if (err && errlen > 0) {
      std::strncpy(err, e.what(), size_t(errlen) - 1);
      err[errlen - 1] = 0;
    }
    rc = dynamic_cast<const std::domain_error*>(&e) ? 1 : dynamic_cast<const std::invalid_argument*>(&e) ? 2 : 3;
  }
  stan::math::recover_memory();
  return rc;
}

}  // extern "C"
