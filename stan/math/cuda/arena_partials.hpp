#ifndef STAN_MATH_CUDA_ARENA_PARTIALS_HPP
#define STAN_MATH_CUDA_ARENA_PARTIALS_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/marshal.hpp>
#include <stan/math/rev/core/precomputed_gradients.hpp>
#include <atomic>
#include <cstdint>
#include <vector>

namespace stan {
namespace math {
namespace cuda {

/** How many reverse-mode operands an argument contributes: 0 for arithmetic arguments, 1 for a var,
 * size() for a container of vars (the edge sizes of rev/scal/meta/operands_and_partials.hpp:72-127 and
 * rev/mat/meta/operands_and_partials.hpp:15-73). */
template <typename T>
struct operand_edge {
  static size_t size(const T& /*v*/) { return 0; }
  static void dump(const T& /*v*/, vari** /*dst*/) {}
};
template <>
struct operand_edge<var> {
  static size_t size(const var& /*v*/) { return 1; }
  static void dump(const var& v, vari** dst) { dst[0] = v.vi_; }
};
inline void dump_var_array(const var* v, size_t n, vari** dst) {
  for_each_chunk(n, [&](size_t begin, size_t end) {
    for (size_t i = begin; i != end; ++i) {
      dst[i] = v[i].vi_;
    }
  });
}

/** smallest edge for which the adjoint scatter of chain() is spread over the TBB workers */
constexpr size_t bulk_chain_min = size_t(1) << 17;

/** true when the operand addresses of an edge are strictly ascending -- which proves that no vari occurs twice,
 * so the adjoint updates of different elements never touch the same memory.  It is what a vector of parameters
 * built element after element on the autodiff arena looks like; anything else keeps the serial loop. */
inline bool strictly_ascending(vari* const* v, size_t n) {
#ifdef STAN_THREADS
  std::atomic<bool> ok(true);
  for_each_chunk(n, [&](size_t begin, size_t end) {
    bool mine = true;
    for (size_t i = begin > 0 ? begin : 1; i < end; ++i) {
      mine = mine && (reinterpret_cast<uintptr_t>(v[i - 1]) < reinterpret_cast<uintptr_t>(v[i]));
    }
    if (!mine) {
      ok.store(false, std::memory_order_relaxed);
    }
  });
  return ok.load();
#else
  (void)v;
  (void)n;
  return false;
#endif
}

/**
 * The reference's precomputed_gradients_vari (rev/core/precomputed_gradients.hpp:20-76) with the adjoint scatter of
 * chain() (:71-75, one host thread over all N operands) run edge by edge over the TBB workers.  Same node layout,
 * same arithmetic per operand (one multiply-add into each operand's adjoint, so the result is bit-identical to the
 * serial loop); only edges whose operands are provably distinct (strictly ascending addresses) and large are spread,
 * the others run the reference's loop.  Edges run one after the other: the same vari may sit in two edges.
 */
class bulk_gradients_vari : public precomputed_gradients_vari {
 public:
  bulk_gradients_vari(double val, size_t size, vari** varis, double* gradients, const size_t* edge_offsets,
                      const bool* edge_parallel)
      : precomputed_gradients_vari(val, size, varis, gradients) {
    for (int e = 0; e < 6; ++e) {
      off_[e] = edge_offsets[e];
    }
    for (int e = 0; e < 5; ++e) {
      par_[e] = edge_parallel[e];
    }
  }
  void chain() {
    const double adj = adj_;
    vari** varis = varis_;
    const double* gradients = gradients_;
    for (int e = 0; e < 5; ++e) {
      const size_t b = off_[e], n = off_[e + 1] - off_[e];
      if (par_[e]) {
        for_each_chunk(n, [=](size_t begin, size_t end) {
          for (size_t i = b + begin; i != b + end; ++i) {
            varis[i]->adj_ += adj * gradients[i];
          }
        });
      } else {
        for (size_t i = b; i != b + n; ++i) {
          varis[i]->adj_ += adj * gradients[i];
        }
      }
    }
  }

 private:
  size_t off_[6];
  bool par_[5];
};
template <>
struct operand_edge<std::vector<var>> {
  static size_t size(const std::vector<var>& v) { return v.size(); }
  static void dump(const std::vector<var>& v, vari** dst) { dump_var_array(v.data(), v.size(), dst); }
};
template <int R, int C, int O, int MR, int MC>
struct operand_edge<Eigen::Matrix<var, R, C, O, MR, MC>> {
  typedef Eigen::Matrix<var, R, C, O, MR, MC> T;
  static size_t size(const T& v) { return v.size(); }
  static void dump(const T& v, vari** dst) { dump_var_array(v.data(), v.size(), dst); }
};

/**
 * The tape node of an elementwise density with N-element operands, assembled where it will live.
 *
 * operands_and_partials keeps the partials of a container edge in a zero-initialised Eigen vector of its own
 * (rev/mat/meta/operands_and_partials.hpp:21-27) and build() then copies operands and partials into two arena
 * arrays (rev/scal/meta/operands_and_partials.hpp:106-126): at N = 1e6 that is an 8 MB allocation that
 * page-faults on every evaluation plus two 8 MB copies, a third of the host time of C1.  Here the two arena
 * arrays are allocated first -- same allocator, same layout: edge 1, edge 2, edge 3 -- the device results are
 * written straight into the partials array, and the node is the reference's own precomputed_gradients_vari
 * (rev/core/precomputed_gradients.hpp:20-76, its (value, size, varis, gradients) constructor).  The arena keeps
 * its blocks across recover_memory(), so after the first evaluation nothing is allocated from the system.
 * A scalar operand has ONE slot that receives the sum over n, which is what broadcast_array accumulates.
 */
template <typename T1, typename T2 = double, typename T3 = double, typename T4 = double, typename T5 = double>
class arena_partials {
 public:
  explicit arena_partials(const T1& o1, const T2& o2 = T2(), const T3& o3 = T3(), const T4& o4 = T4(),
                          const T5& o5 = T5()) {
    n_[0] = operand_edge<T1>::size(o1);
    n_[1] = operand_edge<T2>::size(o2);
    n_[2] = operand_edge<T3>::size(o3);
    n_[3] = operand_edge<T4>::size(o4);
    n_[4] = operand_edge<T5>::size(o5);
    off_[0] = 0;
    for (int e = 1; e <= 5; ++e) {
      off_[e] = off_[e - 1] + n_[e - 1];
    }
    varis_ = nullptr;
    partials_ = nullptr;
    if (off_[5] > 0) {
      varis_ = ChainableStack::instance_->memalloc_.alloc_array<vari*>(off_[5]);
      partials_ = ChainableStack::instance_->memalloc_.alloc_array<double>(off_[5]);
      operand_edge<T1>::dump(o1, varis_ + off_[0]);
      operand_edge<T2>::dump(o2, varis_ + off_[1]);
      operand_edge<T3>::dump(o3, varis_ + off_[2]);
      operand_edge<T4>::dump(o4, varis_ + off_[3]);
      operand_edge<T5>::dump(o5, varis_ + off_[4]);
      for (int e = 0; e < 5; ++e) {  // small edges start from zero like operands_and_partials' (cheap insurance);
        if (n_[e] <= 4096) {         // an N-element edge is written entirely by the device
          zero(e);
        }
      }
    }
  }
  /** where the device writes the partials of edge e (0-based, in argument order); nullptr for an arithmetic
   * operand.  Every slot the caller asks the device for is written by it. */
  double* partials(int e) const { return n_[e] == 0 ? nullptr : partials_ + off_[e]; }
  size_t size(int e) const { return n_[e]; }
  /** sets the partials of edge e to zero (operands_and_partials starts from zeroed partials) */
  void zero(int e) const {
    for (size_t i = 0; i < n_[e]; ++i) {
      partials_[off_[e] + i] = 0.0;
    }
  }
  return_type_t<T1, T2, T3, T4, T5> build(double value) const {
    return build(value, std::integral_constant<bool, is_constant_all<T1, T2, T3, T4, T5>::value>());
  }

 private:
  double build(double value, std::true_type /*all arithmetic*/) const { return value; }
  var build(double value, std::false_type) const {
    bool par[5];
    bool any = false;
    for (int e = 0; e < 5; ++e) {
      par[e] = n_[e] >= bulk_chain_min && strictly_ascending(varis_ + off_[e], n_[e]);
      any = any || par[e];
    }
    if (any) {  // N-element edges: the adjoint scatter of chain() runs over the TBB workers
      return var(new bulk_gradients_vari(value, off_[5], varis_, partials_, off_, par));
    }
    return var(new precomputed_gradients_vari(value, off_[5], varis_, partials_));
  }
  size_t n_[5];
  size_t off_[6];
  vari** varis_;
  double* partials_;
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
