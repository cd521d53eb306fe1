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
  static bool dump(const T& /*v*/, vari** /*dst*/, double* /*vals*/ = nullptr) { return false; }
};
template <>
struct operand_edge<var> {
  static size_t size(const var& /*v*/) { return 1; }
  static bool dump(const var& v, vari** dst, double* vals = nullptr) {
    dst[0] = v.vi_;
    if (vals) {
      vals[0] = v.vi_->val_;
    }
    return false;
  }
};

/** smallest edge for which the adjoint scatter of chain() is spread over the TBB workers */
constexpr size_t bulk_chain_min = size_t(1) << 17;

/**
 * Copies the vari pointers of an edge into the tape node's operand array.  For a large edge the same pass also
 * answers whether the addresses are strictly ascending -- which proves that no vari occurs twice, so the adjoint
 * updates of different elements never touch the same memory (it is what a vector of parameters built element after
 * element on the autodiff arena looks like).  Returns that answer (false for small edges: they keep the serial loop).
 */
inline bool dump_var_array(const var* v, size_t n, vari** dst, double* vals = nullptr) {
  // `vals` (optional): the same pass also gathers the operand values, e.g. straight into the page-locked staging
  // area the device call uploads from -- one walk over the N varis instead of two
#ifdef STAN_THREADS
  if (n >= bulk_chain_min) {
    std::atomic<bool> ascending(true);
    for_each_chunk(n, [&](size_t begin, size_t end) {
      bool mine = true;
      uintptr_t prev = begin > 0 ? reinterpret_cast<uintptr_t>(v[begin - 1].vi_) : 0;
      for (size_t i = begin; i != end; ++i) {
        vari* p = v[i].vi_;
        dst[i] = p;
        if (vals) {
          vals[i] = p->val_;
        }
        mine = mine && prev < reinterpret_cast<uintptr_t>(p);
        prev = reinterpret_cast<uintptr_t>(p);
      }
      if (!mine) {
        ascending.store(false, std::memory_order_relaxed);
      }
    });
    return ascending.load();
  }
#endif
  for_each_chunk(n, [&](size_t begin, size_t end) {
    for (size_t i = begin; i != end; ++i) {
      dst[i] = v[i].vi_;
      if (vals) {
        vals[i] = v[i].vi_->val_;
      }
    }
  });
  return false;
}

/** true for the operand kinds whose values the tape-node assembly can gather on its way (containers of var) */
template <typename T>
struct gathered_by_node : std::false_type {};
template <>
struct gathered_by_node<std::vector<var>> : std::true_type {};
template <int R, int C, int O, int MR, int MC>
struct gathered_by_node<Eigen::Matrix<var, R, C, O, MR, MC>> : std::true_type {};

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
  static bool dump(const std::vector<var>& v, vari** dst, double* vals = nullptr) {
    return dump_var_array(v.data(), v.size(), dst, vals);
  }
};
template <int R, int C, int O, int MR, int MC>
struct operand_edge<Eigen::Matrix<var, R, C, O, MR, MC>> {
  typedef Eigen::Matrix<var, R, C, O, MR, MC> T;
  static size_t size(const T& v) { return v.size(); }
  static bool dump(const T& v, vari** dst, double* vals = nullptr) { return dump_var_array(v.data(), v.size(), dst, vals); }
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
                          const T5& o5 = T5(), double* const* value_sinks = nullptr) {
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
    for (int e = 0; e < 5; ++e) {
      distinct_[e] = false;
    }
    if (off_[5] > 0) {
      varis_ = ChainableStack::instance_->memalloc_.alloc_array<vari*>(off_[5]);
      partials_ = ChainableStack::instance_->memalloc_.alloc_array<double>(off_[5]);
      // value_sinks[e] (optional): where the values of edge e go in the same pass (gathered_by_node kinds only)
      distinct_[0] = operand_edge<T1>::dump(o1, varis_ + off_[0], value_sinks ? value_sinks[0] : nullptr);
      distinct_[1] = operand_edge<T2>::dump(o2, varis_ + off_[1], value_sinks ? value_sinks[1] : nullptr);
      distinct_[2] = operand_edge<T3>::dump(o3, varis_ + off_[2], value_sinks ? value_sinks[2] : nullptr);
      distinct_[3] = operand_edge<T4>::dump(o4, varis_ + off_[3], value_sinks ? value_sinks[3] : nullptr);
      distinct_[4] = operand_edge<T5>::dump(o5, varis_ + off_[4], value_sinks ? value_sinks[4] : nullptr);
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
    bool any = false;
    for (int e = 0; e < 5; ++e) {
      any = any || distinct_[e];
    }
    if (any) {  // N-element edges of provably distinct operands: the adjoint scatter of chain() runs over the TBB workers
      return var(new bulk_gradients_vari(value, off_[5], varis_, partials_, off_, distinct_));
    }
    return var(new precomputed_gradients_vari(value, off_[5], varis_, partials_));
  }
  size_t n_[5];
  size_t off_[6];
  bool distinct_[5];  // edge e is large and its operand addresses ascend strictly (see dump_var_array)
  vari** varis_;
  double* partials_;
};

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
