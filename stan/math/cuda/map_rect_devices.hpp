#ifndef STAN_MATH_CUDA_MAP_RECT_DEVICES_HPP
#define STAN_MATH_CUDA_MAP_RECT_DEVICES_HPP
#ifdef STAN_CUDA

#include <stan/math/cuda/context.hpp>
#include <stan/math/prim/mat/fun/typedefs.hpp>
#include <stan/math/prim/mat/functor/map_rect_combine.hpp>
#include <stan/math/prim/mat/functor/map_rect_reduce.hpp>
#include <stan/math/prim/meta.hpp>

#ifdef STAN_THREADS
#include <tbb/blocked_range.h>
#include <tbb/parallel_for.h>
#endif

#include <vector>

namespace stan {
namespace math {
namespace cuda {

/**
 * Shard -> device dispatch for map_rect.  Job i runs inside a shard scope (call_id, i, device
 * i mod G): every GLM / normal_lpdf evaluated by the user functor for that job lands on that
 * device, and its data is cached under (call_id, i) -- the functor rebuilds its design matrix
 * from x_r[i] at a fresh address on every call, so a pointer is no identity there; (call_id, job)
 * is the identity the reference's MPI backend already uses for the immutable job data
 * (stan/math/prim/mat/functor/mpi_parallel_call.hpp:40-89, map_rect.hpp:97-100).
 *
 * Takes the place of the loop in stan/math/rev/mat/functor/map_rect_concurrent.hpp:39-54; the
 * nested-autodiff reduce per job and the combine step are the reference's own classes, untouched.
 * With STAN_THREADS the jobs still fan out over TBB workers (each has its own tape and its own
 * stream per device), so G devices are kept busy by >= G host threads.
 */
template <int call_id, typename F, typename T_shared_param, typename T_job_param>
Eigen::Matrix<typename stan::return_type<T_shared_param, T_job_param>::type, Eigen::Dynamic, 1>
map_rect_on_devices(const Eigen::Matrix<T_shared_param, Eigen::Dynamic, 1>& shared_params,
                    const std::vector<Eigen::Matrix<T_job_param, Eigen::Dynamic, 1>>& job_params,
                    const std::vector<std::vector<double>>& x_r, const std::vector<std::vector<int>>& x_i,
                    std::ostream* msgs) {
  using reduce_t = internal::map_rect_reduce<F, T_shared_param, T_job_param>;
  using combine_t = internal::map_rect_combine<F, T_shared_param, T_job_param>;

  const std::size_t num_jobs = job_params.size();
  int n_devices = device_count();
  if (n_devices < 1) {
    n_devices = 1;
  }
  const vector_d shared_dbl = value_of(shared_params);
  std::vector<matrix_d> per_job(num_jobs);

  auto run_job = [&](std::size_t i) {
    shard_scope scope(call_id, static_cast<int>(i), static_cast<int>(i % n_devices));
    per_job[i] = reduce_t()(shared_dbl, value_of(job_params[i]), x_r[i], x_i[i], msgs);
  };

#ifdef STAN_THREADS
  tbb::parallel_for(tbb::blocked_range<std::size_t>(0, num_jobs), [&](const tbb::blocked_range<std::size_t>& r) {
    for (std::size_t i = r.begin(); i != r.end(); ++i) {
      run_job(i);
    }
  });
#else
  for (std::size_t i = 0; i < num_jobs; ++i) {
    run_job(i);
  }
#endif

  std::vector<int> outputs_per_job(num_jobs, 0);
  int total_outputs = 0;
  for (std::size_t i = 0; i < num_jobs; ++i) {
    outputs_per_job[i] = static_cast<int>(per_job[i].cols());
    total_outputs += outputs_per_job[i];
  }
  matrix_d gathered(num_jobs > 0 ? per_job[0].rows() : 0, total_outputs);  // (map_rect.hpp:164-166 returns before 0 jobs)
  int col = 0;
  for (std::size_t i = 0; i < num_jobs; ++i) {
    if (outputs_per_job[i] > 0) {
      gathered.middleCols(col, outputs_per_job[i]) = per_job[i];
    }
    col += outputs_per_job[i];
  }
  combine_t combine(shared_params, job_params);
  return combine(gathered, outputs_per_job);
}

}  // namespace cuda
}  // namespace math
}  // namespace stan
#endif
#endif
