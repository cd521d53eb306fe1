// bench/host_phases.cpp -- where the HOST time of an N-operand elementwise density goes (C1: normal_lpdf, y data,
// mu a var vector of N = 1e6, sigma a var), phase by phase, for the two ways of assembling the tape node:
//
//   reference style   operands_and_partials edges (zero-initialised Eigen partials), values gathered through
//                     scalar_seq_view (which copies the container), build() copying operands and partials
//   device path       stan/math/cuda/arena_partials.hpp + seq_access gather (what normal_lpdf_device does)
//
// The device call in the middle is whatever libstanmath_cuda.so the binary finds: on a GPU box the real library,
// in the CPU container tests/mock_device (LD_LIBRARY_PATH=build/mock), whose time is printed as its own column and
// is NOT a device number.  Everything else is host work and does not depend on the device.
//
//   make -C test ../build/cxx/host_phases && build/cxx/host_phases [N] [reps]
#include <stan/math.hpp>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace stan::math;  // NOLINT
typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_var;

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Phases {
  double node = 0, gather = 0, eval = 0, build = 0, grad = 0, read = 0;
};

static void fill_args(stan_cuda_normal_args* a, const Eigen::VectorXd& y, const double* mu, int64_t n, const double* sigma,
                      double* logp, double* dmu, double* dsigma) {
  *a = stan_cuda_normal_args();
  a->n = n;
  a->y = y.data();
  a->y_len = n;
  a->y_is_data = 1;
  a->mu = mu;
  a->mu_len = n;
  a->sigma = sigma;
  a->sigma_len = 1;
  a->lp = logp;
  a->flags = STAN_CUDA_LP_ALL | STAN_CUDA_NEED_DALPHA | STAN_CUDA_NEED_DSIGMA;
  a->dmu = dmu;
  a->dsigma = dsigma;
}

template <bool ARENA>
static double once(const Eigen::VectorXd& y, const Eigen::VectorXd& mu_val, std::vector<double>* adj, Phases* ph) {
  const int64_t n = y.size();
  start_nested();
  vector_var mu(n);
  for (int64_t i = 0; i < n; ++i) mu[i] = mu_val[i];
  var sigma = 1.3;
  const double sg = 1.3;
  double logp = 0;
  double* stage = static_cast<double*>(stan_cuda_host_staging((n + 2) * sizeof(double)));
  stan_cuda_normal_args a;
  var lp;
  double t0, t1, t2, t3, t4;
  if (ARENA) {
    t0 = now_ms();
    cuda::arena_partials<Eigen::VectorXd, vector_var, var> node(y, mu, sigma);
    t1 = now_ms();
    cuda::staged_operand<vector_var>::fill(mu, stage);
    t2 = now_ms();
    fill_args(&a, y, stage, n, &sg, &logp, node.partials(1), node.partials(2));
    stan_cuda_normal_lpdf_eval(&a);
    t3 = now_ms();
    lp = node.build(logp);
    t4 = now_ms();
  } else {
    t0 = now_ms();
    operands_and_partials<Eigen::VectorXd, vector_var, var> node(y, mu, sigma);
    t1 = now_ms();
    {
      stan::scalar_seq_view<vector_var> view(mu);  // prim/scal/meta/scalar_seq_view.hpp:36 copies the container
      for (int64_t i = 0; i < n; ++i) stage[i] = value_of_rec(view[i]);
    }
    t2 = now_ms();
    fill_args(&a, y, stage, n, &sg, &logp, cuda::partials_data(node.edge2_), cuda::partials_data(node.edge3_));
    stan_cuda_normal_lpdf_eval(&a);
    t3 = now_ms();
    lp = node.build(logp);
    t4 = now_ms();
  }
  grad(lp.vi_);
  const double t5 = now_ms();
  for (int64_t i = 0; i < n; ++i) (*adj)[i] = mu[i].adj();
  (*adj)[n] = sigma.adj();
  const double t6 = now_ms();
  recover_memory_nested();
  ph->node += t1 - t0;
  ph->gather += t2 - t1;
  ph->eval += t3 - t2;
  ph->build += t4 - t3;
  ph->grad += t5 - t4;
  ph->read += t6 - t5;
  return logp;
}

int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? std::atoll(argv[1]) : 1000000;
  const int reps = argc > 2 ? std::atoi(argv[2]) : 15;
  init_threadpool_tbb();
  Eigen::VectorXd y(n), mu(n);
  for (int64_t i = 0; i < n; ++i) {
    y[i] = std::sin(0.37 * i);
    mu[i] = 0.1 * std::cos(0.11 * i);
  }
  std::vector<double> adj_a(n + 1), adj_r(n + 1);
  Phases pa, pr, warm;
  double la = 0, lr = 0;
  for (int r = 0; r < 3; ++r) {
    once<true>(y, mu, &adj_a, &warm);
    once<false>(y, mu, &adj_r, &warm);
  }
  for (int r = 0; r < reps; ++r) {
    la = once<true>(y, mu, &adj_a, &pa);
    lr = once<false>(y, mu, &adj_r, &pr);
  }
  double worst = std::fabs(la - lr);
  for (int64_t i = 0; i <= n; ++i) worst = std::max(worst, std::fabs(adj_a[i] - adj_r[i]));
  std::printf("{\"what\": \"host phases of normal_lpdf, mu var vector N=%lld, ms per evaluation, %d reps\", \"device\": \"%s\",\n",
              (long long)n, reps, stan_cuda_describe());
  const Phases* p[2] = {&pr, &pa};
  const char* name[2] = {"reference_style", "device_path"};
  for (int k = 0; k < 2; ++k) {
    std::printf(" \"%s\": {\"node_setup\": %.2f, \"gather_values\": %.2f, \"device_call\": %.2f, \"build\": %.2f, \"grad\": %.2f, "
                "\"read_adjoints\": %.2f, \"host_forward\": %.2f},\n",
                name[k], p[k]->node / reps, p[k]->gather / reps, p[k]->eval / reps, p[k]->build / reps, p[k]->grad / reps,
                p[k]->read / reps, (p[k]->node + p[k]->gather + p[k]->build) / reps);
  }
  std::printf(" \"max_abs_difference_of_results\": %.3e}\n", worst);
  std::fflush(stdout);
  std::_Exit(0);  // as bench/stan_bench.cpp: the reference's tape observer races with TBB's shutdown
}
