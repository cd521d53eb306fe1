// CPU-suite test of map_rect's shard -> device dispatch (stan/math/cuda/map_rect_devices.hpp) against the TEST DOUBLE
// of the device library (tests/mock_device/mock_stan_cuda.c), which records the shard scope every evaluation ran in
// and reports STAN_CUDA_MOCK_DEVICES devices.  With the real library the tests return at once: the GPU suite checks
// the same dispatch on hardware through results and cache statistics (glm_parity_test.cpp, CudaMapRect).
#include <stan/math.hpp>
#include <gtest/gtest.h>
#include <dlfcn.h>
#include <cstdlib>
#include <set>
#include <stdexcept>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::MatrixXd;
using stan::math::var;

typedef int (*scope_log_fn)(int*, int);
typedef int (*scope_active_fn)(void);
static scope_log_fn mock_scope_log() { return reinterpret_cast<scope_log_fn>(dlsym(RTLD_DEFAULT, "stan_cuda_mock_scope_log")); }

struct shard_glm {
  template <typename T1, typename T2>
  Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(const Matrix<T1, Dynamic, 1>& shared,
                                                                         const Matrix<T2, Dynamic, 1>& job,
                                                                         const std::vector<double>& x_r,
                                                                         const std::vector<int>& x_i,
                                                                         std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int rows = x_i[0], k = x_i[1];
    if (x_i[2] < 0) throw std::domain_error("shard_glm: poisoned job");
    MatrixXd x = Eigen::Map<const MatrixXd>(x_r.data(), rows, k);
    std::vector<int> y(x_i.begin() + 2, x_i.end());
    Matrix<T, Dynamic, 1> res(1);
    res(0) = stan::math::poisson_log_glm_lpmf(y, x, job(0), shared);
    return res;
  }
};

static void make_jobs(int J, int rows, int k, std::vector<std::vector<double>>* x_r, std::vector<std::vector<int>>* x_i) {
  x_r->assign(J, std::vector<double>());
  x_i->assign(J, std::vector<int>());
  for (int j = 0; j < J; ++j) {
    for (int q = 0; q < rows * k; ++q) (*x_r)[j].push_back(0.1 * std::sin(0.3 * q + j));
    (*x_i)[j].push_back(rows);
    (*x_i)[j].push_back(k);
    for (int r = 0; r < rows; ++r) (*x_i)[j].push_back((r + j) % 4);
  }
}

TEST(CudaMapRectDispatch, job_i_runs_in_scope_call_id_i_device_i_mod_G) {
  scope_log_fn log = mock_scope_log();
  if (!log) return;  // real library: nothing to introspect
  const int G = stan::math::cuda::device_count();
  ASSERT_EQ(3, G) << "run with STAN_CUDA_MOCK_DEVICES=3";
  const int J = 11, rows = 50, k = 4;
  std::vector<std::vector<double>> x_r;
  std::vector<std::vector<int>> x_i;
  make_jobs(J, rows, k, &x_r, &x_i);
  std::vector<int> scratch(3 * 4096);
  log(scratch.data(), 4096);  // reset
  Matrix<var, Dynamic, 1> beta(k);
  for (int q = 0; q < k; ++q) beta[q] = 0.05 * q;
  std::vector<Matrix<var, Dynamic, 1>> job(J, Matrix<var, Dynamic, 1>(1));
  for (int j = 0; j < J; ++j) job[j][0] = 0.01 * j;
  Matrix<var, Dynamic, 1> out = stan::math::map_rect<41, shard_glm>(beta, job, x_r, x_i);
  ASSERT_EQ(J, out.size());
  const int n = log(scratch.data(), 4096);
  ASSERT_EQ(J, n);  // one GLM evaluation per job, each inside a scope
  std::set<int> jobs;
  for (int i = 0; i < n; ++i) {
    EXPECT_EQ(41, scratch[3 * i]);
    EXPECT_EQ(scratch[3 * i + 1] % G, scratch[3 * i + 2]);
    jobs.insert(scratch[3 * i + 1]);
  }
  EXPECT_EQ(static_cast<size_t>(J), jobs.size());
  stan::math::recover_memory();
}

TEST(CudaMapRectDispatch, scope_is_left_when_a_job_throws_and_outside_map_rect) {
  scope_log_fn log = mock_scope_log();
  if (!log) return;
  scope_active_fn active = reinterpret_cast<scope_active_fn>(dlsym(RTLD_DEFAULT, "stan_cuda_mock_scope_active"));
  ASSERT_TRUE(active != nullptr);
  const int J = 5, rows = 20, k = 3;
  std::vector<std::vector<double>> x_r;
  std::vector<std::vector<int>> x_i;
  make_jobs(J, rows, k, &x_r, &x_i);
  x_i[3][2] = -1;  // the functor throws for job 3
  Matrix<var, Dynamic, 1> beta(k);
  for (int q = 0; q < k; ++q) beta[q] = 0.05 * q;
  std::vector<Matrix<var, Dynamic, 1>> job(J, Matrix<var, Dynamic, 1>(1));
  for (int j = 0; j < J; ++j) job[j][0] = 0.01 * j;
  EXPECT_THROW((stan::math::map_rect<42, shard_glm>(beta, job, x_r, x_i)), std::domain_error);
  EXPECT_EQ(0, active());  // this thread ran some of the jobs: its scope must be closed again
  // an evaluation outside map_rect is not attributed to any shard
  std::vector<int> scratch(3 * 4096);
  log(scratch.data(), 4096);
  MatrixXd x = Eigen::Map<const MatrixXd>(x_r[0].data(), rows, k);
  std::vector<int> y(x_i[0].begin() + 2, x_i[0].end());
  var lp = stan::math::poisson_log_glm_lpmf(y, x, job[0][0], beta);
  EXPECT_TRUE(std::isfinite(lp.val()));
  EXPECT_EQ(0, log(scratch.data(), 4096));
  stan::math::recover_memory();
}

// A job whose functor evaluates a vectorised normal_lpdf large enough for the parallel operand gather
// (stan/math/cuda/normal_lpdf.hpp: N >= 65536).  The gather is a nested tbb::parallel_for: while a worker waits in
// it, TBB may hand it ANOTHER map_rect job, whose evaluation uses the same thread-local staging buffer.  The gathers
// therefore run isolated (tbb::this_task_arena::isolate); this test fails without the isolation.
struct shard_normal {
  template <typename T1, typename T2>
  Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(const Matrix<T1, Dynamic, 1>& shared,
                                                                         const Matrix<T2, Dynamic, 1>& job,
                                                                         const std::vector<double>& x_r,
                                                                         const std::vector<int>& x_i,
                                                                         std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int n = x_i[0];
    Matrix<T, Dynamic, 1> mu(n), sg(n);
    for (int i = 0; i < n; ++i) {
      mu[i] = shared[0] + job[0] * x_r[i];
      sg[i] = shared[1] + 0.001 * (i % 7);
    }
    std::vector<double> y(x_r.begin(), x_r.end());
    Matrix<T, Dynamic, 1> res(1);
    res(0) = stan::math::normal_lpdf(y, mu, sg);
    return res;
  }
};

TEST(CudaMapRectDispatch, nested_parallel_gather_does_not_mix_jobs) {
  const int J = 24, n = 70000;
  std::vector<std::vector<double>> x_r(J, std::vector<double>(n));
  std::vector<std::vector<int>> x_i(J, std::vector<int>(1, n));
  for (int j = 0; j < J; ++j)
    for (int i = 0; i < n; ++i) x_r[j][i] = std::sin(0.01 * i + j);
  for (int rep = 0; rep < 3; ++rep) {
    Matrix<var, Dynamic, 1> shared(2);
    shared << 0.1, 1.2;
    std::vector<Matrix<var, Dynamic, 1>> job(J, Matrix<var, Dynamic, 1>(1));
    for (int j = 0; j < J; ++j) job[j][0] = 0.5 + 0.1 * j;
    Matrix<var, Dynamic, 1> out = stan::math::map_rect<43, shard_normal>(shared, job, x_r, x_i);
    for (int j = 0; j < J; ++j) {
      double ref = 0;  // normal_lpdf.hpp:89-97 by hand
      for (int i = 0; i < n; ++i) {
        const double mu = 0.1 + (0.5 + 0.1 * j) * x_r[j][i], sg = 1.2 + 0.001 * (i % 7);
        const double z = (x_r[j][i] - mu) / sg;
        ref += -0.91893853320467274178 - std::log(sg) - 0.5 * z * z;
      }
      EXPECT_NEAR(ref, out[j].val(), 1e-9 * std::fabs(ref)) << "job " << j << " rep " << rep;
    }
    stan::math::recover_memory();
  }
}

TEST(CudaMapRectDispatch, scopes_nest) {
  scope_log_fn log = mock_scope_log();
  if (!log) return;
  std::vector<int> scratch(3 * 4096);
  log(scratch.data(), 4096);
  const int rows = 10, k = 2;
  std::vector<std::vector<double>> x_r;
  std::vector<std::vector<int>> x_i;
  make_jobs(1, rows, k, &x_r, &x_i);
  MatrixXd x = Eigen::Map<const MatrixXd>(x_r[0].data(), rows, k);
  std::vector<int> y(x_i[0].begin() + 2, x_i[0].end());
  Eigen::VectorXd beta = Eigen::VectorXd::Constant(k, 0.1);
  {
    stan::math::cuda::shard_scope outer(7, 1, 1);
    {
      stan::math::cuda::shard_scope inner(7, 2, 2);
      stan::math::poisson_log_glm_lpmf(y, x, 0.3, beta);
    }
    stan::math::poisson_log_glm_lpmf(y, x, 0.3, beta);  // back in the outer scope
  }
  stan::math::poisson_log_glm_lpmf(y, x, 0.3, beta);  // no scope
  ASSERT_EQ(2, log(scratch.data(), 4096));
  EXPECT_EQ(2, scratch[1]);
  EXPECT_EQ(2, scratch[2]);
  EXPECT_EQ(1, scratch[4]);
  EXPECT_EQ(1, scratch[5]);
}
