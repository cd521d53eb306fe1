// GPU-vs-CPU parity of the STAN_CUDA path through the reference's own C++ signatures.
// Modelled on the reference's OpenCL-vs-CPU tests (test/unit/math/opencl/rev/*_glm_*_test.cpp):
// same sizes (3x2 fixtures, N = 0, M = 0, vector alpha / sigma, "big" 153 x 71) and the same
// error cases, but at the north star's tolerance of 1e-10 instead of 1e-8.  The CPU side is the
// unmodified reference body, reachable in the same binary as *_reference_cpu (see overlay/).
#ifdef STAN_CUDA
#include <stan/math.hpp>
#include <gtest/gtest.h>
#include <random>
#include <vector>

using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::MatrixXd;
using Eigen::VectorXd;
using stan::math::var;
typedef Matrix<var, Dynamic, 1> vector_v;

namespace {

void expect_rel(double got, double ref, double scale, const char* what) {
  const double denom = std::max(std::fabs(ref), 1e-6 * scale);
  if (denom == 0) {
    EXPECT_EQ(got, ref) << what;
  } else {
    EXPECT_LE(std::fabs(got - ref) / denom, 1e-10) << what << " got " << got << " ref " << ref;
  }
}

void expect_vec(const VectorXd& got, const VectorXd& ref, const char* what) {
  ASSERT_EQ(got.size(), ref.size()) << what;
  const double scale = ref.size() ? ref.cwiseAbs().maxCoeff() : 0.0;
  for (int i = 0; i < ref.size(); ++i) {
    expect_rel(got[i], ref[i], scale, what);
  }
}

VectorXd adj(const vector_v& v) {
  VectorXd a(v.size());
  for (int i = 0; i < v.size(); ++i) a[i] = v[i].adj();
  return a;
}

vector_v to_v(const VectorXd& d) {
  vector_v v(d.size());
  for (int i = 0; i < d.size(); ++i) v[i] = d[i];
  return v;
}

struct Problem {
  MatrixXd x;
  std::vector<int> y_bern, y_pois;
  VectorXd y_norm, beta, alpha_vec, sigma_vec;
  double alpha, sigma;
};

Problem make_problem(int n, int k, unsigned seed) {
  std::mt19937_64 rng(seed);
  std::normal_distribution<double> nd(0, 1);
  std::uniform_real_distribution<double> ud(0, 1);
  Problem p;
  p.x.resize(n, k);
  for (int j = 0; j < k; ++j)
    for (int i = 0; i < n; ++i) p.x(i, j) = nd(rng);
  VectorXd beta_true(k);
  for (int j = 0; j < k; ++j) beta_true[j] = nd(rng) * 0.6 / std::sqrt(std::max(k, 1));
  p.beta = beta_true;
  for (int j = 0; j < k; ++j) p.beta[j] += 0.05 * nd(rng);
  p.alpha = 0.3;
  p.sigma = 1.3;
  p.y_bern.resize(n);
  p.y_pois.resize(n);
  p.y_norm.resize(n);
  p.alpha_vec.resize(n);
  p.sigma_vec.resize(n);
  for (int i = 0; i < n; ++i) {
    double eta = 0.2 + (k ? p.x.row(i).dot(beta_true) : 0.0);
    p.y_bern[i] = ud(rng) < 1 / (1 + std::exp(-eta)) ? 1 : 0;
    p.y_pois[i] = std::poisson_distribution<int>(std::exp(eta))(rng);
    p.y_norm[i] = eta + 0.7 * nd(rng);
    p.alpha_vec[i] = 0.3 + 0.1 * nd(rng);
    p.sigma_vec[i] = 0.5 + ud(rng);
  }
  return p;
}

}  // namespace

#define STAN_CUDA_PARITY(NAME, GPU_CALL, CPU_CALL, ...)                                     \
  {                                                                                         \
    var lp_gpu = GPU_CALL;                                                                  \
    lp_gpu.grad();                                                                          \
    const double v_gpu = lp_gpu.val();                                                      \
    std::vector<VectorXd> g_gpu = {__VA_ARGS__};                                            \
    stan::math::set_zero_all_adjoints();                                                    \
    var lp_cpu = CPU_CALL;                                                                  \
    lp_cpu.grad();                                                                          \
    std::vector<VectorXd> g_cpu = {__VA_ARGS__};                                            \
    expect_rel(v_gpu, lp_cpu.val(), std::fabs(lp_cpu.val()), NAME " value");                \
    for (size_t q = 0; q < g_cpu.size(); ++q) expect_vec(g_gpu[q], g_cpu[q], NAME " grad"); \
    stan::math::set_zero_all_adjoints();                                                    \
  }

class CudaGlm : public ::testing::TestWithParam<std::pair<int, int>> {};

TEST_P(CudaGlm, bernoulli_poisson_normal_match_reference_body) {
  const int n = GetParam().first, k = GetParam().second;
  Problem p = make_problem(n, k, 1000 + n + k);
  vector_v beta = to_v(p.beta);
  var alpha = p.alpha, sigma = p.sigma;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("bernoulli", bernoulli_logit_glm_lpmf(p.y_bern, p.x, alpha, beta),
                   bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("bernoulli propto", bernoulli_logit_glm_lpmf<true>(p.y_bern, p.x, alpha, beta),
                   bernoulli_logit_glm_lpmf_reference_cpu<true>(p.y_bern, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("poisson", poisson_log_glm_lpmf(p.y_pois, p.x, alpha, beta),
                   poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("poisson propto", poisson_log_glm_lpmf<true>(p.y_pois, p.x, alpha, beta),
                   poisson_log_glm_lpmf_reference_cpu<true>(p.y_pois, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("normal_id", normal_id_glm_lpdf(p.y_norm, p.x, alpha, beta, sigma),
                   normal_id_glm_lpdf_reference_cpu(p.y_norm, p.x, alpha, beta, sigma), adj(beta),
                   VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, sigma.adj()));
  STAN_CUDA_PARITY("normal_id propto", normal_id_glm_lpdf<true>(p.y_norm, p.x, alpha, beta, sigma),
                   normal_id_glm_lpdf_reference_cpu<true>(p.y_norm, p.x, alpha, beta, sigma), adj(beta),
                   VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, sigma.adj()));
  var phi = 2.5;
  std::vector<int> y_nb = p.y_pois;
  for (size_t i = 0; i < y_nb.size(); ++i) y_nb[i] = (y_nb[i] * 7 + static_cast<int>(i % 5)) % 200;
  STAN_CUDA_PARITY("neg_binomial_2_log", neg_binomial_2_log_glm_lpmf(y_nb, p.x, alpha, beta, phi),
                   neg_binomial_2_log_glm_lpmf_reference_cpu(y_nb, p.x, alpha, beta, phi), adj(beta),
                   VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, phi.adj()));
  STAN_CUDA_PARITY("neg_binomial_2_log propto", neg_binomial_2_log_glm_lpmf<true>(y_nb, p.x, alpha, beta, phi),
                   neg_binomial_2_log_glm_lpmf_reference_cpu<true>(y_nb, p.x, alpha, beta, phi), adj(beta),
                   VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, phi.adj()));
  // all-double call returns a double through the same device path
  EXPECT_NEAR(poisson_log_glm_lpmf(p.y_pois, p.x, p.alpha, p.beta),
              poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, p.alpha, p.beta),
              1e-10 * std::fabs(poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, p.alpha, p.beta)) + 1e-300);
  stan::math::recover_memory();
}

INSTANTIATE_TEST_CASE_P(Shapes, CudaGlm,
                        ::testing::Values(std::make_pair(3, 2), std::make_pair(1, 1), std::make_pair(153, 71),
                                          std::make_pair(5000, 32), std::make_pair(20000, 128),
                                          std::make_pair(3000, 256), std::make_pair(40, 0),
                                          // wider than one warp tile (k > 512): the column-block path
                                          std::make_pair(300, 513), std::make_pair(800, 1100)));

TEST(CudaGlmVariants, vector_alpha_vector_sigma_var_y) {
  Problem p = make_problem(153, 71, 7);
  vector_v beta = to_v(p.beta), alpha_v = to_v(p.alpha_vec), sigma_v = to_v(p.sigma_vec), y_v = to_v(p.y_norm);
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("bernoulli vector alpha", bernoulli_logit_glm_lpmf(p.y_bern, p.x, alpha_v, beta),
                   bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, p.x, alpha_v, beta), adj(beta), adj(alpha_v));
  STAN_CUDA_PARITY("poisson vector alpha", poisson_log_glm_lpmf(p.y_pois, p.x, alpha_v, beta),
                   poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, alpha_v, beta), adj(beta), adj(alpha_v));
  STAN_CUDA_PARITY("normal all vectors + var y", normal_id_glm_lpdf(y_v, p.x, alpha_v, beta, sigma_v),
                   normal_id_glm_lpdf_reference_cpu(y_v, p.x, alpha_v, beta, sigma_v), adj(beta), adj(alpha_v),
                   adj(sigma_v), adj(y_v));
  vector_v phi_v = to_v(p.sigma_vec);
  STAN_CUDA_PARITY("neg_binomial_2_log vector alpha and phi", neg_binomial_2_log_glm_lpmf(p.y_pois, p.x, alpha_v, beta, phi_v),
                   neg_binomial_2_log_glm_lpmf_reference_cpu(p.y_pois, p.x, alpha_v, beta, phi_v), adj(beta),
                   adj(alpha_v), adj(phi_v));
  // std::vector<var> and row-vector operands
  std::vector<var> beta_std(p.beta.data(), p.beta.data() + p.beta.size());
  Matrix<var, 1, Dynamic> beta_row = beta.transpose();
  var alpha = 0.3;
  var a = poisson_log_glm_lpmf(p.y_pois, p.x, alpha, beta_std);
  var b = poisson_log_glm_lpmf(p.y_pois, p.x, alpha, beta_row);
  var c = poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, alpha, beta);
  expect_rel(a.val(), c.val(), std::fabs(c.val()), "std::vector<var> beta");
  expect_rel(b.val(), c.val(), std::fabs(c.val()), "row-vector beta");
  // Eigen int outcome vector
  Matrix<int, Dynamic, 1> y_eig = Eigen::Map<Matrix<int, Dynamic, 1>>(p.y_pois.data(), p.y_pois.size());
  var d = poisson_log_glm_lpmf(y_eig, p.x, alpha, beta);
  expect_rel(d.val(), c.val(), std::fabs(c.val()), "Eigen int y");
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, wide_matrix_vector_alpha_vector_sigma_var_y) {
  // k > 512 with per-row operands: the carry chain starts from the intercept vector, d comes back per row
  Problem p = make_problem(211, 600, 9);
  vector_v beta = to_v(p.beta), alpha_v = to_v(p.alpha_vec), sigma_v = to_v(p.sigma_vec), y_v = to_v(p.y_norm);
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("wide bernoulli vector alpha", bernoulli_logit_glm_lpmf(p.y_bern, p.x, alpha_v, beta),
                   bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, p.x, alpha_v, beta), adj(beta), adj(alpha_v));
  STAN_CUDA_PARITY("wide normal all vectors + var y", normal_id_glm_lpdf(y_v, p.x, alpha_v, beta, sigma_v),
                   normal_id_glm_lpdf_reference_cpu(y_v, p.x, alpha_v, beta, sigma_v), adj(beta), adj(alpha_v),
                   adj(sigma_v), adj(y_v));
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, var_design_matrix) {
  // x as a parameter -- d/dx = d * beta^T (bernoulli_logit_glm_lpmf.hpp:125-128,
  // poisson_log_glm_lpmf.hpp:119-122, normal_id_glm_lpdf.hpp:121-124), the reference's 3 x 2 fixture shape and larger
  const int shapes[4][2] = {{3, 2}, {153, 71}, {2000, 128}, {300, 600}};
  for (int q = 0; q < 4; ++q) {
    Problem p = make_problem(shapes[q][0], shapes[q][1], 40 + q);
    Matrix<var, Dynamic, Dynamic> xv(p.x.rows(), p.x.cols());
    for (int64_t i = 0; i < p.x.size(); ++i) xv.data()[i] = p.x.data()[i];
    auto adj_x = [&]() {
      VectorXd g(xv.size());
      for (int64_t i = 0; i < xv.size(); ++i) g[i] = xv.data()[i].adj();
      return g;
    };
    vector_v beta = to_v(p.beta), alpha_v = to_v(p.alpha_vec);
    var alpha = p.alpha, sigma = p.sigma, phi = 2.5;
    using namespace stan::math;  // NOLINT
    STAN_CUDA_PARITY("bernoulli var x", bernoulli_logit_glm_lpmf(p.y_bern, xv, alpha, beta),
                     bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, xv, alpha, beta), adj(beta),
                     VectorXd::Constant(1, alpha.adj()), adj_x());
    STAN_CUDA_PARITY("poisson var x, vector alpha", poisson_log_glm_lpmf(p.y_pois, xv, alpha_v, beta),
                     poisson_log_glm_lpmf_reference_cpu(p.y_pois, xv, alpha_v, beta), adj(beta), adj(alpha_v), adj_x());
    STAN_CUDA_PARITY("poisson var x only", poisson_log_glm_lpmf(p.y_pois, xv, p.alpha, p.beta),
                     poisson_log_glm_lpmf_reference_cpu(p.y_pois, xv, p.alpha, p.beta), adj_x());
    STAN_CUDA_PARITY("normal_id var x", normal_id_glm_lpdf(p.y_norm, xv, alpha, beta, sigma),
                     normal_id_glm_lpdf_reference_cpu(p.y_norm, xv, alpha, beta, sigma), adj(beta),
                     VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, sigma.adj()), adj_x());
    STAN_CUDA_PARITY("neg_binomial_2_log var x", neg_binomial_2_log_glm_lpmf(p.y_pois, xv, alpha, beta, phi),
                     neg_binomial_2_log_glm_lpmf_reference_cpu(p.y_pois, xv, alpha, beta, phi), adj(beta),
                     VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, phi.adj()), adj_x());
    // the values of x change between calls: nothing may be served from a cache
    for (int64_t i = 0; i < xv.size(); ++i) xv.data()[i] = 0.5 * p.x.data()[i];
    STAN_CUDA_PARITY("bernoulli var x, new values at the same shape", bernoulli_logit_glm_lpmf(p.y_bern, xv, alpha, beta),
                     bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, xv, alpha, beta), adj(beta),
                     VectorXd::Constant(1, alpha.adj()), adj_x());
    stan::math::recover_memory();
  }
}

TEST(CudaGlmVariants, scalar_weight_broadcast_over_several_columns) {
  // a scalar beta passes check_consistent_size for any number of columns; the device path broadcasts it and sums the K
  // gradient components into its ONE partials slot (round 1 copied K doubles over that slot: advisor finding)
  Problem p = make_problem(301, 7, 21);
  using namespace stan::math;  // NOLINT
  var alpha = 0.3, b = 0.11;
  var lp = poisson_log_glm_lpmf(p.y_pois, p.x, alpha, b);
  lp.grad();
  const double v = lp.val(), gb = b.adj(), ga = alpha.adj();
  set_zero_all_adjoints();
  vector_v beta = to_v(VectorXd::Constant(7, 0.11));
  var alpha2 = 0.3;
  var lp2 = poisson_log_glm_lpmf_reference_cpu(p.y_pois, p.x, alpha2, beta);
  lp2.grad();
  expect_rel(v, lp2.val(), std::fabs(lp2.val()), "scalar beta value");
  expect_rel(gb, adj(beta).sum(), std::fabs(adj(beta).sum()), "scalar beta gradient = sum of the components");
  expect_rel(ga, alpha2.adj(), std::fabs(alpha2.adj()), "alpha gradient");
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, ordered_logistic) {
  using stan::math::ordered_logistic_glm_lpmf;
  using stan::math::ordered_logistic_glm_lpmf_reference_cpu;
  Problem p = make_problem(5003, 37, 11);
  vector_v beta = to_v(p.beta);
  VectorXd cuts_d(4);
  cuts_d << -0.8, -0.1, 0.4, 1.3;
  vector_v cuts = to_v(cuts_d);
  std::vector<int> y(p.y_pois.size());
  for (size_t i = 0; i < y.size(); ++i) y[i] = 1 + (p.y_pois[i] + static_cast<int>(i)) % 5;
  STAN_CUDA_PARITY("ordered_logistic", ordered_logistic_glm_lpmf(y, p.x, beta, cuts),
                   ordered_logistic_glm_lpmf_reference_cpu(y, p.x, beta, cuts), adj(beta), adj(cuts));
  STAN_CUDA_PARITY("ordered_logistic propto, data cuts", ordered_logistic_glm_lpmf<true>(y, p.x, beta, cuts_d),
                   ordered_logistic_glm_lpmf_reference_cpu<true>(y, p.x, beta, cuts_d), adj(beta));
  STAN_CUDA_PARITY("ordered_logistic data beta", ordered_logistic_glm_lpmf(y, p.x, p.beta, cuts),
                   ordered_logistic_glm_lpmf_reference_cpu(y, p.x, p.beta, cuts), adj(cuts));
  EXPECT_NEAR(ordered_logistic_glm_lpmf(y, p.x, p.beta, cuts_d),
              ordered_logistic_glm_lpmf_reference_cpu(y, p.x, p.beta, cuts_d), 1e-8);
  EXPECT_EQ(ordered_logistic_glm_lpmf<true>(y, p.x, p.beta, cuts_d), 0.0);
  // out-of-range class: same exception as the reference, thrown from the device's cached range of y
  std::vector<int> y_high = y, y_low = y;
  y_high[17] = 6;
  y_low[17] = 0;
  EXPECT_THROW(ordered_logistic_glm_lpmf(y_high, p.x, beta, cuts), std::domain_error);
  EXPECT_THROW(ordered_logistic_glm_lpmf(y_low, p.x, beta, cuts), std::domain_error);
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, ordered_logistic_with_more_cut_points_than_the_fused_kernel_holds) {
  // the reference takes any number of classes (ordered_logistic_glm_lpmf.hpp:69-102); beyond 1024 cut-points the device
  // evaluates location / rows / gradient as separate passes
  using stan::math::ordered_logistic_glm_lpmf;
  using stan::math::ordered_logistic_glm_lpmf_reference_cpu;
  Problem p = make_problem(4001, 19, 12);
  const int ncuts = 1500;
  VectorXd cuts_d(ncuts);
  for (int j = 0; j < ncuts; ++j) cuts_d[j] = -3.0 + 6.0 * j / ncuts + 0.001 * ((j * 7) % 3);
  vector_v cuts = to_v(cuts_d);
  vector_v beta = to_v(p.beta);
  std::vector<int> y(p.y_pois.size());
  for (size_t i = 0; i < y.size(); ++i) y[i] = 1 + static_cast<int>((i * 131 + p.y_pois[i]) % (ncuts + 1));
  STAN_CUDA_PARITY("ordered_logistic, 1500 cut-points", ordered_logistic_glm_lpmf(y, p.x, beta, cuts),
                   ordered_logistic_glm_lpmf_reference_cpu(y, p.x, beta, cuts), adj(beta), adj(cuts));
  STAN_CUDA_PARITY("ordered_logistic, 1500 data cut-points", ordered_logistic_glm_lpmf(y, p.x, beta, cuts_d),
                   ordered_logistic_glm_lpmf_reference_cpu(y, p.x, beta, cuts_d), adj(beta));
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, categorical_logit) {
  using stan::math::categorical_logit_glm_lpmf;
  using stan::math::categorical_logit_glm_lpmf_reference_cpu;
  typedef Eigen::Matrix<var, Dynamic, Dynamic> matrix_v;
  Problem p = make_problem(4001, 29, 12);
  const int C = 5;
  std::mt19937_64 rng(77);
  std::normal_distribution<double> nd(0.0, 0.3);
  MatrixXd beta_d(29, C);
  VectorXd alpha_d(C);
  for (int q = 0; q < beta_d.size(); ++q) beta_d(q) = nd(rng);
  for (int c = 0; c < C; ++c) alpha_d[c] = nd(rng);
  std::vector<int> y(p.y_pois.size());
  for (size_t i = 0; i < y.size(); ++i) y[i] = 1 + (p.y_pois[i] + static_cast<int>(i)) % C;
  matrix_v beta = beta_d.cast<var>();
  vector_v alpha = to_v(alpha_d);
  auto adj_m = [](const matrix_v& m) {
    VectorXd g(m.size());
    for (int q = 0; q < m.size(); ++q) g[q] = m(q).adj();
    return g;
  };
  STAN_CUDA_PARITY("categorical_logit", categorical_logit_glm_lpmf(y, p.x, alpha, beta),
                   categorical_logit_glm_lpmf_reference_cpu(y, p.x, alpha, beta), adj_m(beta), adj(alpha));
  STAN_CUDA_PARITY("categorical_logit propto, data alpha", categorical_logit_glm_lpmf<true>(y, p.x, alpha_d, beta),
                   categorical_logit_glm_lpmf_reference_cpu<true>(y, p.x, alpha_d, beta), adj_m(beta));
  STAN_CUDA_PARITY("categorical_logit data beta", categorical_logit_glm_lpmf(y, p.x, alpha, beta_d),
                   categorical_logit_glm_lpmf_reference_cpu(y, p.x, alpha, beta_d), adj(alpha));
  EXPECT_NEAR(categorical_logit_glm_lpmf(y, p.x, alpha_d, beta_d),
              categorical_logit_glm_lpmf_reference_cpu(y, p.x, alpha_d, beta_d), 1e-8);
  EXPECT_EQ(categorical_logit_glm_lpmf<true>(y, p.x, alpha_d, beta_d), 0.0);
  std::vector<int> y_high = y, y_low = y;
  y_high[17] = C + 1;
  y_low[17] = 0;
  EXPECT_THROW(categorical_logit_glm_lpmf(y_high, p.x, alpha, beta), std::domain_error);
  EXPECT_THROW(categorical_logit_glm_lpmf(y_low, p.x, alpha, beta), std::domain_error);
}

TEST(CudaGlmVariants, categorical_logit_beyond_the_fused_kernels) {
  // more than 16 classes, classes * attributes > 4096, more than 512 attributes: the general path
  // (csrc/glm_categorical_wide.cu); round 1 raised std::system_error for these shapes
  using stan::math::categorical_logit_glm_lpmf;
  using stan::math::categorical_logit_glm_lpmf_reference_cpu;
  typedef Eigen::Matrix<var, Dynamic, Dynamic> matrix_v;
  const int shapes[3][3] = {{2001, 29, 17}, {1501, 300, 16}, {701, 600, 4}};
  for (int q = 0; q < 3; ++q) {
    const int n = shapes[q][0], k = shapes[q][1], C = shapes[q][2];
    Problem p = make_problem(n, k, 50 + q);
    std::mt19937_64 rng(78 + q);
    std::normal_distribution<double> nd(0.0, 0.6 / std::sqrt(double(k)));
    MatrixXd beta_d(k, C);
    VectorXd alpha_d(C);
    for (int i = 0; i < beta_d.size(); ++i) beta_d(i) = nd(rng);
    for (int c = 0; c < C; ++c) alpha_d[c] = nd(rng);
    std::vector<int> y(n);
    for (int i = 0; i < n; ++i) y[i] = 1 + (p.y_pois[i] + i) % C;
    matrix_v beta = beta_d.cast<var>();
    vector_v alpha = to_v(alpha_d);
    auto adj_m = [](const matrix_v& m) {
      VectorXd g(m.size());
      for (int i = 0; i < m.size(); ++i) g[i] = m(i).adj();
      return g;
    };
    STAN_CUDA_PARITY("categorical_logit general path", categorical_logit_glm_lpmf(y, p.x, alpha, beta),
                     categorical_logit_glm_lpmf_reference_cpu(y, p.x, alpha, beta), adj_m(beta), adj(alpha));
    stan::math::recover_memory();
  }
}

TEST(CudaGlmVariants, zero_copy_data_view) {
  Problem p = make_problem(4001, 37, 9);
  vector_v beta = to_v(p.beta);
  var alpha = 0.3, sigma = 1.3;
  stan::math::cuda::data_matrix_view xv(p.x.data(), p.x.rows(), p.x.cols());
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("bernoulli view", bernoulli_logit_glm_lpmf(p.y_bern, xv, alpha, beta),
                   bernoulli_logit_glm_lpmf_reference_cpu(p.y_bern, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("poisson view propto", poisson_log_glm_lpmf<true>(p.y_pois, xv, alpha, beta),
                   poisson_log_glm_lpmf_reference_cpu<true>(p.y_pois, p.x, alpha, beta), adj(beta),
                   VectorXd::Constant(1, alpha.adj()));
  STAN_CUDA_PARITY("normal view", normal_id_glm_lpdf(p.y_norm, xv, alpha, beta, sigma),
                   normal_id_glm_lpdf_reference_cpu(p.y_norm, p.x, alpha, beta, sigma), adj(beta),
                   VectorXd::Constant(1, alpha.adj()), VectorXd::Constant(1, sigma.adj()));
  stan::math::recover_memory();
}

TEST(CudaGlmVariants, zero_instances) {
  MatrixXd x(0, 2);
  std::vector<int> y;
  VectorXd yn(0);
  vector_v beta(2);
  beta << 0.3, 2;
  var alpha = 0.3, sigma = 2.0;
  EXPECT_EQ(0, stan::math::bernoulli_logit_glm_lpmf(y, x, alpha, beta).val());
  EXPECT_EQ(0, stan::math::poisson_log_glm_lpmf(y, x, alpha, beta).val());
  EXPECT_EQ(0, stan::math::normal_id_glm_lpdf(yn, x, alpha, beta, sigma).val());
  stan::math::recover_memory();
}

TEST(CudaGlmErrors, same_exceptions_as_the_reference) {
  using stan::math::bernoulli_logit_glm_lpmf;
  using stan::math::normal_id_glm_lpdf;
  using stan::math::poisson_log_glm_lpmf;
  MatrixXd x(3, 2);
  x << -12, 46, -42, 24, 25, 27;
  x *= 0.01;
  std::vector<int> y{1, 0, 1}, y_size{1, 0, 1, 0}, y_val{1, 0, 23}, y_neg{1, 0, -1};
  VectorXd yn(3), beta(2), beta_size(3), beta_inf(2), sigma_v(3);
  yn << 0.1, 0.2, 0.3;
  beta << 0.3, 2;
  beta_size << 0.3, 2, 1;
  beta_inf << 0.3, INFINITY;
  double alpha = 0.3;
  EXPECT_THROW(bernoulli_logit_glm_lpmf(y_size, x, alpha, beta), std::invalid_argument);
  EXPECT_THROW(bernoulli_logit_glm_lpmf(y, x, alpha, beta_size), std::invalid_argument);
  EXPECT_THROW(bernoulli_logit_glm_lpmf(y_val, x, alpha, beta), std::domain_error);
  EXPECT_THROW(poisson_log_glm_lpmf(y_neg, x, alpha, beta), std::domain_error);
  EXPECT_THROW(poisson_log_glm_lpmf(y, x, alpha, beta_inf), std::domain_error);
  EXPECT_THROW(poisson_log_glm_lpmf(y, x, NAN, beta), std::domain_error);
  EXPECT_THROW(normal_id_glm_lpdf(yn, x, alpha, beta, 0.0), std::domain_error);
  using stan::math::neg_binomial_2_log_glm_lpmf;
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y_size, x, alpha, beta, 2.0), std::invalid_argument);
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y_neg, x, alpha, beta, 2.0), std::domain_error);
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y, x, alpha, beta_inf, 2.0), std::domain_error);
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y, x, alpha, beta, 0.0), std::domain_error);
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y, x, NAN, beta, 2.0), std::domain_error);
  EXPECT_THROW(normal_id_glm_lpdf(yn, x, alpha, beta, INFINITY), std::domain_error);
  MatrixXd x_bad = x;
  x_bad(0, 1) = NAN;
  EXPECT_THROW(bernoulli_logit_glm_lpmf(y, x_bad, alpha, beta), std::domain_error);
  EXPECT_THROW(poisson_log_glm_lpmf(y, x_bad, alpha, beta), std::domain_error);
  EXPECT_THROW(normal_id_glm_lpdf(yn, x_bad, alpha, beta, 1.0), std::domain_error);
  EXPECT_THROW(neg_binomial_2_log_glm_lpmf(y, x_bad, alpha, beta, 1.0), std::domain_error);
  // messages agree with the reference body on the lazy path
  std::string m_gpu, m_cpu;
  try {
    poisson_log_glm_lpmf(y, x_bad, alpha, beta);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    stan::math::poisson_log_glm_lpmf_reference_cpu(y, x_bad, alpha, beta);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
}

TEST(CudaNormalLpdf, vectorised_matches_reference_loop) {
  const int n = 100000;
  Problem p = make_problem(n, 1, 5);
  vector_v mu = to_v(0.1 * p.y_norm), sigma_v = to_v(p.sigma_vec), y_v = to_v(p.y_norm);
  var sigma = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("normal_lpdf vec mu", normal_lpdf(p.y_norm, mu, sigma),
                   normal_lpdf_reference_cpu(p.y_norm, mu, sigma), adj(mu), VectorXd::Constant(1, sigma.adj()));
  STAN_CUDA_PARITY("normal_lpdf propto", normal_lpdf<true>(p.y_norm, mu, sigma),
                   normal_lpdf_reference_cpu<true>(p.y_norm, mu, sigma), adj(mu),
                   VectorXd::Constant(1, sigma.adj()));
  STAN_CUDA_PARITY("normal_lpdf all var vectors", normal_lpdf(y_v, mu, sigma_v),
                   normal_lpdf_reference_cpu(y_v, mu, sigma_v), adj(mu), adj(sigma_v), adj(y_v));
  var mu0 = 0.25;
  STAN_CUDA_PARITY("normal_lpdf scalar mu", normal_lpdf(p.y_norm, mu0, sigma),
                   normal_lpdf_reference_cpu(p.y_norm, mu0, sigma), VectorXd::Constant(1, mu0.adj()),
                   VectorXd::Constant(1, sigma.adj()));
  // known answers of test/prob/normal/normal_test.hpp:11-40 through the vectorised entry
  std::vector<double> y1{-3.5}, m1{1.9}, s1{7.2};
  EXPECT_NEAR(-3.174269559226682080322, normal_lpdf(y1, m1, s1), 1e-14);
  stan::math::recover_memory();
}

TEST(CudaNormalLpdf, bulk_adjoint_scatter_matches_the_reference_node) {
  // N above the threshold of the parallel chain() (stan/math/cuda/arena_partials.hpp, bulk_gradients_vari): operands
  // built one after the other (strictly ascending addresses: scattered in parallel), the same vector used for two
  // arguments (edges run one after the other), and a vector that repeats ONE var (not ascending: serial loop)
  const int n = 300001;
  Problem p = make_problem(n, 1, 6);
  vector_v mu = to_v(0.1 * p.y_norm), y_v = to_v(p.y_norm);
  var sigma = 1.3;
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("normal_lpdf bulk", normal_lpdf(p.y_norm, mu, sigma), normal_lpdf_reference_cpu(p.y_norm, mu, sigma),
                   adj(mu), VectorXd::Constant(1, sigma.adj()));
  vector_v pos = to_v(p.sigma_vec);  // positive: usable as location and as scale
  STAN_CUDA_PARITY("normal_lpdf bulk, one vector in two edges", normal_lpdf(y_v, pos, pos), normal_lpdf_reference_cpu(y_v, pos, pos),
                   adj(pos), adj(y_v));
  var m = 0.25;
  vector_v repeated(n);
  for (int i = 0; i < n; ++i) repeated[i] = m;
  STAN_CUDA_PARITY("normal_lpdf one var repeated", normal_lpdf(p.y_norm, repeated, sigma),
                   normal_lpdf_reference_cpu(p.y_norm, repeated, sigma), VectorXd::Constant(1, m.adj()),
                   VectorXd::Constant(1, sigma.adj()));
  stan::math::recover_memory();
}

TEST(CudaNormalLpdf, same_exceptions_as_the_reference) {
  using stan::math::normal_lpdf;
  const int n = 1000;
  VectorXd y = VectorXd::LinSpaced(n, -1, 1), mu_d = VectorXd::Constant(n, 0.1);
  vector_v mu = to_v(mu_d);
  var sigma = 1.5;
  VectorXd y_nan = y;
  y_nan[777] = NAN;
  EXPECT_THROW(normal_lpdf(y_nan, mu, sigma), std::domain_error);
  vector_v mu_inf = to_v(mu_d);
  mu_inf[3] = INFINITY;
  EXPECT_THROW(normal_lpdf(y, mu_inf, sigma), std::domain_error);
  var sigma0 = 0.0, sigma_neg = -1.0;
  EXPECT_THROW(normal_lpdf(y, mu, sigma0), std::domain_error);
  EXPECT_THROW(normal_lpdf(y, mu, sigma_neg), std::domain_error);
  vector_v sigma_v = to_v(VectorXd::Constant(n, 2.0));
  sigma_v[999] = -2.0;
  EXPECT_THROW(normal_lpdf(y, mu, sigma_v), std::domain_error);
  vector_v mu_short = to_v(VectorXd::Constant(n - 1, 0.1));
  EXPECT_THROW(normal_lpdf(y, mu_short, sigma), std::invalid_argument);
  // same message as the reference loop
  std::string m_gpu, m_cpu;
  try {
    normal_lpdf(y_nan, mu, sigma);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    stan::math::normal_lpdf_reference_cpu<false>(y_nan, mu, sigma);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  stan::math::recover_memory();
}

// --- vectorised elementwise densities -------------------------------------------
TEST(CudaIntDensity, bernoulli_logit_and_poisson_log_match_reference_loop) {
  const int n = 100000;
  Problem p = make_problem(n, 1, 9);
  VectorXd th_d = 3.0 * p.y_norm, al_d = 0.5 * p.y_norm;
  th_d[5] = 25.0;   // the cutoff branches of bernoulli_logit_lpmf.hpp:63-79
  th_d[6] = -25.0;
  vector_v theta = to_v(th_d), alpha = to_v(al_d);
  std::vector<var> theta_sv(theta.data(), theta.data() + n);
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("bernoulli_logit_lpmf", bernoulli_logit_lpmf(p.y_bern, theta),
                   bernoulli_logit_lpmf_reference_cpu<false>(p.y_bern, theta), adj(theta));
  STAN_CUDA_PARITY("bernoulli_logit_lpmf propto", bernoulli_logit_lpmf<true>(p.y_bern, theta),
                   bernoulli_logit_lpmf_reference_cpu<true>(p.y_bern, theta), adj(theta));
  STAN_CUDA_PARITY("bernoulli_logit_lpmf scalar n", bernoulli_logit_lpmf(1, theta),
                   bernoulli_logit_lpmf_reference_cpu<false>(1, theta), adj(theta));
  STAN_CUDA_PARITY("poisson_log_lpmf", poisson_log_lpmf(p.y_pois, alpha),
                   poisson_log_lpmf_reference_cpu<false>(p.y_pois, alpha), adj(alpha));
  STAN_CUDA_PARITY("poisson_log_lpmf propto", poisson_log_lpmf<true>(p.y_pois, alpha),
                   poisson_log_lpmf_reference_cpu<true>(p.y_pois, alpha), adj(alpha));
  {  // std::vector<var> operand
    var lp = bernoulli_logit_lpmf(p.y_bern, theta_sv);
    var lp_ref = bernoulli_logit_lpmf_reference_cpu<false>(p.y_bern, theta);
    expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "std::vector<var> theta");
  }
  // double operands: value only, double return type
  std::vector<double> th_sd(th_d.data(), th_d.data() + n);
  const double v = bernoulli_logit_lpmf(p.y_bern, th_sd);
  expect_rel(v, bernoulli_logit_lpmf_reference_cpu<false>(p.y_bern, th_sd), std::fabs(v), "double theta");
  const double w = poisson_log_lpmf(p.y_pois, al_d);
  expect_rel(w, poisson_log_lpmf_reference_cpu<false>(p.y_pois, al_d), std::fabs(w), "double alpha");
  EXPECT_EQ(0.0, poisson_log_lpmf<true>(p.y_pois, al_d));  // propto with a constant operand
  // known answers of test/prob/bernoulli/bernoulli_logit_test.hpp:11-48 and test/prob/poisson/poisson_log_test.hpp:11-31
  std::vector<int> n1{1}, n17{17};
  std::vector<double> t1{std::log(0.25 / 0.75)}, a13{std::log(13.0)};
  EXPECT_NEAR(std::log(0.25), bernoulli_logit_lpmf(n1, t1), 1e-14);
  EXPECT_NEAR(-2.900934373290765311282, poisson_log_lpmf(n17, a13), 1e-13);
  stan::math::recover_memory();
}

TEST(CudaIntDensity, same_exceptions_and_log_zero_as_the_reference) {
  using namespace stan::math;  // NOLINT
  const int n = 2000;
  std::vector<int> y(n, 1), y_bad = y, y_neg = y;
  y_bad[1500] = 2;
  y_neg[3] = -1;
  vector_v theta = to_v(VectorXd::LinSpaced(n, -2, 2));
  vector_v theta_nan = to_v(VectorXd::LinSpaced(n, -2, 2));
  theta_nan[1999] = NAN;
  EXPECT_THROW(bernoulli_logit_lpmf(y_bad, theta), std::domain_error);
  EXPECT_THROW(bernoulli_logit_lpmf(y, theta_nan), std::domain_error);
  EXPECT_THROW(poisson_log_lpmf(y_neg, theta), std::domain_error);
  EXPECT_THROW(poisson_log_lpmf(y, theta_nan), std::domain_error);
  vector_v theta_short = to_v(VectorXd::LinSpaced(n - 1, -2, 2));
  EXPECT_THROW(bernoulli_logit_lpmf(y, theta_short), std::invalid_argument);
  EXPECT_THROW(poisson_log_lpmf(y, theta_short), std::invalid_argument);
  std::string m_gpu, m_cpu;
  try {
    bernoulli_logit_lpmf(y_bad, theta);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    bernoulli_logit_lpmf_reference_cpu<false>(y_bad, theta);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // LOG_ZERO early returns of poisson_log_lpmf.hpp:51-62
  vector_v a_inf = to_v(VectorXd::Constant(n, 0.1)), a_minf = to_v(VectorXd::Constant(n, 0.1));
  a_inf[7] = INFINITY;
  a_minf[7] = -INFINITY;
  EXPECT_EQ(LOG_ZERO, poisson_log_lpmf(y, a_inf).val());
  EXPECT_EQ(LOG_ZERO, poisson_log_lpmf(y, a_minf).val());
  std::vector<int> y0 = y;
  y0[7] = 0;  // -inf next to a zero count contributes nothing (:77)
  var lp = poisson_log_lpmf(y0, a_minf);
  var lp_ref = poisson_log_lpmf_reference_cpu<false>(y0, a_minf);
  EXPECT_TRUE(std::isfinite(lp.val()));
  expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "-inf with zero count");
  std::vector<int> none;
  vector_v empty(0);
  EXPECT_EQ(0.0, bernoulli_logit_lpmf(none, empty).val());
  stan::math::recover_memory();
}

TEST(CudaIntDensity, poisson_and_bernoulli_match_reference_loop) {
  const int n = 100000;
  Problem p = make_problem(n, 1, 10);
  VectorXd lam_d(n), th_d(n);
  for (int i = 0; i < n; ++i) {
    lam_d[i] = 0.05 + std::exp(0.7 * p.y_norm[i]);                    // rates
    th_d[i] = 1.0 / (1.0 + std::exp(-1.5 * p.y_norm[i]));             // probabilities
    th_d[i] = std::min(std::max(th_d[i], 1e-6), 1.0 - 1e-6);
  }
  lam_d[5] = 1e-300;  // the corners
  lam_d[6] = 1e6;
  th_d[5] = 1e-300;
  th_d[6] = 1.0 - 1e-16;
  std::vector<int> y_pois = p.y_pois, y_bern = p.y_bern;
  y_pois[5] = 0;
  y_bern[5] = 0;
  y_bern[6] = 1;
  vector_v lambda = to_v(lam_d), theta = to_v(th_d);
  using namespace stan::math;  // NOLINT
  STAN_CUDA_PARITY("poisson_lpmf", poisson_lpmf(y_pois, lambda), poisson_lpmf_reference_cpu<false>(y_pois, lambda), adj(lambda));
  STAN_CUDA_PARITY("poisson_lpmf propto", poisson_lpmf<true>(y_pois, lambda),
                   poisson_lpmf_reference_cpu<true>(y_pois, lambda), adj(lambda));
  STAN_CUDA_PARITY("poisson_lpmf scalar n", poisson_lpmf(2, lambda), poisson_lpmf_reference_cpu<false>(2, lambda), adj(lambda));
  STAN_CUDA_PARITY("bernoulli_lpmf", bernoulli_lpmf(y_bern, theta), bernoulli_lpmf_reference_cpu<false>(y_bern, theta), adj(theta));
  STAN_CUDA_PARITY("bernoulli_lpmf propto", bernoulli_lpmf<true>(y_bern, theta),
                   bernoulli_lpmf_reference_cpu<true>(y_bern, theta), adj(theta));
  STAN_CUDA_PARITY("bernoulli_lpmf scalar n", bernoulli_lpmf(1, theta), bernoulli_lpmf_reference_cpu<false>(1, theta), adj(theta));
  // the deprecated aliases forward to the routed functions (prim/scal/prob/poisson_log.hpp, bernoulli_log.hpp)
  STAN_CUDA_PARITY("poisson_log alias", poisson_log(y_pois, lambda), poisson_lpmf_reference_cpu<false>(y_pois, lambda), adj(lambda));
  // double operands: value only, double return type
  const double v = poisson_lpmf(y_pois, lam_d);
  expect_rel(v, poisson_lpmf_reference_cpu<false>(y_pois, lam_d), std::fabs(v), "double lambda");
  const double w = bernoulli_lpmf(y_bern, th_d);
  expect_rel(w, bernoulli_lpmf_reference_cpu<false>(y_bern, th_d), std::fabs(w), "double theta");
  EXPECT_EQ(0.0, poisson_lpmf<true>(y_pois, lam_d));  // propto with a constant operand
  // a container of length one takes bernoulli_lpmf's closed form (:58-91)
  vector_v one = to_v(VectorXd::Constant(1, 0.3));
  STAN_CUDA_PARITY("bernoulli_lpmf, one probability", bernoulli_lpmf(1, one), bernoulli_lpmf_reference_cpu<false>(1, one), adj(one));
  // known answers of test/prob/poisson/poisson_test.hpp:11-30 and test/prob/bernoulli/bernoulli_test.hpp:11-35
  std::vector<int> n17{17}, n1{1};
  std::vector<double> l13{13.0}, t25{0.25};
  EXPECT_NEAR(-2.900934373290765311282, poisson_lpmf(n17, l13), 1e-13);
  EXPECT_NEAR(std::log(0.25), bernoulli_lpmf(n1, t25), 1e-14);
  stan::math::recover_memory();
}

TEST(CudaIntDensity, poisson_and_bernoulli_same_exceptions_and_log_zero_as_the_reference) {
  using namespace stan::math;  // NOLINT
  const int n = 2000;
  std::vector<int> y(n, 1), y_two = y, y_neg = y;
  y_two[1500] = 2;
  y_neg[3] = -1;
  vector_v lambda = to_v(VectorXd::LinSpaced(n, 0.5, 3.0)), theta = to_v(VectorXd::LinSpaced(n, 0.05, 0.95));
  vector_v lam_nan = to_v(VectorXd::LinSpaced(n, 0.5, 3.0)), lam_neg = to_v(VectorXd::LinSpaced(n, 0.5, 3.0));
  lam_nan[1999] = NAN;
  lam_neg[17] = -1e-5;
  vector_v th_hi = to_v(VectorXd::LinSpaced(n, 0.05, 0.95)), th_lo = to_v(VectorXd::LinSpaced(n, 0.05, 0.95)),
           th_inf = to_v(VectorXd::LinSpaced(n, 0.05, 0.95));
  th_hi[9] = 1.001;
  th_lo[9] = -0.001;
  th_inf[9] = INFINITY;
  EXPECT_THROW(poisson_lpmf(y_neg, lambda), std::domain_error);
  EXPECT_THROW(poisson_lpmf(y, lam_nan), std::domain_error);
  EXPECT_THROW(poisson_lpmf(y, lam_neg), std::domain_error);
  EXPECT_THROW(bernoulli_lpmf(y_two, theta), std::domain_error);
  EXPECT_THROW(bernoulli_lpmf(y_neg, theta), std::domain_error);
  EXPECT_THROW(bernoulli_lpmf(y, th_hi), std::domain_error);
  EXPECT_THROW(bernoulli_lpmf(y, th_lo), std::domain_error);
  EXPECT_THROW(bernoulli_lpmf(y, th_inf), std::domain_error);
  vector_v shorter = to_v(VectorXd::LinSpaced(n - 1, 0.1, 0.9));
  EXPECT_THROW(poisson_lpmf(y, shorter), std::invalid_argument);
  EXPECT_THROW(bernoulli_lpmf(y, shorter), std::invalid_argument);
  std::string m_gpu, m_cpu;
  try {
    poisson_lpmf(y, lam_neg);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    poisson_lpmf_reference_cpu<false>(y, lam_neg);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  m_gpu.clear();
  m_cpu.clear();
  try {
    bernoulli_lpmf(y, th_hi);
  } catch (const std::domain_error& e) {
    m_gpu = e.what();
  }
  try {
    bernoulli_lpmf_reference_cpu<false>(y, th_hi);
  } catch (const std::domain_error& e) {
    m_cpu = e.what();
  }
  EXPECT_EQ(m_gpu, m_cpu);
  EXPECT_FALSE(m_gpu.empty());
  // LOG_ZERO early returns of poisson_lpmf.hpp:45-54
  vector_v l_inf = to_v(VectorXd::Constant(n, 1.1)), l_zero = to_v(VectorXd::Constant(n, 1.1));
  l_inf[7] = INFINITY;
  l_zero[7] = 0.0;
  EXPECT_EQ(LOG_ZERO, poisson_lpmf(y, l_inf).val());
  EXPECT_EQ(LOG_ZERO, poisson_lpmf(y, l_zero).val());
  std::vector<int> y0 = y;
  y0[7] = 0;  // a zero rate next to a zero count contributes nothing (:59)
  var lp = poisson_lpmf(y0, l_zero);
  var lp_ref = poisson_lpmf_reference_cpu<false>(y0, l_zero);
  EXPECT_TRUE(std::isfinite(lp.val()));
  expect_rel(lp.val(), lp_ref.val(), std::fabs(lp_ref.val()), "zero rate with zero count");
  std::vector<int> none;
  vector_v empty(0);
  EXPECT_EQ(0.0, poisson_lpmf(none, empty).val());
  EXPECT_EQ(0.0, bernoulli_lpmf(none, empty).val());
  stan::math::recover_memory();
}

// --- map_rect: shards dispatched to devices ------------------------------------------------------
struct glm_shard {
  template <typename T1, typename T2>
  Eigen::Matrix<typename stan::return_type<T1, T2>::type, Dynamic, 1> operator()(
      const Matrix<T1, Dynamic, 1>& shared, const Matrix<T2, Dynamic, 1>& job, const std::vector<double>& x_r,
      const std::vector<int>& x_i, std::ostream* msgs = 0) const {
    typedef typename stan::return_type<T1, T2>::type T;
    const int rows = x_i[0], k = x_i[1];
    MatrixXd x = Eigen::Map<const MatrixXd>(x_r.data(), rows, k);
    std::vector<int> y(x_i.begin() + 2, x_i.end());
    Matrix<T, Dynamic, 1> res(1);
    res(0) = stan::math::bernoulli_logit_glm_lpmf(y, x, job(0), shared);
    return res;
  }
};

TEST(CudaMapRect, hierarchical_logistic_matches_direct_sum) {
  const int J = 12, rows = 3000, k = 32;
  Problem p = make_problem(J * rows, k, 77);
  std::vector<std::vector<double>> x_r(J);
  std::vector<std::vector<int>> x_i(J);
  for (int j = 0; j < J; ++j) {
    MatrixXd xs = p.x.middleRows(j * rows, rows);
    x_r[j].assign(xs.data(), xs.data() + xs.size());
    x_i[j] = {rows, k};
    x_i[j].insert(x_i[j].end(), p.y_bern.begin() + j * rows, p.y_bern.begin() + (j + 1) * rows);
  }
  for (int rep = 0; rep < 2; ++rep) {  // second pass hits the (call_id, job) cache
    vector_v beta = to_v(p.beta);
    std::vector<vector_v> job(J, vector_v(1));
    for (int j = 0; j < J; ++j) job[j][0] = 0.01 * j;
    vector_v out = stan::math::map_rect<3, glm_shard>(beta, job, x_r, x_i);
    var total = stan::math::sum(out);
    total.grad();
    const double v = total.val();
    VectorXd g = adj(beta);
    VectorXd ga(J);
    for (int j = 0; j < J; ++j) ga[j] = job[j][0].adj();
    stan::math::set_zero_all_adjoints();
    var ref = 0;
    for (int j = 0; j < J; ++j) {
      MatrixXd xs = p.x.middleRows(j * rows, rows);
      std::vector<int> ys(p.y_bern.begin() + j * rows, p.y_bern.begin() + (j + 1) * rows);
      ref += stan::math::bernoulli_logit_glm_lpmf_reference_cpu(ys, xs, job[j][0], beta);
    }
    ref.grad();
    expect_rel(v, ref.val(), std::fabs(ref.val()), "map_rect value");
    expect_vec(g, adj(beta), "map_rect shared grad");
    VectorXd ga_ref(J);
    for (int j = 0; j < J; ++j) ga_ref[j] = job[j][0].adj();
    expect_vec(ga, ga_ref, "map_rect job grad");
    stan::math::recover_memory();
  }
  stan_cuda_stats_t st;
  stan_cuda_stats(&st);
  EXPECT_GE(st.cache_hits, static_cast<uint64_t>(J));
}
#endif
