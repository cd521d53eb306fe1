// The reference's distribution-test fixture (test/prob/test_fixture_distr.hpp) and its test classes for
// poisson and bernoulli (test/prob/poisson/poisson_test.hpp, test/prob/bernoulli/bernoulli_test.hpp: they call the
// deprecated aliases poisson_log(n, lambda) / bernoulli_log(n, theta), which forward to poisson_lpmf / bernoulli_lpmf),
// all three included UNMODIFIED from the reference tree, instantiated -- the way the
// reference's generator (test/prob/generate_tests.cpp:235-340) writes its "_v_" files -- for the argument
// kinds that the overlay routes to the B200 kernel (container operand of var / double) plus one scalar kind
// that stays on the reference loop.  Known answers, invalid values, propto behaviour, finite-difference and
// hand-written-function gradients, vector-vs-repeated-scalar consistency: the fixture's own checks.
#ifdef STAN_CUDA
#include <gtest/gtest.h>
#include <boost/mpl/vector.hpp>
#include <test/prob/test_fixture_distr.hpp>
#include <test/prob/bernoulli/bernoulli_test.hpp>
#include <test/prob/poisson/poisson_test.hpp>
#include <vector>

typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_var;
typedef Eigen::Matrix<var, 1, Eigen::Dynamic> row_vector_var;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_dbl;

typedef boost::mpl::vector<int, std::vector<var>, empty, empty, empty, empty> type_v_0;
typedef boost::mpl::vector<std::vector<int>, std::vector<var>, empty, empty, empty, empty> type_v_1;
typedef boost::mpl::vector<std::vector<int>, vector_var, empty, empty, empty, empty> type_v_2;
typedef boost::mpl::vector<std::vector<int>, row_vector_var, empty, empty, empty, empty> type_v_3;
typedef boost::mpl::vector<Eigen::Matrix<int, Eigen::Dynamic, 1>, vector_var, empty, empty, empty, empty> type_v_4;
typedef boost::mpl::vector<std::vector<int>, std::vector<double>, empty, empty, empty, empty> type_v_5;
typedef boost::mpl::vector<int, vector_dbl, empty, empty, empty, empty> type_v_6;
typedef boost::mpl::vector<std::vector<int>, var, empty, empty, empty, empty> type_v_7;  // reference loop

#define STAN_CUDA_DISTR(TEST_CLASS, N)                                                      \
  typedef boost::mpl::vector<TEST_CLASS, type_v_##N> TEST_CLASS##_v_##N;                    \
  INSTANTIATE_TYPED_TEST_CASE_P(TEST_CLASS##_v_##N, AgradDistributionTestFixture, TEST_CLASS##_v_##N);

STAN_CUDA_DISTR(AgradDistributionsBernoulli, 0)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 1)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 2)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 3)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 4)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 5)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 6)
STAN_CUDA_DISTR(AgradDistributionsBernoulli, 7)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 0)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 1)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 2)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 3)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 4)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 5)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 6)
STAN_CUDA_DISTR(AgradDistributionsPoisson, 7)
#endif
