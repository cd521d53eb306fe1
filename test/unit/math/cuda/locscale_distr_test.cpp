// The reference's distribution-test fixture (test/prob/test_fixture_distr.hpp) and its cauchy and lognormal test
// classes and those of gamma, exponential, beta and weibull (test/prob/{cauchy,lognormal,gamma,exponential,beta,weibull}/*_test.hpp: known answers,
// invalid values),
// all included UNMODIFIED from the reference tree and instantiated -- the way the reference's generator
// (test/prob/generate_tests.cpp:235-340) writes its "_v_" files -- for argument kinds the overlay routes to the
// B200 kernel (at least one container operand of double / var) plus one all-scalar kind that stays on the
// reference loop.  Known answers, invalid values, propto behaviour, finite-difference and hand-written-function
// gradients, vector-vs-repeated-scalar consistency: the fixture's own checks.
#ifdef STAN_CUDA
#include <gtest/gtest.h>
#include <boost/mpl/vector.hpp>
#include <test/prob/test_fixture_distr.hpp>
#include <test/prob/cauchy/cauchy_test.hpp>
#include <test/prob/lognormal/lognormal_test.hpp>
#include <test/prob/gamma/gamma_test.hpp>
#include <test/prob/exponential/exponential_test.hpp>
#include <test/prob/beta/beta_test.hpp>
#include <test/prob/weibull/weibull_test.hpp>
#include <test/prob/inv_gamma/inv_gamma_test.hpp>
#include <vector>

typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_var;
typedef Eigen::Matrix<var, 1, Eigen::Dynamic> row_vector_var;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_dbl;

// the model statement's kinds: double y vector, var vector mu, var scalar sigma
typedef boost::mpl::vector<std::vector<double>, std::vector<var>, var, empty, empty, empty> type_v_0;
typedef boost::mpl::vector<vector_dbl, vector_var, var, empty, empty, empty> type_v_1;
typedef boost::mpl::vector<std::vector<var>, std::vector<var>, std::vector<var>, empty, empty, empty> type_v_2;
typedef boost::mpl::vector<vector_var, double, vector_var, empty, empty, empty> type_v_3;
typedef boost::mpl::vector<double, row_vector_var, double, empty, empty, empty> type_v_4;
typedef boost::mpl::vector<std::vector<double>, var, std::vector<var>, empty, empty, empty> type_v_5;
typedef boost::mpl::vector<var, vector_dbl, var, empty, empty, empty> type_v_6;
typedef boost::mpl::vector<var, var, var, empty, empty, empty> type_v_7;  // all scalars: reference loop

#define STAN_CUDA_DISTR(TEST_CLASS, N)                                                      \
  typedef boost::mpl::vector<TEST_CLASS, type_v_##N> TEST_CLASS##_v_##N;                    \
  INSTANTIATE_TYPED_TEST_CASE_P(TEST_CLASS##_v_##N, AgradDistributionTestFixture, TEST_CLASS##_v_##N);

STAN_CUDA_DISTR(AgradDistributionsCauchy, 0)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 1)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 2)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 3)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 4)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 5)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 6)
STAN_CUDA_DISTR(AgradDistributionsCauchy, 7)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 0)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 1)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 2)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 3)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 4)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 5)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 6)
STAN_CUDA_DISTR(AgradDistributionsLognormal, 7)
STAN_CUDA_DISTR(AgradDistributionsGamma, 0)
STAN_CUDA_DISTR(AgradDistributionsGamma, 1)
STAN_CUDA_DISTR(AgradDistributionsGamma, 2)
STAN_CUDA_DISTR(AgradDistributionsGamma, 3)
STAN_CUDA_DISTR(AgradDistributionsGamma, 4)
STAN_CUDA_DISTR(AgradDistributionsGamma, 5)
STAN_CUDA_DISTR(AgradDistributionsGamma, 6)
STAN_CUDA_DISTR(AgradDistributionsGamma, 7)

STAN_CUDA_DISTR(AgradDistributionsBeta, 0)
STAN_CUDA_DISTR(AgradDistributionsBeta, 1)
STAN_CUDA_DISTR(AgradDistributionsBeta, 2)
STAN_CUDA_DISTR(AgradDistributionsBeta, 3)
STAN_CUDA_DISTR(AgradDistributionsBeta, 4)
STAN_CUDA_DISTR(AgradDistributionsBeta, 5)
STAN_CUDA_DISTR(AgradDistributionsBeta, 6)
STAN_CUDA_DISTR(AgradDistributionsBeta, 7)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 0)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 1)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 2)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 3)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 4)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 5)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 6)
STAN_CUDA_DISTR(AgradDistributionsWeibull, 7)

STAN_CUDA_DISTR(AgradDistributionsInvGamma, 0)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 1)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 2)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 3)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 4)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 5)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 6)
STAN_CUDA_DISTR(AgradDistributionsInvGamma, 7)

// exponential_lpdf(y | beta): two operands
typedef boost::mpl::vector<std::vector<double>, var, empty, empty, empty, empty> type_e_0;   // data y, var rate: the model statement
typedef boost::mpl::vector<vector_dbl, vector_var, empty, empty, empty, empty> type_e_1;
typedef boost::mpl::vector<std::vector<var>, std::vector<var>, empty, empty, empty, empty> type_e_2;
typedef boost::mpl::vector<vector_var, double, empty, empty, empty, empty> type_e_3;
typedef boost::mpl::vector<double, row_vector_var, empty, empty, empty, empty> type_e_4;
typedef boost::mpl::vector<var, var, empty, empty, empty, empty> type_e_5;  // all scalars: reference loop
#define STAN_CUDA_DISTR_E(TEST_CLASS, N)                                                    \
  typedef boost::mpl::vector<TEST_CLASS, type_e_##N> TEST_CLASS##_e_##N;                    \
  INSTANTIATE_TYPED_TEST_CASE_P(TEST_CLASS##_e_##N, AgradDistributionTestFixture, TEST_CLASS##_e_##N);
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 0)
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 1)
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 2)
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 3)
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 4)
STAN_CUDA_DISTR_E(AgradDistributionsExponential, 5)
#endif
