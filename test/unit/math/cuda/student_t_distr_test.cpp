// The reference's distribution-test fixture (test/prob/test_fixture_distr.hpp) and its student_t test class
// (test/prob/student_t/student_t_test.hpp:8-90 -- the golden vectors and invalid values the reference itself holds
// for student_t_lpdf), both included UNMODIFIED from the reference tree and instantiated -- the way the reference's
// generator (test/prob/generate_tests.cpp:235-340) writes its "_v_" files -- for argument kinds the overlay routes
// to the B200 kernel (at least one container operand of double / var) plus one all-scalar kind that stays on the
// reference loops.  Known answers, invalid values, propto behaviour, finite-difference and hand-written-function
// gradients, vector-vs-repeated-scalar consistency: the fixture's own checks.
#ifdef STAN_CUDA
#include <gtest/gtest.h>
#include <boost/mpl/vector.hpp>
#include <test/prob/test_fixture_distr.hpp>
#include <test/prob/student_t/student_t_test.hpp>
#include <vector>

typedef Eigen::Matrix<var, Eigen::Dynamic, 1> vector_var;
typedef Eigen::Matrix<var, 1, Eigen::Dynamic> row_vector_var;
typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_dbl;

typedef boost::mpl::vector<std::vector<double>, var, std::vector<var>, var, empty, empty> type_v_0;  // the model statement
typedef boost::mpl::vector<vector_dbl, var, vector_var, var, empty, empty> type_v_1;
typedef boost::mpl::vector<std::vector<var>, std::vector<var>, std::vector<var>, std::vector<var>, empty, empty> type_v_2;
typedef boost::mpl::vector<vector_var, double, vector_var, double, empty, empty> type_v_3;
typedef boost::mpl::vector<double, row_vector_var, double, var, empty, empty> type_v_4;
typedef boost::mpl::vector<std::vector<double>, std::vector<var>, double, std::vector<var>, empty, empty> type_v_5;
typedef boost::mpl::vector<var, vector_dbl, var, vector_dbl, empty, empty> type_v_6;
typedef boost::mpl::vector<var, var, var, var, empty, empty> type_v_7;  // all scalars: reference loops

#define STAN_CUDA_DISTR(TEST_CLASS, N)                                                      \
  typedef boost::mpl::vector<TEST_CLASS, type_v_##N> TEST_CLASS##_v_##N;                    \
  INSTANTIATE_TYPED_TEST_CASE_P(TEST_CLASS##_v_##N, AgradDistributionTestFixture, TEST_CLASS##_v_##N);

STAN_CUDA_DISTR(AgradDistributionsStudentT, 0)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 1)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 2)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 3)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 4)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 5)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 6)
STAN_CUDA_DISTR(AgradDistributionsStudentT, 7)
#endif
