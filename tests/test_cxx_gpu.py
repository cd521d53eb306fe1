"""GPU tests of the C++ host layer: gtest binaries built in the dev container by test/Makefile
(they embed the reference's header-only code, so they run where /root/reference is absent).

* cuda_glm_parity_test: the STAN_CUDA path through the reference's own C++ signatures (var and
  double arguments) against the unmodified reference body, 1e-10.
* cuda_int_density_distr_test: the reference's distribution-test fixture and its bernoulli_logit / poisson_log test
  classes (included unmodified) instantiated for the argument kinds routed to the device.
* ref_*: the reference's OWN unit tests, compiled unmodified with the overlay include path and
  -DSTAN_CUDA -- the drop-in contract."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CXX_DIR = os.path.join(ROOT, "build", "cxx")
BINARIES = ["cuda_glm_parity_test", "ref_bernoulli_logit_glm_lpmf_test", "ref_poisson_log_glm_lpmf_test",
            "ref_normal_id_glm_lpdf_test", "ref_neg_binomial_2_log_glm_lpmf_test", "ref_ordered_logistic_glm_lpmf_test", "ref_categorical_logit_glm_lpmf_test", "ref_operands_and_partials_test", "ref_precomputed_gradients_test",
            "ref_map_rect_concurrent_test", "cuda_int_density_distr_test", "ref_bernoulli_logit_log_test",
            "ref_poisson_log_log_test", "ref_bernoulli_logit_test", "ref_poisson_log_test",
            "cuda_normal_distr_test", "ref_map_rect_concurrent_threads_test", "ref_map_rect_prim_test",
            "ref_map_rect_concurrent_prim_test", "ref_normal_log_test", "ref_normal_test",
            "cuda_student_t_parity_test", "cuda_student_t_distr_test", "ref_student_t_log_test",
            "cuda_multiply_test", "ref_multiply_test", "ref_mix_multiply_test",
            "cuda_cholesky_test", "ref_cholesky_decompose_test", "ref_mix_cholesky_decompose_test",
            "cuda_tri_solve_test", "ref_mdivide_left_tri_test", "ref_mdivide_right_tri_test", "ref_mix_mdivide_left_tri_test",
            "cuda_gp_exp_quad_cov_test", "ref_gp_exp_quad_cov_test", "ref_rev_gp_exp_quad_cov_test",
            "ref_mix_mdivide_left_tri_low_test",
            "ref_multi_normal_cholesky_test", "ref_rev_multi_normal_cholesky_test",
            "cuda_int_density2_distr_test", "ref_poisson_test", "ref_bernoulli_test", "ref_bernoulli_log_test",
            "cuda_locscale_distr_test", "cuda_locscale_parity_test", "ref_cauchy_test", "ref_lognormal_test",
            "ref_cauchy_log_test", "ref_lognormal_log_test", "ref_gamma_test", "ref_exponential_test", "ref_gamma_log_test",
            "ref_exponential_log_test", "ref_beta_test", "ref_weibull_test", "ref_beta_log_test", "ref_weibull_log_test",
            "ref_inv_gamma_test", "ref_inv_gamma_log_test", "ref_mix_normal_test", "ref_rev_normal_log_test", "ref_fwd_normal_test"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", BINARIES)
def test_cxx_binary(name):
    exe = os.path.join(CXX_DIR, name)
    if not os.path.exists(exe):
        pytest.skip("%s not built (test/Makefile needs /root/reference)" % name)
    env = dict(os.environ)    # the real library, through RUNPATH: never the CPU suite's test double (build/mock)
    kept = [p for p in env.get("LD_LIBRARY_PATH", "").split(os.pathsep) if p and os.path.join("build", "mock") not in p]
    env.pop("LD_LIBRARY_PATH", None)
    if kept:
        env["LD_LIBRARY_PATH"] = os.pathsep.join(kept)
    ldd = subprocess.run(["ldd", exe], stdout=subprocess.PIPE, text=True, env=env).stdout
    if "libstanmath_cuda.so" in ldd:   # (tests whose argument kinds never route to the device do not link it)
        assert os.path.join("math_b200", "lib", "libstanmath_cuda.so") in ldd, ldd
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, env=env)
    tail = "\n".join(r.stdout.splitlines()[-40:])
    assert r.returncode == 0, tail
    assert "PASSED" in r.stdout, tail


def test_cxx_binaries_link_the_cuda_library_not_the_oracle():
    """The product's host layer links libstanmath_cuda.so and nothing under oracle/ except the
    vendored TBB runtime (map_rect threads)."""
    exe = os.path.join(CXX_DIR, "cuda_glm_parity_test")
    if not os.path.exists(exe):
        pytest.skip("not built")
    out = subprocess.run(["ldd", exe], stdout=subprocess.PIPE, text=True).stdout
    assert "libstanmath_cuda.so" in out
    assert "libstanref" not in out and "liboracle" not in out
