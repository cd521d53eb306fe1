"""math_b200 -- B200-native (sm_100a) GLM log-density + gradient path for Stan Math.

The product is libstanmath_cuda.so (math_b200/csrc, C ABI in include/stan_cuda.h) and the C++
host layer under stan/math/cuda/.  This Python package is the thin ctypes binding used by the
tests and the benchmark; it mirrors the reference's function names and argument meaning.
"""
from ._abi import (BERNOULLI_LOGIT, NORMAL_ID, POISSON_LOG, LP_ALL, LP_CONST, LP_MAIN, LP_SCALE,  # noqa: F401
                   NEED_ALL, NEED_DALPHA, NEED_DBETA, NEED_DSIGMA, NEED_DY, NONFINITE, OK)


def __getattr__(name):
    # lazy: importing the package must not need the built library
    if name in ("bernoulli_logit_glm_lpmf", "poisson_log_glm_lpmf", "normal_id_glm_lpdf", "neg_binomial_2_log_glm_lpmf", "normal_lpdf", "student_t_lpdf",
                "GlmData", "glm_eval", "init", "init_device", "describe", "stats", "cache_clear",
                "cache_invalidate", "shard_scope", "DomainError", "comm_init_from_torch"):
        from . import glm
        return getattr(glm, name)
    raise AttributeError(name)
