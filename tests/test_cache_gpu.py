"""GPU test of the device cache's byte budget (stan_cuda_cache_set_limit): least-recently-used data sets are
dropped when a new one does not fit, results stay right, a dropped data set is simply uploaded again."""
import numpy as np
import pytest

from tests.helpers import check_glm_result, synth_glm

pytestmark = pytest.mark.gpu


def test_cache_budget_evicts_least_recently_used():
    from math_b200 import glm
    from math_b200._lib import lib
    from oracle import loader as L
    glm.init(1)
    glm.cache_clear()
    lib().stan_cuda_stats_reset()
    cases = [synth_glm("poisson_log", 40_000, 32, seed=s) for s in (1, 2, 3)]   # ~10.4 MB each on the device
    refs = [L.oracle_glm(c["family"], c["x"], c["y"], c["alpha"], c["beta"], c["sigma"]) for c in cases]

    def run(i):
        c = cases[i]
        r = glm.glm_eval(c["family"], c["y"], c["x"], c["alpha"], c["beta"], c["sigma"])
        check_glm_result(r, refs[i], "case %d" % i)

    try:
        lib().stan_cuda_cache_set_limit(25 << 20)   # room for two of the three
        run(0)
        run(1)
        assert glm.stats()["cache_evictions"] == 0 and glm.stats()["cache_misses"] == 2
        run(0)                                       # hit: case 0 is now the most recently used
        assert glm.stats()["cache_hits"] == 1
        run(2)                                       # does not fit: case 1 (least recently used) goes
        st = glm.stats()
        assert st["cache_evictions"] == 1 and st["cache_misses"] == 3
        assert lib().stan_cuda_cache_bytes() <= (25 << 20)
        run(0)                                       # still resident
        assert glm.stats()["cache_hits"] == 2
        run(1)                                       # was dropped: uploaded again, evicting case 2
        st = glm.stats()
        assert st["cache_misses"] == 4 and st["cache_evictions"] == 2
        lib().stan_cuda_cache_set_limit(1 << 20)     # smaller than any data set: everything resident is dropped
        assert lib().stan_cuda_cache_bytes() == 0
        run(2)                                       # a data set larger than the budget still evaluates
    finally:
        lib().stan_cuda_cache_set_limit(0)
        glm.cache_clear()
