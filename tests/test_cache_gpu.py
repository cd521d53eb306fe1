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


def test_two_glm_calls_of_one_map_rect_job_do_not_share_device_data():
    """Advisor finding (round 1, high): inside a shard scope the key was (call_id, job, shape, family) + 4 samples,
    so two GLM calls of one functor with the same design matrix and outcome vectors that differ only at unsampled
    positions shared one device copy.  The ordinal of the call inside the job is now part of the identity."""
    from math_b200 import glm
    from oracle import loader as L
    glm.init(1)
    glm.cache_clear()
    c = synth_glm("bernoulli_logit", 30_011, 32, seed=11)
    y1 = c["y"].copy()
    y2 = y1.copy()
    y2[3:-3] = 1 - y2[3:-3]          # differs everywhere except at the ends (which any sampling scheme reads)
    refs = [L.oracle_glm("bernoulli_logit", c["x"], y, c["alpha"], c["beta"]) for y in (y1, y2)]
    assert abs(refs[0]["lp"] - refs[1]["lp"]) > 1.0
    for evaluation in range(3):       # every evaluation of the job makes the same two calls in the same order
        st0 = glm.stats()
        with glm.shard_scope(call_id=77, job=5, device=0):
            x_copy = np.asfortranarray(c["x"].copy())   # a functor rebuilds x at a fresh address every time
            r1 = glm.glm_eval("bernoulli_logit", y1.copy(), x_copy, c["alpha"], c["beta"])
            r2 = glm.glm_eval("bernoulli_logit", y2.copy(), x_copy, c["alpha"], c["beta"])
        st1 = glm.stats()
        check_glm_result(r1, refs[0], "first call of the job, evaluation %d" % evaluation)
        check_glm_result(r2, refs[1], "second call of the job, evaluation %d" % evaluation)
        if evaluation > 0:            # both data sets stay resident: no eviction ping-pong, no upload
            assert st1["cache_misses"] == st0["cache_misses"] and st1["cache_hits"] - st0["cache_hits"] == 2
            assert st1["bytes_uploaded"] == st0["bytes_uploaded"]
    glm.cache_clear()


def test_cache_verify_mode_notices_in_place_mutation_the_samples_miss():
    """stan_cuda_cache_set_verify(1) hashes every element of x and y per call: a host array rewritten in place at a
    position the 64-sample fingerprint does not read is re-uploaded instead of evaluated stale."""
    from math_b200 import glm
    from math_b200._lib import lib
    from oracle import loader as L
    glm.init(1)
    glm.cache_clear()
    c = synth_glm("poisson_log", 50_003, 24, seed=12)
    x, y = np.asfortranarray(c["x"]), c["y"].copy()
    try:
        lib().stan_cuda_cache_set_verify(1)
        r = glm.glm_eval("poisson_log", y, x, c["alpha"], c["beta"])
        check_glm_result(r, L.oracle_glm("poisson_log", x, y, c["alpha"], c["beta"]), "before the mutation")
        x[12_345, 7] += 0.5                      # not a sampled position
        y[23_456] += 3
        st0 = glm.stats()
        r = glm.glm_eval("poisson_log", y, x, c["alpha"], c["beta"])
        st1 = glm.stats()
        assert st1["cache_misses"] - st0["cache_misses"] == 1
        check_glm_result(r, L.oracle_glm("poisson_log", x, y, c["alpha"], c["beta"]), "after the mutation")
        # the explicit route without verify mode: stan_cuda_cache_invalidate after an in-place change
        lib().stan_cuda_cache_set_verify(0)
        x[3_456, 3] -= 0.25
        assert glm.cache_invalidate(x) >= 1
        r = glm.glm_eval("poisson_log", y, x, c["alpha"], c["beta"])
        check_glm_result(r, L.oracle_glm("poisson_log", x, y, c["alpha"], c["beta"]), "after invalidate")
    finally:
        lib().stan_cuda_cache_set_verify(0)
        glm.cache_clear()


@pytest.mark.parametrize("family", ["normal_id", "poisson_log", "neg_binomial_2_log"])
def test_per_call_outcomes_do_not_refill_x(family):
    """y_per_call (a var outcome vector, a broadcast scalar outcome): the cache is keyed on x alone, y travels with
    every call together with its range / sum lgamma(y + 1); x is uploaded once."""
    from math_b200 import glm
    from oracle import loader as L
    glm.init(1)
    glm.cache_clear()
    c = synth_glm(family, 20_005, 40, seed=13)
    x = np.asfortranarray(c["x"])
    rng = np.random.default_rng(3)
    uploaded = []
    for it in range(3):
        y = (c["y"] + rng.standard_normal(c["y"].shape)) if family == "normal_id" else rng.permutation(c["y"])
        y = np.ascontiguousarray(y.copy())       # a fresh address every call, like the wrapper's temporary
        st0 = glm.stats()
        r = glm.glm_eval(family, y, x, c["alpha"], c["beta"], c["sigma"], y_per_call=True, need_dy=(family == "normal_id"))
        st1 = glm.stats()
        uploaded.append(st1["bytes_uploaded"] - st0["bytes_uploaded"])
        ref = L.oracle_glm(family, x, y, c["alpha"], c["beta"], c["sigma"], need_dy=(family == "normal_id"))
        check_glm_result(r, ref, "%s per-call y, call %d" % (family, it))
    assert uploaded[0] > x.nbytes * 0.9 and uploaded[1] == 0 and uploaded[2] == 0, uploaded
    # a per-call outcome outside the family's domain is reported from the range recomputed for THIS call
    if family != "normal_id":
        bad = c["y"].copy()
        bad[777] = -1
        assert glm.glm_eval(family, bad, x, c["alpha"], c["beta"], c["sigma"], y_per_call=True)["rc"] == 2
    glm.cache_clear()
