"""GPU parity tests of the batched evaluation (stan_cuda_glm_eval_batch: the jobs of a map_rect call whose
functor is a single GLM call, one launch per device): every item against the plain-C oracle and against its own
single evaluation; ragged shard sizes; cache keyed by (call_id, job); error codes."""
import numpy as np
import pytest

from math_b200 import _abi
from tests.helpers import assert_close, check_glm_result, synth_glm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    g.init(1)
    g.cache_clear()
    yield g
    g.cache_clear()


@pytest.fixture(scope="module")
def L():
    from oracle import loader
    return loader


@pytest.mark.parametrize("family,k,sizes", [
    ("bernoulli_logit", 32, [100_000] * 8),                                   # a slice of C5
    ("bernoulli_logit", 32, [1, 7, 129, 1000, 4097, 20_000, 333, 50_001, 64, 2]),   # ragged, down to a single row
    ("poisson_log", 128, [5000, 12_345, 700, 30_000, 32]),
    ("normal_id", 7, [999, 1000, 1001, 40_000]),
    ("poisson_log", 1, [3000, 10]),
    ("bernoulli_logit", 32, [2000] * 70),                                      # more items than fit one wave of 2 CTAs per SM at gx > 4
])
def test_batch_matches_oracle_and_single_evaluations(glm, L, family, k, sizes):
    cases = [synth_glm(family, n, k, seed=17 * j + n) for j, n in enumerate(sizes)]
    alphas = [0.3 + 0.01 * j for j in range(len(cases))]
    sigmas = [1.0 + 0.1 * j for j in range(len(cases))] if family == "normal_id" else None
    betas = np.stack([c["beta"] * (1 + 0.05 * j) for j, c in enumerate(cases)])
    batch = glm.GlmBatch(family, [c["x"] for c in cases], [c["y"] for c in cases], call_id=len(sizes) + k)
    rc, res = batch.eval(alphas, betas, sigmas)
    assert rc == 0
    for j, c in enumerate(cases):
        sg = [sigmas[j]] if sigmas else None
        ref = L.oracle_glm(c["family"], c["x"], c["y"], [alphas[j]], betas[j], sg)
        check_glm_result(res[j], ref, "%s batch item %d (n=%d)" % (family, j, sizes[j]))
        if j < 4:
            one = glm.glm_eval(c["family"], c["y"], c["x"], [alphas[j]], betas[j], sg)
            assert_close(res[j]["lp"], one["lp"], "vs single lp", 1e-12)
            assert_close(res[j]["dbeta"], one["dbeta"], "vs single dbeta", 1e-11)
    hits0 = glm.stats()["cache_hits"]
    rc2, res2 = batch.eval(alphas, betas, sigmas)                   # second call: every item cached, same bits
    assert glm.stats()["cache_hits"] - hits0 == len(cases)
    for a, b in zip(res, res2):
        assert a["lp"] == b["lp"] and np.array_equal(a["dbeta"], b["dbeta"])
    rc3, res3 = batch.eval([a + 0.5 for a in alphas], betas * 0.5, sigmas)   # new parameters, same data
    ref = L.oracle_glm(cases[0]["family"], cases[0]["x"], cases[0]["y"], [alphas[0] + 0.5], betas[0] * 0.5,
                       [sigmas[0]] if sigmas else None)
    check_glm_result(res3[0], ref, "new parameters")


def test_batch_codes(glm):
    c = [synth_glm("poisson_log", 500, 4, seed=s) for s in (1, 2, 3)]
    ys = [x["y"].copy() for x in c]
    ys[1][7] = -3
    batch = glm.GlmBatch("poisson_log", [x["x"] for x in c], ys, call_id=901)
    rc, res = batch.eval([0.1] * 3, c[0]["beta"])
    assert rc == _abi.BAD_OUTCOME and [r["rc"] for r in res] == [0, _abi.BAD_OUTCOME, 0]
    batch = glm.GlmBatch("poisson_log", [x["x"] for x in c], [x["y"] for x in c], call_id=902)
    beta = c[0]["beta"].copy()
    rc, res = batch.eval([0.1, np.inf, 0.1], beta)
    assert rc == _abi.NONFINITE and [r["rc"] for r in res] == [0, _abi.NONFINITE, 0]
    mixed = glm.GlmBatch("poisson_log", [c[0]["x"], c[1]["x"][:, :3]], [c[0]["y"], c[1]["y"]], call_id=903)
    with pytest.raises(Exception):
        mixed.eval([0.1, 0.1], np.zeros(3))
