"""GPU: the CUDA path at the FULL sizes of the BASELINE configs against the stored outputs of the compiled
reference (tests/golden/configs.npz; inputs regenerated from seeds, checked against the manifest's checksums).
C3 is compared inside tests/test_glm_gpu.py::test_full_size_c3_row_additivity_and_sample_parity, which already
generates its 10 GB of inputs."""
import numpy as np
import pytest

from tests import config_cases as cc
from tests.helpers import assert_close, check_glm_result

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    g.init(1)
    g.cache_clear()
    yield g
    g.cache_clear()


@pytest.fixture(scope="module")
def stored():
    return cc.load()


def test_c1_normal_lpdf_against_stored_reference(glm, stored):
    golden, man = stored
    y, mu = cc.c1_inputs()
    if not cc.inputs_reproduce(man, "c1", mu, y):
        pytest.skip("inputs do not regenerate bit for bit on this host")
    lp, g = glm.normal_lpdf(y, mu, [cc.C1["sigma"]])
    assert_close(lp, golden["c1/lp"][0], "C1 lp")
    assert_close(g["sigma"], golden["c1/dsigma"], "C1 dsigma")
    assert_close(g["mu"][golden["c1/dmu_idx"]], golden["c1/dmu_samples"], "C1 dmu samples")
    assert_close(np.abs(g["mu"]).sum(), golden["c1/dmu_abs_sum"][0], "C1 sum |dmu|")
    glm.cache_clear()


@pytest.mark.parametrize("name", ["c2", "c4s"])
def test_glm_config_against_stored_reference(glm, stored, name):
    import torch
    golden, man = stored
    if name == "c4s" and torch.cuda.mem_get_info()[0] < 20e9:
        pytest.skip("needs 20 GB of free device memory")
    fam, x, y, alpha, beta, sigma = cc.glm_inputs(name)
    if not cc.inputs_reproduce(man, name, x, y):
        pytest.skip("inputs do not regenerate bit for bit on this host")
    got = glm.glm_eval(fam, y, x, alpha, beta, sigma)
    check_glm_result(got, cc.stored_glm(golden, name), name + " vs compiled reference")
    glm.cache_clear()


def test_c5_all_shards_batched_and_single_against_stored_reference(glm, stored):
    """the 64 shards of C5: one batched launch, and shard by shard, both against the reference's per-shard outputs"""
    golden, man = stored
    shards = [cc.c5_shard(j) for j in range(cc.C5["shards"])]
    if not cc.c5_inputs_reproduce(man, shards):
        pytest.skip("inputs do not regenerate bit for bit on this host")
    batch = glm.GlmBatch(cc.C5["family"], [s[0] for s in shards], [s[1] for s in shards], call_id=5005)
    rc, res = batch.eval([s[2][0] for s in shards], np.stack([s[3] for s in shards]))
    assert rc == 0
    for j, r in enumerate(res):
        assert r["rc"] == 0
        assert_close(r["lp"], golden["c5/lp"][j], "C5 batched shard %d lp" % j)
        assert_close(r["dbeta"], golden["c5/dbeta"][j], "C5 batched shard %d dbeta" % j)
        assert_close(r["dalpha"][0], golden["c5/dalpha"][j], "C5 batched shard %d dalpha" % j)
    for j in (0, 17, 63):
        x, y, alpha, beta = shards[j]
        one = glm.glm_eval(cc.C5["family"], y, x, alpha, beta)
        assert_close(one["lp"], golden["c5/lp"][j], "C5 shard %d lp" % j)
        assert_close(one["dbeta"], golden["c5/dbeta"][j], "C5 shard %d dbeta" % j)
    glm.cache_clear()
