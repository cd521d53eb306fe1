"""CPU: the stored full-size reference outputs (tests/golden/configs.npz, made by the COMPILED REFERENCE) against
the plain-C oracle on regenerated inputs — pins the oracle at BASELINE sizes and proves the fixture reproduces.
c3 and c4s (10 GB / 5 GB of inputs, a minute of generation each) were compared when the fixture was made
(manifest key oracle_vs_reference) and are compared again by the GPU test; here c1, c2 and three c5 shards keep
the CPU suite short."""
import numpy as np
import pytest

from tests import config_cases as cc
from tests.helpers import RTOL, assert_close, check_glm_result


@pytest.fixture(scope="module")
def stored():
    return cc.load()


@pytest.fixture(scope="module")
def L():
    from oracle import loader
    return loader


def test_manifest_matches_fixture(stored):
    golden, man = stored
    assert set(man["configs"]) == {"c1", "c2", "c3", "c4s", "c5"}
    for name, (fam, n, k, seed) in cc.CONFIGS.items():
        m = man["configs"][name]
        assert (m["family"], m["n"], m["k"], m["seed"]) == (fam, n, k, seed)
        assert golden[name + "/dbeta"].shape == (k,) and golden[name + "/lp"].shape == (1,)
        assert np.all(np.isfinite(golden[name + "/dbeta"]))
        assert max(m["oracle_vs_reference"].values()) < RTOL          # margin recorded at full size
    assert golden["c4s/dsigma"].shape == (1,)
    assert golden["c5/dbeta"].shape == (cc.C5["shards"], cc.C5["k"]) and golden["c5/lp"].shape == (cc.C5["shards"],)
    assert man["configs"]["c5"]["oracle_vs_reference"] < RTOL
    assert "3.0.0" in man["reference"]


def test_c2_oracle_against_stored_reference(stored, L):
    golden, man = stored
    fam, x, y, alpha, beta, sigma = cc.glm_inputs("c2")
    if not cc.inputs_reproduce(man, "c2", x, y):
        pytest.skip("inputs do not regenerate bit for bit on this host (fixture made with torch %s)" % man["torch"])
    check_glm_result(L.oracle_glm(fam, x, y, alpha, beta, sigma), cc.stored_glm(golden, "c2"), "C2 oracle vs reference")


def test_c1_oracle_against_stored_reference(stored, L):
    golden, man = stored
    y, mu = cc.c1_inputs()
    if not cc.inputs_reproduce(man, "c1", mu, y):
        pytest.skip("inputs do not regenerate bit for bit on this host")
    r = L.oracle_normal_lpdf(y, mu, [cc.C1["sigma"]])
    assert_close(r["lp"], golden["c1/lp"][0], "C1 lp")
    assert_close(r["dsigma"], golden["c1/dsigma"], "C1 dsigma")
    assert_close(r["dmu"][golden["c1/dmu_idx"]], golden["c1/dmu_samples"], "C1 dmu samples")
    assert_close(np.abs(r["dmu"]).sum(), golden["c1/dmu_abs_sum"][0], "C1 sum |dmu|")


@pytest.mark.parametrize("j", [0, 31, 63])
def test_c5_shard_oracle_against_stored_reference(stored, L, j):
    golden, man = stored
    x, y, alpha, beta = cc.c5_shard(j)
    if j == 0 and not cc.c5_inputs_reproduce(man, [cc.c5_shard(q) for q in range(cc.C5["shards"])]):
        pytest.skip("inputs do not regenerate bit for bit on this host")
    r = L.oracle_glm(cc.C5["family"], x, y, alpha, beta)
    assert_close(r["lp"], golden["c5/lp"][j], "C5 shard lp")
    assert_close(r["dbeta"], golden["c5/dbeta"][j], "C5 shard dbeta")
    assert_close(r["dalpha"][0], golden["c5/dalpha"][j], "C5 shard dalpha")
