"""Multi-GPU tests (need >= 2 B200s: `gpurun --gpus 2`): rows sharded over the devices of one
process with an NCCL all-reduce of the result block, and shard scopes landing on distinct devices."""
import numpy as np
import pytest

from tests.helpers import check_glm_result, synth_glm

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def glm2():
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from math_b200 import glm
    from math_b200._lib import lib
    glm.init(0)
    if lib().stan_cuda_device_count() < 2:
        pytest.skip("library was initialised with one device earlier in this process")
    return glm


@pytest.mark.parametrize("family", ["poisson_log", "normal_id"])
def test_in_process_row_sharding_with_nccl(glm2, family):
    from math_b200._lib import lib
    from oracle import loader as L
    lib().stan_cuda_set_shard_min_elems(1 << 16)
    n, k = 200_003, 48
    s = synth_glm(family, n, k, seed=17)
    ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
    st0 = glm2.stats()
    got = glm2.glm_eval(family, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
    st1 = glm2.stats()
    check_glm_result(got, ref, "sharded " + family)
    assert st1["allreduces"] - st0["allreduces"] == 1
    assert st1["kernel_launches"] - st0["kernel_launches"] >= 2
    # vector alpha: per-row partials come back from each device's row block
    alpha = 0.3 + 0.05 * np.random.default_rng(1).standard_normal(n)
    ref = L.oracle_glm(family, s["x"], s["y"], alpha, s["beta"], s["sigma"])
    got = glm2.glm_eval(family, s["y"], s["x"], alpha, s["beta"], s["sigma"])
    check_glm_result(got, ref, "sharded vector alpha " + family)
    glm2.cache_clear()
    lib().stan_cuda_set_shard_min_elems(1 << 24)


def test_shard_scopes_use_distinct_devices(glm2):
    from oracle import loader as L
    s = synth_glm("bernoulli_logit", 30000, 32, seed=4)
    ref = L.oracle_glm("bernoulli_logit", s["x"], s["y"], s["alpha"], s["beta"])
    for job in range(4):
        with glm2.shard_scope(call_id=9, job=job, device=job % 2):
            got = glm2.glm_eval("bernoulli_logit", s["y"], s["x"], s["alpha"], s["beta"])
        check_glm_result(got, ref, "scope job %d" % job)
    glm2.cache_clear()
