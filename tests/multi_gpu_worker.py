"""Bodies of the multi-GPU tests.  Run as a script in a FRESH process by tests/test_multi_gpu.py:
the library binds its device set once per process (stan_cuda_init), and the single-GPU tests of
the same pytest session have already bound one device."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests.helpers import (check_cat_result, check_glm_result, check_ord_result, synth_categorical, synth_glm,  # noqa: E402
                           synth_ordered)


def in_process_row_sharding_with_nccl(family):
    from math_b200 import glm
    from math_b200._lib import lib
    from oracle import loader as L
    glm.init(0)
    assert lib().stan_cuda_device_count() >= 2
    lib().stan_cuda_set_shard_min_elems(1 << 16)
    n, k = 200_003, 48
    s = synth_glm(family, n, k, seed=17)
    ref = L.oracle_glm(family, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
    st0 = glm.stats()
    got = glm.glm_eval(family, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
    st1 = glm.stats()
    check_glm_result(got, ref, "sharded " + family)
    # one all-reduce per evaluation: inside the kernel over peer memory, or ncclAllReduce
    assert (st1["allreduces"] + st1["fused_allreduces"]) - (st0["allreduces"] + st0["fused_allreduces"]) == 1
    print("collective:", lib().stan_cuda_comm_describe().decode(), "fused", st1["fused_allreduces"] - st0["fused_allreduces"])
    assert st1["kernel_launches"] - st0["kernel_launches"] >= 2
    again = glm.glm_eval(family, got["_keep"][1], got["_keep"][0], s["alpha"], s["beta"], s["sigma"])
    assert again["lp"] == got["lp"] and np.array_equal(again["dbeta"], got["dbeta"])  # reproducible
    # vector alpha: per-row partials come back from each device's row block
    alpha = 0.3 + 0.05 * np.random.default_rng(1).standard_normal(n)
    ref = L.oracle_glm(family, s["x"], s["y"], alpha, s["beta"], s["sigma"])
    got = glm.glm_eval(family, s["y"], s["x"], alpha, s["beta"], s["sigma"])
    check_glm_result(got, ref, "sharded vector alpha " + family)
    glm.cache_clear()


def widened_families_row_sharded():
    """neg_binomial_2_log, ordered_logistic (result block with d/dcuts: fused exchange) and categorical_logit
    (result block wider than an exchange slot: ncclAllReduce behind the kernels) across the devices."""
    from math_b200 import _abi, glm
    from math_b200._lib import lib
    from oracle import loader as L
    glm.init(0)
    assert lib().stan_cuda_device_count() >= 2
    lib().stan_cuda_set_shard_min_elems(1 << 16)
    n, k = 150_001, 48
    s = synth_glm("neg_binomial_2_log", n, k, seed=23)
    ref = L.oracle_glm("neg_binomial_2_log", s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
    got = glm.glm_eval("neg_binomial_2_log", s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
    check_glm_result(got, ref, "sharded neg_binomial_2_log")
    so = synth_ordered(n, k, 6, seed=29)
    flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
    st0 = glm.stats()
    got = glm.ordered_logistic_glm_eval(so["y"], so["x"], so["beta"], so["cuts"], flags)
    st1 = glm.stats()
    check_ord_result(got, L.oracle_ordered_logistic_glm(so["x"], so["y"], so["beta"], so["cuts"]), "sharded ordered_logistic")
    assert st1["kernel_launches"] - st0["kernel_launches"] >= 2
    for ncls in (3, 16):   # 16 classes x k_pad 48: 788 doubles > one exchange slot
        sc = synth_categorical(n, k, ncls, seed=31 + ncls)
        st0 = glm.stats()
        got = glm.categorical_logit_glm_eval(sc["y"], sc["x"], sc["alpha"], sc["beta"],
                                             _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA)
        st1 = glm.stats()
        check_cat_result(got, L.oracle_categorical_logit_glm(sc["x"], sc["y"], sc["alpha"], sc["beta"]),
                         "sharded categorical_logit C=%d" % ncls)
        assert (st1["allreduces"] + st1["fused_allreduces"]) - (st0["allreduces"] + st0["fused_allreduces"]) == 1
        print("categorical C=%d: nccl %d fused %d" % (ncls, st1["allreduces"] - st0["allreduces"],
                                                      st1["fused_allreduces"] - st0["fused_allreduces"]))
    glm.cache_clear()


def shard_scopes_use_distinct_devices():
    from math_b200 import glm
    from oracle import loader as L
    glm.init(0)
    s = synth_glm("bernoulli_logit", 30000, 32, seed=4)
    ref = L.oracle_glm("bernoulli_logit", s["x"], s["y"], s["alpha"], s["beta"])
    for job in range(4):
        with glm.shard_scope(call_id=9, job=job, device=job % 2):
            got = glm.glm_eval("bernoulli_logit", s["y"], s["x"], s["alpha"], s["beta"])
        check_glm_result(got, ref, "scope job %d" % job)
    glm.cache_clear()


if __name__ == "__main__":
    case = sys.argv[1]
    if case == "sharding":
        in_process_row_sharding_with_nccl(sys.argv[2])
    elif case == "widened":
        widened_families_row_sharded()
    elif case == "scopes":
        shard_scopes_use_distinct_devices()
    else:
        raise SystemExit("unknown case " + case)
    print("OK", case)
