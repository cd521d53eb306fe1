"""GPU stress tests of the hand-rolled synchronisation in csrc/glm_warp.cu (compute-sanitizer is closed on this GPU
pool, profiles/r02_sanitizer_unavailable.txt): the TMA stage ring with arrival counters, the ticketed reductions that
share one workspace across launches, and the early start of evaluation n + 1 behind the tail of evaluation n
(programmatic dependent launch).  A race shows as a wrong sum or as a changed low bit between repeats."""
import numpy as np
import pytest

from tests.helpers import assert_close, synth_glm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    from math_b200._lib import lib
    g.init(1)
    lib().stan_cuda_set_kernel(1)
    yield g
    lib().stan_cuda_set_tuning(0, 0, 0)
    g.cache_clear()


@pytest.mark.parametrize("family,n,k", [("poisson_log", 400_003, 128), ("bernoulli_logit", 1_000_001, 32),
                                        ("normal_id", 150_001, 256), ("poisson_log", 70_001, 16)])
def test_back_to_back_launches_overlap_without_interfering(glm, family, n, k):
    """40 evaluations with 40 different parameter blocks enqueued back to back on one stream (they share the data
    handle's workspace and, with programmatic dependent launch, overlap on the device) must each equal the same
    evaluation run alone -- bit for bit -- and the oracle to 1e-10."""
    import torch
    from oracle import loader as L
    s = synth_glm(family, n, k, seed=3)
    fam = s["family"]
    data = glm.GlmData(fam, s["x"], s["y"])
    flags = glm.glm_flags(fam, False)
    K = 40
    try:
        params = np.zeros((K, data.params_len))
        for i in range(K):
            params[i, :k] = s["beta"] * (1.0 + 0.01 * i)
            params[i, data.kpad] = s["alpha"][0] + 0.001 * i
            params[i, data.kpad + 1] = s["sigma"][0] if s["sigma"] is not None else 1.0
        d_params = torch.from_numpy(params).cuda()
        d_res = torch.zeros((K, data.result_len), dtype=torch.float64, device="cuda")
        d_alone = torch.zeros((K, data.result_len), dtype=torch.float64, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        for rep in range(3):
            d_res.zero_()
            torch.cuda.synchronize()
            for i in range(K):   # back to back: nothing between the launches
                data.launch(flags, d_params[i].data_ptr(), d_res[i].data_ptr(), stream.cuda_stream)
            stream.synchronize()
            if rep == 0:
                for i in range(K):   # one at a time
                    data.launch(flags, d_params[i].data_ptr(), d_alone[i].data_ptr(), stream.cuda_stream)
                    stream.synchronize()
            assert torch.equal(d_res, d_alone), "overlapping launches differ from launches run alone (rep %d)" % rep
        r = d_res.cpu().numpy()
        for i in (0, 17, K - 1):
            ref = L.oracle_glm(family, s["x"], s["y"], [params[i, data.kpad]], params[i, :k], s["sigma"])
            assert_close(r[i, 0], ref["lp"], "launch %d lp" % i)
            assert_close(r[i, 4:4 + k], ref["dbeta"], "launch %d dbeta" % i)
            assert_close(r[i, 1], ref["dalpha"][0], "launch %d dalpha" % i)
    finally:
        data.close()


def test_many_shapes_stage_counts_and_grids(glm):
    """200 evaluations over changing shapes, pipeline depths and CTAs per SM: each against the oracle, each repeated
    once and required to be bit-identical."""
    from math_b200._lib import lib
    from oracle import loader as L
    rng = np.random.default_rng(11)
    fams = ["bernoulli_logit", "poisson_log", "normal_id"]
    count = 0
    try:
        for it in range(50):
            fam = fams[it % 3]
            n = int(rng.integers(1, 60_000))
            k = int(rng.choice([1, 3, 16, 17, 32, 33, 64, 100, 128, 200, 256, 300, 512]))
            s = synth_glm(fam, n, k, seed=100 + it)
            if k > 64:
                s["beta"] = s["beta"] * (8.0 / np.sqrt(k))
            ref = L.oracle_glm(fam, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"])
            for stages, ctas in ((0, 0), (2, 1), (3, 2), (6, 1)):
                lib().stan_cuda_set_tuning(-1, stages, ctas)
                a = glm.glm_eval(fam, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
                b = glm.glm_eval(fam, a["_keep"][1], a["_keep"][0], s["alpha"], s["beta"], s["sigma"])
                what = "%s %dx%d stages=%d ctas=%d" % (fam, n, k, stages, ctas)
                assert_close(a["lp"], ref["lp"], what + " lp")
                assert_close(a["dbeta"], ref["dbeta"], what + " dbeta")
                assert_close(a["dalpha"], ref["dalpha"], what + " dalpha")
                assert a["lp"] == b["lp"] and np.array_equal(a["dbeta"], b["dbeta"]), what + ": repeat differs"
                count += 1
            glm.cache_clear()
    finally:
        lib().stan_cuda_set_tuning(0, 0, 0)
    assert count == 200


@pytest.mark.parametrize("family", ["neg_binomial_2_log", "ordered_logistic", "categorical_mma", "categorical_vector"])
def test_back_to_back_launches_of_the_widened_families(glm, family):
    """The same check for the families that joined the dependent-launch chain later: neg_binomial_2_log (whose
    histogram pre-kernel sits between two GLM kernels and writes its two sums only after the previous evaluation
    has completed), ordered_logistic and both categorical_logit kernels."""
    import torch
    from math_b200 import _abi
    from math_b200._lib import lib
    from tests.helpers import synth_categorical, synth_ordered
    n, k, K = 300_007, 32, 30
    ncls = 5
    lib().stan_cuda_set_categorical_kernel(0 if family == "categorical_vector" else 1)
    try:
        if family == "neg_binomial_2_log":
            s = synth_glm(family, n, k, seed=4)
            data = glm.GlmData(s["family"], s["x"], s["y"])
            flags = glm.glm_flags(s["family"], False)
            plen, rlen = data.params_len, data.result_len
        elif family == "ordered_logistic":
            s = synth_ordered(n, k, ncls, seed=4)
            data = glm.GlmData(_abi.ORDERED_LOGISTIC, s["x"], s["y"])
            flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
            plen, rlen = data.params_len + ncls - 1, data.result_len + ncls - 1
        else:
            s = synth_categorical(n, k, ncls, seed=4)
            data = glm.GlmData(_abi.CATEGORICAL_LOGIT, s["x"], s["y"])
            flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA
            plen = ncls * (data.kpad + 1)
            rlen = 4 + plen
        params = np.zeros((K, plen))
        for i in range(K):
            f = 1.0 + 0.01 * i
            if family == "neg_binomial_2_log":
                params[i, :k], params[i, data.kpad], params[i, data.kpad + 1] = s["beta"] * f, 0.3, 2.5 + 0.01 * i
            elif family == "ordered_logistic":
                params[i, :k] = s["beta"] * f
                params[i, data.kpad + 2:] = s["cuts"] * f
            else:
                for c in range(ncls):
                    params[i, c * data.kpad:c * data.kpad + k] = s["beta"][:, c] * f
                params[i, ncls * data.kpad:] = s["alpha"] * f
        d_params = torch.from_numpy(params).cuda()
        d_res = torch.zeros((K, rlen), dtype=torch.float64, device="cuda")
        d_alone = torch.zeros((K, rlen), dtype=torch.float64, device="cuda")
        stream = torch.cuda.Stream()

        def go(i, out):
            if family == "neg_binomial_2_log":
                data.launch(flags, d_params[i].data_ptr(), out[i].data_ptr(), stream.cuda_stream)
            elif family == "ordered_logistic":
                data.launch_ordered(flags, d_params[i].data_ptr(), ncls - 1, out[i].data_ptr(), stream.cuda_stream)
            else:
                data.launch_categorical(flags, d_params[i].data_ptr(), ncls, out[i].data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        for i in range(K):
            go(i, d_alone)
            stream.synchronize()
        for rep in range(3):
            d_res.zero_()
            torch.cuda.synchronize()
            for i in range(K):
                go(i, d_res)
            stream.synchronize()
            assert torch.equal(d_res, d_alone), "%s: overlapping launches differ from launches run alone (rep %d)" % (family, rep)
        assert torch.isfinite(d_res[:, 0]).all() and float(d_res[:, 0].abs().min()) > 0
        data.close()
    finally:
        lib().stan_cuda_set_categorical_kernel(1)
