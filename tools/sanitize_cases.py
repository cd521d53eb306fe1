"""Small evaluations of every kernel family for compute-sanitizer (tools/sanitize.sh): memcheck, racecheck and
synccheck over the TMA ring + arrival counters of glm_warp_kernel, the batched launch, the FP64 tensor-core
categorical kernel, the elementwise kernels and -- with two devices -- the in-kernel peer exchange.
Every result is also compared with the plain-C oracle, so a sanitizer run doubles as a parity run."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from math_b200 import _abi, glm  # noqa: E402
from math_b200._lib import lib  # noqa: E402
from oracle import loader as L  # noqa: E402
from tests.helpers import (check_cat_result, check_glm_result, check_ord_result, synth_categorical, synth_glm,  # noqa: E402
                           synth_ordered)


def main():
    n_dev = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    which = sys.argv[2] if len(sys.argv) > 2 else "all"
    glm.init(n_dev)
    if n_dev > 1:
        lib().stan_cuda_set_shard_min_elems(1 << 12)   # row-shard even these small problems: the fused exchange runs
    print(glm.describe())
    done = []
    if which in ("all", "glm"):
        for fam, n, k in (("bernoulli_logit", 4097, 33), ("poisson_log", 30011, 128), ("normal_id", 30011, 128),
                          ("neg_binomial_2_log", 4097, 33), ("poisson_log", 8191, 256), ("bernoulli_logit", 5000, 7)):
            s = synth_glm(fam, n, k, seed=5)
            got = glm.glm_eval(fam, s["y"], s["x"], s["alpha"], s["beta"], s["sigma"])
            check_glm_result(got, L.oracle_glm(fam, s["x"], s["y"], s["alpha"], s["beta"], s["sigma"]), "%s %dx%d" % (fam, n, k))
            glm.glm_eval(fam, got["_keep"][1], got["_keep"][0], s["alpha"], s["beta"] * 1.01, s["sigma"])   # cache hit, second launch
            done.append("%s %dx%d" % (fam, n, k))
        # ROWIO: vector intercept (per-row d/dalpha) and vector scale
        s = synth_glm("normal_id", 150001, 40, seed=6)
        alpha = 0.3 + 0.05 * np.random.default_rng(1).standard_normal(150001)
        sigma = 1.0 + 0.1 * np.random.default_rng(2).random(150001)
        got = glm.glm_eval("normal_id", s["y"], s["x"], alpha, s["beta"], sigma, need_dy=True)
        check_glm_result(got, L.oracle_glm("normal_id", s["x"], s["y"], alpha, s["beta"], sigma, need_dy=True), "ROWIO 150001x40")
        done.append("normal_id ROWIO 150001x40")
        so = synth_ordered(20011, 48, 6, seed=7)
        got = glm.ordered_logistic_glm_eval(so["y"], so["x"], so["beta"], so["cuts"], _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS)
        check_ord_result(got, L.oracle_ordered_logistic_glm(so["x"], so["y"], so["beta"], so["cuts"]), "ordered 20011x48")
        done.append("ordered_logistic 20011x48x6")
    if which in ("all", "cat"):
        for ncls, k in ((5, 32), (16, 40)):
            sc = synth_categorical(20011, k, ncls, seed=8)
            got = glm.categorical_logit_glm_eval(sc["y"], sc["x"], sc["alpha"], sc["beta"], _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA)
            check_cat_result(got, L.oracle_categorical_logit_glm(sc["x"], sc["y"], sc["alpha"], sc["beta"]), "categorical")
            done.append("categorical_logit (DMMA) 20011x%dx%d" % (k, ncls))
    if which in ("all", "batch") and n_dev == 1:
        cases = [synth_glm("bernoulli_logit", 6000 + 17 * j, 32, seed=20 + j) for j in range(6)]
        b = glm.GlmBatch("bernoulli_logit", [c["x"] for c in cases], [c["y"] for c in cases], call_id=3)
        rc, res = b.eval([0.1 * j for j in range(6)], cases[0]["beta"])
        for j, c in enumerate(cases):
            ref = L.oracle_glm("bernoulli_logit", c["x"], c["y"], [0.1 * j], cases[0]["beta"])
            check_glm_result(res[j], ref, "batch item %d" % j)
        done.append("glm_warp_batch_kernel 6 shards")
    if which in ("all", "dens") and n_dev == 1:
        rng = np.random.default_rng(9)
        th = rng.standard_normal(50001)
        yb = (rng.random(50001) < 0.5).astype(np.int32)
        lp, g = glm.bernoulli_logit_lpmf(yb, th)
        o = L.oracle_density("bernoulli_logit", yb, th)
        assert abs(lp - o["lp"]) <= 1e-10 * abs(o["lp"])
        lp, g = glm.normal_lpdf(th, th * 0.5, 1.3)
        o = L.oracle_normal_lpdf(th, th * 0.5, 1.3)
        assert abs(lp - o["lp"]) <= 1e-10 * abs(o["lp"])
        done.append("density_kernel + normal_lpdf_kernel 50001")
    st = glm.stats()
    print("sanitize_cases OK: %d cases, launches=%d fused_allreduces=%d nccl_allreduces=%d" %
          (len(done), st["kernel_launches"], st["fused_allreduces"], st["allreduces"]))
    for d in done:
        print("  ", d)
    glm.cache_clear()


if __name__ == "__main__":
    main()
