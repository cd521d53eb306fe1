"""Wall time of one stan_cuda_density_eval / stan_cuda_normal_lpdf_eval call through the C ABI (host buffers in
pageable memory, partials returned to the host): where the end-to-end time of the C++ path goes."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from math_b200 import _abi, glm  # noqa: E402

glm.init(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rng = np.random.default_rng(0)
theta = rng.standard_normal(n)
y = (rng.random(n) < 0.5).astype(np.int32)
F = _abi.LP_CONST | _abi.LP_MAIN | _abi.NEED_DALPHA
lam, prob = np.exp(0.5 * theta), 1.0 / (1.0 + np.exp(-theta))
for name, f in (("bernoulli_logit value+partials", lambda: glm.density_eval("bernoulli_logit", y, theta, F)),
                ("bernoulli_logit value only", lambda: glm.density_eval("bernoulli_logit", y, theta, _abi.LP_MAIN)),
                ("poisson_log value+partials", lambda: glm.density_eval("poisson_log", y, theta, F)),
                ("poisson value+partials", lambda: glm.density_eval("poisson", y, lam, F)),
                ("bernoulli value+partials", lambda: glm.density_eval("bernoulli", y, prob, F)),
                ("normal_lpdf mu vector", lambda: glm.normal_lpdf(theta, theta * 0.1, [1.3]))):
    for _ in range(3):
        f()
    t = []
    for _ in range(20):
        t0 = time.perf_counter()
        f()
        t.append(time.perf_counter() - t0)
    print("%-34s n=%d median %.3f ms best %.3f ms" % (name, n, 1e3 * np.median(t), 1e3 * min(t)), flush=True)
from oracle import loader as L  # noqa: E402
if L.have_ref():  # the compiled reference on one core: forward + grad() over a vector of var
    for d, op in (("poisson", lam), ("bernoulli", prob)):
        r = L.ref_density(d, y, op)
        print("%-34s n=%d reference, one core %.3f ms" % (d, n, 1e3 * r["seconds"]), flush=True)
