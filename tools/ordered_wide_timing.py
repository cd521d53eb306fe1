"""Times ordered_logistic_glm_lpmf on the device either side of the 1024-cut-point boundary (fused kernel up to it,
location / rows / gradient passes beyond: glm_ordered_wide.cu) beside the oracle port on one host core.  Host buffers in
and out, data matrix cached on the device after the first call.  One JSON line per shape; run on the GPU box."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from math_b200 import _abi, glm  # noqa: E402
from oracle import loader as L  # noqa: E402

glm.init(1)
flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
for n, k, ncuts in ((1_000_000, 32, 1000), (1_000_000, 32, 1100), (1_000_000, 32, 20000), (200_000, 600, 1000),
                    (200_000, 600, 1100)):
    rng = np.random.default_rng(n + k + ncuts)
    x = np.asfortranarray(rng.standard_normal((n, k)))
    beta = rng.standard_normal(k) / np.sqrt(k)
    gap = 8.0 / ncuts
    cuts = np.linspace(-4.0, 4.0, ncuts) + rng.uniform(0.0, 0.3 * gap, ncuts)
    y = (1 + np.searchsorted(cuts, x @ beta + rng.logistic(size=n))).astype(np.int32)
    for _ in range(3):
        r = glm.ordered_logistic_glm_eval(y, x, beta, cuts, flags)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter()
        r = glm.ordered_logistic_glm_eval(y, x, beta, cuts, flags)
        ts.append(time.perf_counter() - t0)
    t0 = time.perf_counter()
    o = L.oracle_ordered_logistic_glm(x, y, beta, cuts)
    t_cpu = time.perf_counter() - t0
    err = max(abs(r["lp"] - o["lp"]) / abs(o["lp"]), float(np.max(np.abs(r["dcuts"] - o["dcuts"])) / np.max(np.abs(o["dcuts"]))),
              float(np.max(np.abs(r["dbeta"] - o["dbeta"])) / np.max(np.abs(o["dbeta"]))))
    print(json.dumps({"n": n, "k": k, "cut_points": ncuts, "path": "fused" if ncuts <= 1024 else "general (3 passes)",
                      "device_call_ms_median": 1e3 * float(np.median(ts)), "device_call_ms_min": 1e3 * min(ts),
                      "x_bytes": n * k * 8, "oracle_one_core_ms": 1e3 * t_cpu, "max_rel_err": err, "rc": r["rc"]}), flush=True)
