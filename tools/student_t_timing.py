"""Times stan_cuda_student_t_lpdf_eval (host buffers in, host buffers out: the call a Stan model makes) beside the
compiled reference's student_t_lpdf on one host core, for the model statement (data y, scalar nu and sigma, a vector
of locations) and the all-vector call.  Writes one JSON line per shape; run on the GPU box."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from math_b200 import glm  # noqa: E402
from oracle import loader as L  # noqa: E402

glm.init(1)
rng = np.random.default_rng(3)
for n in (1_000_000, 10_000_000):
    y, mu = rng.standard_normal(n), 0.1 * rng.standard_normal(n)
    nu, sg = rng.uniform(1, 30, n), rng.uniform(0.5, 2, n)
    for name, ops, vm in (("model statement: y data, nu scalar, mu vector, sigma scalar", (y, [4.0], mu, [1.3]), 0b1110),
                          ("all four operands vectors of variables", (y, nu, mu, sg), 0b1111)):
        flags = L.student_t_flags(False, vm)
        for _ in range(3):
            r = glm.student_t_eval(*ops, flags)
        ts = []
        for _ in range(10):
            t0 = time.perf_counter()
            r = glm.student_t_eval(*ops, flags)
            ts.append(time.perf_counter() - t0)
        line = {"n": n, "shape": name, "device_call_ms_median": 1e3 * float(np.median(ts)), "device_call_ms_min": 1e3 * min(ts),
                "rc": r["rc"]}
        if L.have_ref():
            ref = L.ref_student_t_lpdf(*ops, var_mask=vm)
            line["reference_one_core_ms"] = 1e3 * ref["seconds"]
            line["lp_rel_err"] = abs(r["lp"] - ref["lp"]) / abs(ref["lp"])
        print(json.dumps(line), flush=True)
