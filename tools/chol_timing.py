"""Times stan_cuda_cholesky_decompose / stan_cuda_cholesky_adjoint (host matrices in and out) beside the compiled
reference (Eigen LLT and the cholesky_block reverse sweep on one core).  One JSON line per size; run on the GPU box."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from math_b200 import dense, glm  # noqa: E402
from oracle import loader as L  # noqa: E402

glm.init(1)
sizes = [int(v) for v in sys.argv[1:]] or [512, 1024, 2048, 4096, 8192]
for n in sizes:
    rng = np.random.default_rng(n)
    x = rng.uniform(-1, 1, (n, n))
    a = x @ x.T + n * np.eye(n)
    a = np.asfortranarray(0.5 * (a + a.T))   # column-major like the callers' Eigen matrices
    w = np.asfortranarray(np.tril(rng.uniform(-1, 1, (n, n))))
    for _ in range(2):
        l = dense.cholesky_decompose(a)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        l = dense.cholesky_decompose(a)
        ts.append(time.perf_counter() - t0)
    line = {"n": n, "decompose_ms_min": 1e3 * min(ts), "decompose_gflops": n ** 3 / 3 / min(ts) / 1e9}
    l = np.asfortranarray(l)
    adj = dense.cholesky_adjoint(l, w)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        adj = dense.cholesky_adjoint(l, w)
        ts.append(time.perf_counter() - t0)
    line["adjoint_ms_min"] = 1e3 * min(ts)
    if L.have_ref() and n <= 2048:
        r = L.ref_cholesky_rev(a, w)
        line["reference_forward_ms"] = 1e3 * r["forward_s"]
        line["reference_grad_ms"] = 1e3 * r["grad_s"]
        line["factor_rel_err"] = float(np.max(np.abs(np.tril(l - r["l"]))) / np.max(np.abs(r["l"])))
        line["adjoint_rel_err"] = float(np.max(np.abs(adj - r["a_adj"])) / np.max(np.abs(r["a_adj"])))
    print(json.dumps(line), flush=True)
