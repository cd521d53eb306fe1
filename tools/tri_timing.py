"""Times stan_cuda_tri_solve (host matrices in and out) beside the compiled reference's mdivide_left_tri on one core."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from math_b200 import dense, glm  # noqa: E402
from oracle import loader as L  # noqa: E402

glm.init(1)
for (n, m) in [(1024, 1024), (2048, 2048), (4096, 4096), (4096, 64)]:
    rng = np.random.default_rng(n + m)
    a = np.asfortranarray(rng.uniform(-1, 1, (n, n)))
    a[np.diag_indices(n)] = 0.05 * n
    b = np.asfortranarray(rng.uniform(-1, 1, (n, m)))
    for _ in range(2):
        x = dense.tri_solve(a, b)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        x = dense.tri_solve(a, b)
        ts.append(time.perf_counter() - t0)
    line = {"n": n, "m": m, "solve_ms_min": 1e3 * min(ts), "gflops": n * n * m / min(ts) / 1e9}
    if L.have_ref() and n <= 2048:
        xr, secs = L.ref_mdivide_left_tri(a, b, True)
        line["reference_one_core_ms"] = 1e3 * secs
        line["rel_err"] = float(np.max(np.abs(x - xr)) / np.max(np.abs(xr)))
    print(json.dumps(line), flush=True)
