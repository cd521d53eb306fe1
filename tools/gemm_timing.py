"""Times the device DGEMM: resident operands (stan_cuda_matrix_multiply, returns with the stream idle), host to host
(stan_cuda_multiply: upload + product + download), cuBLAS through torch.matmul(float64) as the library yardstick, and
the compiled reference (Eigen, one core) on a smaller cube.  One JSON line per shape; run on the GPU box."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from math_b200 import dense, glm  # noqa: E402
from oracle import loader as L  # noqa: E402

glm.init(1)
rng = np.random.default_rng(1)
shapes = [(1024, 1024, 1024), (2048, 2048, 2048), (4096, 4096, 4096), (8192, 8192, 8192), (10000, 512, 10000),
          (4096, 64, 4096), (300, 200, 100)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in s.split("x")) for s in sys.argv[1:]]
for (m, k, n) in shapes:
    # column-major like the Eigen matrices of the callers (a C-ordered numpy array would be transposed on every call)
    a, b = np.asfortranarray(rng.standard_normal((m, k))), np.asfortranarray(rng.standard_normal((k, n)))
    da, db = dense.MatrixCuda(a), dense.MatrixCuda(b)
    for _ in range(2):
        c = da @ db
    ts = []
    for _ in range(7):
        t0 = time.perf_counter()
        c = da @ db
        ts.append(time.perf_counter() - t0)
        del c
    flops = 2.0 * m * n * k
    line = {"m": m, "k": k, "n": n, "resident_ms_min": 1e3 * min(ts), "resident_ms_median": 1e3 * float(np.median(ts)),
            "resident_tflops": flops / min(ts) / 1e12}
    # A' B and A B' from resident operands (the two products of the reverse sweep)
    dat = dense.MatrixCuda(np.asfortranarray(a.T))
    t0 = time.perf_counter()
    c = dat.T @ db
    line["transposed_a_ms"] = 1e3 * (time.perf_counter() - t0)
    del c, dat
    # the two directions of the host <-> device path on their own: upload of a touched matrix, download into a fresh
    # (never touched) array and into one that has been written before
    t0 = time.perf_counter()
    tmp = dense.MatrixCuda(a)
    line["upload_a_ms"] = 1e3 * (time.perf_counter() - t0)
    line["upload_a_gbs"] = a.nbytes / (time.perf_counter() - t0) / 1e9
    t0 = time.perf_counter()
    fresh = tmp.to_host()
    line["download_fresh_ms"] = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    tmp.to_host(out=fresh)
    line["download_touched_ms"] = 1e3 * (time.perf_counter() - t0)
    del tmp, fresh
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        ch = dense.multiply_host(a, b)
        ts.append(time.perf_counter() - t0)
    line["host_to_host_ms_min"] = 1e3 * min(ts)
    line["host_to_host_bytes"] = 8 * (m * k + k * n + m * n)
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    for _ in range(2):
        tc = ta @ tb
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e30
    for _ in range(5):
        e0.record()
        tc = ta @ tb
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    line["cublas_f64_ms_min"] = best
    line["cublas_f64_tflops"] = flops / best / 1e9
    line["max_abs_diff_vs_cublas"] = float(np.max(np.abs(ch - tc.cpu().numpy())))
    del ta, tb, tc
    if L.have_ref() and m * k * n <= 2048 ** 3:
        _, secs = L.ref_multiply(a, b)
        line["reference_eigen_one_core_ms"] = 1e3 * secs
    print(json.dumps(line), flush=True)
