"""C5 on one device both ways (SURVEY 8e "measure both"): all J shards in ONE batched launch
(stan_cuda_glm_eval_batch) vs J single evaluations inside shard scopes issued by one host thread.
(The C++ map_rect path issues the single evaluations from 16 TBB workers: bench/stan_bench.cpp --config c5.)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synth_shard  # noqa: E402
from math_b200 import glm  # noqa: E402

J = int(sys.argv[1]) if len(sys.argv) > 1 else 64
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
k = 32
glm.init(1)
xs, ys = [], []
for j in range(J):
    x, y, alpha, beta, _ = synth_shard("bernoulli_logit", rows, k, seed=100 + j)
    xs.append(x)
    ys.append(y)
alphas = [0.01 * j for j in range(J)]
batch = glm.GlmBatch("bernoulli_logit", xs, ys, call_id=5)


def timed(f, reps=30):
    for _ in range(3):
        f()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        t.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(t)), 1e3 * min(t)


def singles():
    for j in range(J):
        with glm.shard_scope(6, j, 0):
            glm.glm_eval("bernoulli_logit", ys[j], xs[j], [alphas[j]], beta)


rc, res = batch.eval(alphas, beta)
one = glm.glm_eval("bernoulli_logit", ys[3], xs[3], [alphas[3]], beta)
mb, bb = timed(batch.call)   # the C-ABI call alone: parameters were written by eval() above
ms, bs = timed(singles, reps=10)
bytes_ = J * (rows * k * 8 + rows * 4)
print(json.dumps({"what": "C5 shards on one device: J=%d x %d rows x K=%d bernoulli_logit_glm" % (J, rows, k),
                  "batched_launch_ms": round(mb, 4), "batched_best_ms": round(bb, 4),
                  "batched_GBps": round(bytes_ / (bb * 1e-3) / 1e9, 1),
                  "single_evals_one_thread_ms": round(ms, 4), "single_best_ms": round(bs, 4),
                  "item3_lp_batch": res[3]["lp"], "item3_lp_single": one["lp"], "algorithmic_bytes": bytes_}))
