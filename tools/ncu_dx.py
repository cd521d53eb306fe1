"""var-x evaluation at the C2 shape (bernoulli_logit 1e6 x 32, d/dx requested): target of ncu -k regex:dx_outer."""
import os
import sys

sys.path.insert(0, os.getcwd())
from bench import synth_shard
from math_b200 import glm

glm.init(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 32
x, y, alpha, beta, sigma = synth_shard("bernoulli_logit", n, k, seed=1)
for _ in range(3):
    r = glm.glm_eval("bernoulli_logit", y, x, alpha, beta, need_dx=True)
print(r["lp"], float(abs(r["dx"]).sum()))
