"""One shape, three launches of the GLM kernel: the target of `ncu --set full -k regex:glm_ -s 2 -c 1`.
usage: ncu_one.py FAMILY N K warp|cta [CLASSES]"""
import os
import sys

sys.path.insert(0, os.getcwd())
import numpy as np
import torch

from bench import synth_shard
from math_b200 import _abi, glm
from math_b200._lib import lib

family, n, k, kern = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
ncls = int(sys.argv[5]) if len(sys.argv) > 5 else 5
glm.init(1)
lib().stan_cuda_set_kernel(1 if kern == "warp" else 0)
fam = _abi.FAMILIES[family]
x, y, alpha, beta, sigma = synth_shard(family if fam <= _abi.NORMAL_ID else "poisson_log", n, k, seed=1)
ncuts = 0
flags = glm.glm_flags(fam, False) if fam != _abi.ORDERED_LOGISTIC else (_abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS)
if fam == _abi.ORDERED_LOGISTIC:
    y = (1 + (y + np.arange(n)) % ncls).astype(np.int32)
    ncuts = ncls - 1
cat = fam == _abi.CATEGORICAL_LOGIT
if cat:
    y = (1 + (y + np.arange(n)) % ncls).astype(np.int32)
    flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA
data = glm.GlmData(fam, x, y)
params = np.zeros(data.params_len + ncuts)
params[:k], params[data.kpad] = beta, alpha[0]
params[data.kpad + 1] = 2.5 if fam == _abi.NEG_BINOMIAL_2_LOG else 1.0
params[data.kpad + 2:] = np.linspace(-1.5, 1.5, ncuts)
dr = torch.zeros(data.result_len + ncuts, dtype=torch.float64, device="cuda")
if cat:
    params = np.zeros(ncls * (data.kpad + 1))
    for c in range(ncls):
        params[c * data.kpad:c * data.kpad + k] = beta * (1 + 0.1 * c)
    params[ncls * data.kpad:] = np.linspace(-0.5, 0.5, ncls)
    dr = torch.zeros(4 + ncls * (data.kpad + 1), dtype=torch.float64, device="cuda")
dp = torch.from_numpy(params).cuda()
s = torch.cuda.Stream()
for _ in range(3):
    if cat:
        data.launch_categorical(flags, dp.data_ptr(), ncls, dr.data_ptr(), s.cuda_stream)
    elif ncuts:
        data.launch_ordered(flags, dp.data_ptr(), ncuts, dr.data_ptr(), s.cuda_stream)
    else:
        data.launch(flags, dp.data_ptr(), dr.data_ptr(), s.cuda_stream)
s.synchronize()
print(float(dr[0]))
