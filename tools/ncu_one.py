import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from bench import FAM_ID, synth_shard
from math_b200 import glm
from math_b200._lib import lib
family, n, k, kern = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
glm.init(1)
lib().stan_cuda_set_kernel(1 if kern == "warp" else 0)
x, y, alpha, beta, sigma = synth_shard(family, n, k, seed=1)
fam = FAM_ID[family]
data = glm.GlmData(fam, x, y)
params = np.zeros(data.params_len); params[:k] = beta; params[data.kpad] = alpha[0]; params[data.kpad+1] = 1.0
dp = torch.from_numpy(params).cuda(); dr = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
s = torch.cuda.Stream()
for _ in range(3):
    data.launch(glm.glm_flags(fam, False), dp.data_ptr(), dr.data_ptr(), s.cuda_stream)
s.synchronize()
print(float(dr[0]))
