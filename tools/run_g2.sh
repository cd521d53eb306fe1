for v in "fused 1" "nccl 1" "fused 0" "fused 1"; do
  set -- $v
  STAN_CUDA_PAIR=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 100 --warmup 10 --collective $1 --no-cpu-baseline 2>/dev/null | grep "^{" > gpurun_out/g2_$1_$2.json
  python -c "
import json,sys
d=json.load(open('gpurun_out/g2_$1_$2.json')); print('$1 pair=$2', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['ms_per_step'], d['clocks'])"
done
nvidia-smi --query-gpu=index,power.limit,clocks.mem,clocks.max.mem --format=csv
