#!/bin/bash
# Round 2, GPU call D (2 GPUs): multi-GPU parity tests, var-x tests, bench at N = 1 and N = 2 (fused / nccl, three runs each).
O=gpurun_out/r02d; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
( time timeout 1200 python -m pytest tests/test_multi_gpu.py tests/test_varx_gpu.py -m gpu -x -q ) > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c3_g1.json 2> $O/bench_c3_g1.err; echo "bench g1 rc=$?" | tee -a $O/rc.txt
port=29600
for coll in fused nccl; do
  for run in 1 2 3; do
    port=$((port+1))
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 20 --warmup 5 --collective $coll > $O/bench_c3_g2_${coll}_$run.json 2> $O/bench_c3_g2_${coll}_$run.err; echo "bench g2 $coll $run rc=$?" | tee -a $O/rc.txt
  done
done
port=$((port+1))
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > $O/bench_ref_g2.json 2> $O/bench_ref_g2.err; echo "bench ref g2 rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02d/bench_c3*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f.split('/')[-1], 'value %.1f'%d['value'], 'ms/step %.5f'%d['ms_per_step'], 'median-step %.5f'%d['ms_per_step_stats']['median'], 'max-step %.5f'%d['ms_per_step_stats']['max'], 'kernel %.5f'%d['roofline']['kernel_ms'], 'e2e %.1f'%d['e2e']['value'], 'parity', d.get('multi_gpu_parity_rel_err'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 $O/pytest_multi.log; cat $O/rc.txt
