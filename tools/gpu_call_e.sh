#!/bin/bash
# Round 2, GPU call E (8 GPUs): the scaling run as the driver does it (N = 1, 2, 4, 8 back to back), multi-GPU tests
# on 8 devices, C5 / C3 through the C++ signatures with 8 devices in one process.
O=gpurun_out/r02e; mkdir -p $O
nvidia-smi -L > $O/gpus.txt; nproc > $O/nproc.txt
timeout 600 python bench.py --impl reference --gpus 1 --steps 5 --warmup 3 > $O/bench_ref_g1.json 2> $O/bench_ref_g1.err; echo "ref g1 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_c3_g1.json 2> $O/bench_c3_g1.err; echo "bench g1 rc=$?" | tee -a $O/rc.txt
port=29700
for n in 2 4 8; do
  port=$((port+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 20 --warmup 5 > $O/bench_c3_g$n.json 2> $O/bench_c3_g$n.err; echo "bench g$n rc=$?" | tee -a $O/rc.txt
done
port=$((port+1))
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 20 --warmup 5 --collective nccl > $O/bench_c3_g8_nccl.json 2> $O/bench_c3_g8_nccl.err; echo "bench g8 nccl rc=$?" | tee -a $O/rc.txt
port=$((port+1))
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --workload c4 --steps 20 --warmup 5 > $O/bench_c4_g8.json 2> $O/bench_c4_g8.err; echo "bench c4 g8 rc=$?" | tee -a $O/rc.txt
port=$((port+1))
STAN_CUDA_PDL=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_c3_g8_pdl0.json 2> $O/bench_c3_g8_pdl0.err; echo "bench g8 pdl0 rc=$?" | tee -a $O/rc.txt
( time timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q ) > $O/pytest_multi8.log 2>&1; echo "pytest multi rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c5 --gpus 8 --steps 20 > $O/bench_c5_g8.json 2> $O/bench_c5_g8.err; echo "bench c5 g8 rc=$?" | tee -a $O/rc.txt
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
timeout 600 build/cxx/stan_bench --config c3 --no-cpu > $O/cxx_c3_g8.json 2> $O/cxx_c3_g8.err; echo "stan_bench c3 g8 rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02e/bench_c*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        if 'roofline' in d:
            print(f.split('/')[-1], 'value %.1f'%d['value'], 'ms/step %.5f'%d['ms_per_step'], 'median-step %.5f'%d['ms_per_step_stats']['median'], 'max-step %.5f'%d['ms_per_step_stats']['max'], 'kernel %.5f'%d['roofline']['kernel_ms'], 'e2e %.1f'%d['e2e']['value'], 'parity', d.get('multi_gpu_parity_rel_err'))
        else:
            print(f.split('/')[-1], json.dumps(d)[:300])
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 $O/pytest_multi8.log; cat $O/rc.txt
