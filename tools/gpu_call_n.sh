#!/bin/bash
# final 2-GPU validation: multi-GPU tests (widened families now launch dependently too), bench N=2 both arms
O=gpurun_out/r02n; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_multi_gpu.py -m gpu -x -q ) > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?" | tee -a $O/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29811 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > $O/bench_ref_g2.json 2> $O/bench_ref_g2.err; echo "ref g2 rc=$?" | tee -a $O/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_c3_g2.json 2> $O/bench_c3_g2.err; echo "bench g2 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c3_g1.json 2>/dev/null; echo "bench g1 rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_multi.log; cat $O/rc.txt
python3 - <<'PY'
import json
for f in ('bench_c3_g1','bench_c3_g2','bench_ref_g2'):
    d=json.loads([l for l in open('gpurun_out/r02n/%s.json'%f) if l.startswith('{')][-1])
    print(f, 'value %.2f'%d['value'], 'e2e %.2f'%d['e2e']['value'], d.get('multi_gpu_parity_rel_err'), d.get('roofline',{}).get('frac'))
PY
