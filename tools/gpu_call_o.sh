#!/bin/bash
# final 8-GPU check of the headline scaling points (N = 1 and N = 8) with the end-of-round code
O=gpurun_out/r02o; mkdir -p $O
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c3_g1.json 2>/dev/null; echo "g1 rc=$?" | tee -a $O/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_c3_g8.json 2> $O/bench_c3_g8.err; echo "g8 rc=$?" | tee -a $O/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29912 bench.py --gpus 4 --steps 20 --warmup 5 > $O/bench_c3_g4.json 2> $O/bench_c3_g4.err; echo "g4 rc=$?" | tee -a $O/rc.txt
python3 - <<'PY'
import json
for f in ('bench_c3_g1','bench_c3_g4','bench_c3_g8'):
    d=json.loads([l for l in open('gpurun_out/r02o/%s.json'%f) if l.startswith('{')][-1])
    print(f, 'value %.2f'%d['value'], 'ms %.5f'%d['ms_per_step'], 'e2e %.2f'%d['e2e']['value'], d.get('multi_gpu_parity_rel_err'), 'frac %.3f'%d['roofline']['frac'], d['clocks'])
PY
cat $O/rc.txt
