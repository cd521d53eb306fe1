#!/bin/bash
# final 1-GPU validation of the round: smoke, full suite, both bench arms as the driver runs them, c2.
O=gpurun_out/r02m; mkdir -p $O
( time timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 ) > $O/bench_ref_c3_g1.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
( time timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 ) > $O/bench_c3_g1.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c2 --steps 50 --warmup 5 > $O/bench_c2_g1.json 2>/dev/null; echo "c2 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c1 > $O/bench_c1.json 2>/dev/null; echo "c1 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c5 > $O/bench_c5.json 2>/dev/null; echo "c5 rc=$?" | tee -a $O/rc.txt
grep -E "passed|failed" $O/pytest_gpu.log | tail -1; cat $O/rc.txt; tail -3 $O/bench.err
python3 - <<'PY'
import json
for f in ('bench_c3_g1','bench_ref_c3_g1','bench_c2_g1','bench_c1','bench_c5'):
    d=json.loads([l for l in open('gpurun_out/r02m/%s.json'%f) if l.startswith('{')][-1])
    print(f, 'value %.2f'%d['value'], 'e2e', d['e2e'].get('value'), 'roofline', {k:d['roofline'][k] for k in ('achieved','frac','kernel_ms')} if 'roofline' in d else None, 'iso', d.get('roofline',{}).get('isolated_launch',{}).get('frac'), 'cpu', d.get('cpu_baseline',{}).get('value'), d.get('parity_rel_err_full_n'), d.get('e2e_view'))
PY
