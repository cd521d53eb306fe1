#!/bin/bash
O=gpurun_out/r02l; mkdir -p $O
for spin in 250 50000; do
  STAN_CUDA_SPIN_US=$spin timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c3_spin$spin.json 2>/dev/null
  STAN_CUDA_SPIN_US=$spin timeout 600 python bench.py --workload c2 --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_c2_spin$spin.json 2>/dev/null
done
python3 - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02l/bench_*.json')):
    d=json.loads([l for l in open(f) if l.startswith('{')][-1])
    print(f.split('/')[-1], 'value %.1f'%d['value'], 'ms/step %.5f'%d['ms_per_step'], 'e2e %.1f'%d['e2e']['value'], 'e2e ms %.5f'%d['e2e']['ms_per_step'], d['e2e']['ms_per_step_stats'])
PY
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
for spin in 250 50000; do for c in c3 c2; do STAN_CUDA_SPIN_US=$spin timeout 300 build/cxx/stan_bench --config $c --no-cpu > $O/cxx_${c}_spin$spin.json 2>/dev/null; echo "$c spin=$spin $(cut -c1-330 $O/cxx_${c}_spin$spin.json)"; done; done
