#!/bin/bash
# Round 2, GPU call C (1 GPU): full GPU suite with the new paths, then A/B of programmatic dependent launch.
O=gpurun_out/r02c; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
for pdl in 1 0; do
  for wl in "--workload c3" "--workload c3 --rows 1250000" "--workload c2" "--workload c4 --rows 2500000"; do
    tag=$(echo "$wl" | tr -d ' -' ); 
    STAN_CUDA_PDL=$pdl timeout 600 python bench.py $wl --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_${tag}_pdl$pdl.json 2> $O/bench_${tag}_pdl$pdl.err; echo "bench $tag pdl=$pdl rc=$?" | tee -a $O/rc.txt
  done
done
STAN_CUDA_PDL_SLACK=0 timeout 600 python bench.py --workload c3 --rows 1250000 --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_c3small_slack0.json 2>/dev/null
STAN_CUDA_PDL_SLACK=2 timeout 600 python bench.py --workload c3 --rows 1250000 --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_c3small_slack2.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02c/bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f.split('/')[-1], 'ms/step %.5f'%d['ms_per_step'], 'median-step %.5f'%d['ms_per_step_stats']['median'], 'kernel %.5f'%d['roofline']['kernel_ms'], 'e2e %.5f'%d['e2e']['ms_per_step'], 'frac %.3f'%d['roofline']['frac'], d['consistent_device_vs_host_api'])
    except Exception as e: print(f, 'ERR', e)
PY
grep -E "passed|failed" $O/pytest_gpu.log | tail -3; cat $O/rc.txt
