#!/bin/bash
# per-thread work matrices in the Cholesky node: C++ tests + k2 timing
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_cxx_gpu.py -q -m gpu -k "cholesky or multiply or tri or gp" 2>&1 | tail -4 > gpurun_out/ak_tests.log
for n in 1024 2048; do timeout 600 build/cxx/stan_bench --config k2 --n $n --reps 10 --cpu-reps 2 >> gpurun_out/ak_cxx.jsonl 2>> gpurun_out/ak.err; done
cat gpurun_out/ak_tests.log; python - <<'PY'
import json
for ln in open("gpurun_out/ak_cxx.jsonl"):
    d=json.loads(ln); print(d['config'], d['what'][:40], 'gpu', round(d['gpu']['median_ms'],2), 'cpu', round(d.get('cpu_reference_1_thread',{}).get('median_ms',0),1), d.get('parity_rel_err'))
PY
