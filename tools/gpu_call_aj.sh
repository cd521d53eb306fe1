#!/bin/bash
# device symmetry check: tests + cholesky through the C++ signatures
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py tests/test_abi.py -q -m gpu -k "dense or cholesky or abi or gp" 2>&1 | tail -6 > gpurun_out/aj_tests.log
for n in 1024 2048; do for c in k1 k2; do timeout 600 build/cxx/stan_bench --config $c --n $n --reps 10 --cpu-reps 2 >> gpurun_out/aj_cxx.jsonl 2>> gpurun_out/aj.err; done; done
for c in k1 k2; do timeout 600 build/cxx/stan_bench --config $c --n 4096 --reps 5 --no-cpu >> gpurun_out/aj_cxx.jsonl 2>> gpurun_out/aj.err; done
cat gpurun_out/aj_tests.log; python - <<'PY'
import json
for ln in open("gpurun_out/aj_cxx.jsonl"):
    d=json.loads(ln); print(d['config'], d['what'][:60], 'gpu', round(d['gpu']['median_ms'],2), 'cpu', round(d.get('cpu_reference_1_thread',{}).get('median_ms',0),1), d.get('parity_rel_err'))
PY
tail -3 gpurun_out/aj.err
