#!/bin/bash
# cholesky through the C++ signatures (k1 prim, k2 rev) at 1024 / 2048 / 4096
set -x
mkdir -p gpurun_out
for n in 1024 2048; do for c in k1 k2; do timeout 600 build/cxx/stan_bench --config $c --n $n --reps 10 --cpu-reps 2 >> gpurun_out/ai_cxx.jsonl 2>> gpurun_out/ai.err; done; done
for c in k1 k2; do timeout 600 build/cxx/stan_bench --config $c --n 4096 --reps 5 --no-cpu >> gpurun_out/ai_cxx.jsonl 2>> gpurun_out/ai.err; done
python - <<'PY'
import json
for ln in open("gpurun_out/ai_cxx.jsonl"):
    d=json.loads(ln); print(d['config'], d['what'][:60], 'gpu', round(d['gpu']['median_ms'],2), 'cpu', round(d.get('cpu_reference_1_thread',{}).get('median_ms',0),1), d.get('parity_rel_err'))
PY
tail -3 gpurun_out/ai.err
