#!/bin/bash
# left-looking in-CTA potrf: tests, timings, launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "cholesky or Cholesky or gp or tri" 2>&1 | tail -6 > gpurun_out/ag_tests.log
timeout 600 python tools/chol_timing.py 512 1024 2048 4096 8192 > gpurun_out/ag_chol.jsonl 2> gpurun_out/af.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"potrf|symbolic|tri_inv" --csv --log-file gpurun_out/ag_ncu_spine.csv python tools/chol_timing.py 2048 > gpurun_out/ag_ncu.log 2>&1
cat gpurun_out/ag_tests.log; cut -c1-300 gpurun_out/ag_chol.jsonl; python - <<'PY'
import csv
from collections import defaultdict
rows=list(csv.reader(open('gpurun_out/ag_ncu_spine.csv')))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
H=rows[hdr]; idx={h:i for i,h in enumerate(H)}
a=defaultdict(list)
for r in rows[hdr+1:]:
    if len(r)>=len(H): a[r[idx['Kernel Name']].split('(')[0][-40:]].append(float(r[idx['Metric Value']].replace(',',''))/1e3)
for k,v in a.items(): print(k, len(v), 'median us', sorted(v)[len(v)//2])
PY
tail -3 gpurun_out/af.err
