#!/bin/bash
# full validation after the dense work: whole GPU suite, smoke, default bench, reference arm
set -x
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -8 > gpurun_out/ah_tests_all.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/ah_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/ah_bench.json 2> gpurun_out/ah_bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/ah_bench_ref.json 2> gpurun_out/ah_bench_ref.err
cat gpurun_out/ah_tests_all.log; tail -2 gpurun_out/ah_smoke.log; python - <<'PY'
import json
for f in ("gpurun_out/ah_bench.json","gpurun_out/ah_bench_ref.json"):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("impl","metric","value","unit","ms_per_step","e2e","roofline","cpu_baseline","gpu_launches","clocks","config")})
PY
