#!/bin/bash
# hoisted GEMM loaders: tests, tile variants, transfer-path probe
set -x
mkdir -p gpurun_out
true
timeout 1500 python -m pytest tests/test_dense.py -q -m gpu -x 2>&1 | tail -5 > gpurun_out/y_tests.log
for v in 0 1; do echo "variant $v" >> gpurun_out/y_variants.jsonl; STAN_CUDA_GEMM_TILE=$v timeout 600 python tools/gemm_timing.py 4096x4096x4096 8192x8192x8192 >> gpurun_out/y_variants.jsonl 2>> gpurun_out/x.err; done
timeout 600 python tools/chol_timing.py 1024 2048 4096 8192 > gpurun_out/y_chol.jsonl 2>> gpurun_out/x.err
cat gpurun_out/y_tests.log gpurun_out/y_chol.jsonl; python - <<'PY'
import json
for ln in open("gpurun_out/y_variants.jsonl"):
    if ln.startswith("{"):
        d=json.loads(ln); print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d.items() if k in ("m","resident_ms_min","resident_tflops","host_to_host_ms_min","cublas_f64_tflops","upload_a_ms","upload_a_gbs","download_fresh_ms","download_touched_ms")})
    else: print(ln.strip())
PY
tail -5 gpurun_out/x.err
