#!/bin/bash
# cholesky + generalised DGEMM: GPU tests, C++ binaries, timing
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or multiply or Multiply or cholesky or Cholesky" -x 2>&1 | tail -40 > gpurun_out/w_tests.log
timeout 600 python tools/chol_timing.py > gpurun_out/w_chol.jsonl 2> gpurun_out/w_chol.err
timeout 600 python tools/gemm_timing.py 4096x4096x4096 8192x8192x8192 2048x2048x2048 > gpurun_out/w_gemm.jsonl 2>> gpurun_out/w_chol.err
cat gpurun_out/w_tests.log gpurun_out/w_chol.jsonl; python - <<'PY'
import json
for ln in open("gpurun_out/w_gemm.jsonl"):
    d=json.loads(ln); print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d.items() if k in ("m","k","n","resident_ms_min","resident_tflops","host_to_host_ms_min","cublas_f64_tflops","transposed_a_ms")})
PY
tail -5 gpurun_out/w_chol.err
