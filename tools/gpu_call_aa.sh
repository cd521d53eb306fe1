#!/bin/bash
# tri solve + split-K + batched inverses: tests, C++ binaries, timings
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or multiply or Multiply or cholesky or Cholesky or tri" 2>&1 | tail -30 > gpurun_out/aa_tests.log
timeout 600 python tools/chol_timing.py 1024 2048 4096 8192 > gpurun_out/aa_chol.jsonl 2> gpurun_out/aa.err
timeout 600 python tools/tri_timing.py > gpurun_out/aa_tri.jsonl 2>> gpurun_out/aa.err
timeout 600 python tools/gemm_timing.py > gpurun_out/aa_gemm.jsonl 2>> gpurun_out/aa.err
cat gpurun_out/aa_tests.log gpurun_out/aa_chol.jsonl gpurun_out/aa_tri.jsonl; python - <<'PY'
import json
for ln in open("gpurun_out/aa_gemm.jsonl"):
    d=json.loads(ln); print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d.items() if k in ("m","k","n","resident_ms_min","resident_tflops","host_to_host_ms_min","cublas_f64_tflops","reference_eigen_one_core_ms")})
PY
tail -5 gpurun_out/aa.err
