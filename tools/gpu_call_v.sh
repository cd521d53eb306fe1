#!/bin/bash
# dense multiply: tile variants A/B, staged copies, tests
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or multiply or Multiply" 2>&1 | tail -40 > gpurun_out/v_tests.log
for v in 0 1; do echo "variant $v" >> gpurun_out/v_variants.jsonl; STAN_CUDA_GEMM_TILE=$v timeout 600 python tools/gemm_timing.py 2048x2048x2048 4096x4096x4096 8192x8192x8192 10000x512x10000 >> gpurun_out/v_variants.jsonl 2>> gpurun_out/v_gemm.err; done
timeout 900 python tools/gemm_timing.py > gpurun_out/v_gemm.jsonl 2>> gpurun_out/v_gemm.err
STAN_CUDA_DENSE_COPY=direct timeout 600 python tools/gemm_timing.py 2048x2048x2048 1000x1000x1000 > gpurun_out/v_gemm_direct_copy.jsonl 2>> gpurun_out/v_gemm.err
cat gpurun_out/v_tests.log; python - <<'PY'
import json
for f in ("gpurun_out/v_variants.jsonl","gpurun_out/v_gemm.jsonl","gpurun_out/v_gemm_direct_copy.jsonl"):
    print(f)
    for ln in open(f):
        if ln.startswith("{"):
            d=json.loads(ln); print({k:(round(v,3) if isinstance(v,float) else v) for k,v in d.items() if k in ("m","k","n","resident_ms_min","resident_tflops","host_to_host_ms_min","cublas_f64_tflops","reference_eigen_one_core_ms","transposed_a_ms")})
        else: print(ln.strip())
PY
tail -5 gpurun_out/v_gemm.err
