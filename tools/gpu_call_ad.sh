#!/bin/bash
# copy helper threads: tests, transfer rates, timings
set -x
mkdir -p gpurun_out
nproc > gpurun_out/ad_nproc.txt
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or cholesky or tri or multiply or gp" 2>&1 | tail -6 > gpurun_out/ad_tests.log
for t in 1 2 4 8; do echo "copy threads $t" >> gpurun_out/ad_gemm.jsonl; STAN_CUDA_COPY_THREADS=$t timeout 600 python tools/gemm_timing.py 4096x4096x4096 2048x2048x2048 >> gpurun_out/ad_gemm.jsonl 2>> gpurun_out/ad.err; done
timeout 600 python tools/chol_timing.py 1024 2048 4096 8192 > gpurun_out/ad_chol.jsonl 2>> gpurun_out/ad.err
timeout 600 python tools/tri_timing.py > gpurun_out/ad_tri.jsonl 2>> gpurun_out/ad.err
for c in m1 m2; do timeout 600 build/cxx/stan_bench --config $c --reps 10 --cpu-reps 2 >> gpurun_out/ad_cxx.jsonl 2>> gpurun_out/ad.err; done
cat gpurun_out/ad_nproc.txt gpurun_out/ad_tests.log; python - <<'PY'
import json
for ln in open("gpurun_out/ad_gemm.jsonl"):
    if ln.startswith("{"):
        d=json.loads(ln); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ("m","resident_ms_min","host_to_host_ms_min","upload_a_ms","upload_a_gbs","download_fresh_ms","download_touched_ms")})
    else: print(ln.strip())
PY
cat gpurun_out/ad_chol.jsonl gpurun_out/ad_tri.jsonl; cut -c1-420 gpurun_out/ad_cxx.jsonl; tail -3 gpurun_out/ad.err
