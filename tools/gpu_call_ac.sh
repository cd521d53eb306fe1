#!/bin/bash
# faster block inverse / potrf spine: tests, timings, launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or cholesky or tri" 2>&1 | tail -6 > gpurun_out/ac_tests.log
timeout 600 python tools/chol_timing.py 512 1024 2048 4096 8192 > gpurun_out/ac_chol.jsonl 2> gpurun_out/ac.err
timeout 600 python tools/tri_timing.py > gpurun_out/ac_tri.jsonl 2>> gpurun_out/ac.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ac_ncu_chol_4096.csv python tools/chol_timing.py 4096 > gpurun_out/ac_ncu.log 2>&1
cat gpurun_out/ac_tests.log gpurun_out/ac_chol.jsonl gpurun_out/ac_tri.jsonl; tail -3 gpurun_out/ac.err
