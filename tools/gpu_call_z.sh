#!/bin/bash
# cholesky adjoint launch list at n = 4096; C++ binaries for multiply / cholesky; stan_bench m1 m2
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cxx_gpu.py -q -m gpu -k "multiply or cholesky" 2>&1 | tail -8 > gpurun_out/z_tests.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/z_ncu_chol_4096.csv python tools/chol_timing.py 4096 > gpurun_out/z_ncu.log 2>&1
for c in m1 m2; do timeout 600 build/cxx/stan_bench --config $c --reps 10 --cpu-reps 2 >> gpurun_out/z_cxx.jsonl 2>> gpurun_out/z.err; done
cat gpurun_out/z_tests.log gpurun_out/z_cxx.jsonl; tail -3 gpurun_out/z.err; tail -3 gpurun_out/z_ncu.log
