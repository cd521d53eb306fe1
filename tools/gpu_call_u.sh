#!/bin/bash
# dense multiply on DMMA + staged copies: GPU tests, C++ binaries, timing, launch list
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or multiply or Multiply" 2>&1 | tail -40 > gpurun_out/u_tests.log
timeout 900 python tools/gemm_timing.py > gpurun_out/u_gemm.jsonl 2> gpurun_out/u_gemm.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_nn -s 2 -c 1 -o gpurun_out/u_dgemm_4096 -f python tools/gemm_timing.py 4096x4096x4096 > gpurun_out/u_ncu.log 2>&1
cat gpurun_out/u_tests.log gpurun_out/u_gemm.jsonl; tail -5 gpurun_out/u_gemm.err
