#!/bin/bash
# dense multiply: GPU tests, C++ binaries, timing beside cuBLAS and the reference
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py tests/test_abi.py -q -m gpu -k "dense or multiply or abi or Multiply" 2>&1 | tail -40 > gpurun_out/s_tests.log
timeout 900 python tools/gemm_timing.py > gpurun_out/s_gemm.jsonl 2> gpurun_out/s_gemm.err
cat gpurun_out/s_tests.log gpurun_out/s_gemm.jsonl; tail -5 gpurun_out/s_gemm.err
