#!/bin/bash
# student_t_lpdf: GPU parity tests, the C++ binaries, host-call timing beside the reference
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_student_t.py tests/test_cxx_gpu.py -q -m gpu -k "student_t or Student" 2>&1 | tail -15 > gpurun_out/p_tests.log
timeout 600 python tools/student_t_timing.py > gpurun_out/p_timing.jsonl 2> gpurun_out/p_timing.err
cat gpurun_out/p_tests.log gpurun_out/p_timing.jsonl; tail -3 gpurun_out/p_timing.err
