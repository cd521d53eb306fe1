#!/bin/bash
# ordered_logistic beyond 1024 cut-points: parity tests, C++ parity binary, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_glm_gpu.py tests/test_wide_gpu.py tests/test_stress_gpu.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/am_pytest.log
cat gpurun_out/am_pytest.log
timeout 300 build/cxx/cuda_glm_parity_test 2>&1 | tail -8 > gpurun_out/am_cxx.log
cat gpurun_out/am_cxx.log
timeout 300 build/cxx/ref_ordered_logistic_glm_lpmf_test 2>&1 | tail -3
timeout 600 python tools/ordered_wide_timing.py > gpurun_out/am_timing.jsonl 2> gpurun_out/am_timing.err
cat gpurun_out/am_timing.jsonl; tail -3 gpurun_out/am_timing.err
