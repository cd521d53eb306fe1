#!/bin/bash
# staged copies for the pageable vectors of the elementwise densities: tests + c1 / d1 / d2 / s1 through C++
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_student_t.py tests/test_density_gpu.py tests/test_configs_gpu.py tests/test_glm_gpu.py tests/test_cxx_gpu.py tests/test_stress_gpu.py -q -m gpu -k "student_t or Student or density or normal or c1 or distr or stress" 2>&1 | tail -6 > gpurun_out/ae_tests.log
for c in c1 d1 d2 s1; do timeout 300 build/cxx/stan_bench --config $c --reps 30 --cpu-reps 5 >> gpurun_out/ae_cxx.jsonl 2>> gpurun_out/ae.err; done
timeout 300 python tools/student_t_timing.py > gpurun_out/ae_student_t_timing.jsonl 2>> gpurun_out/ae.err
cat gpurun_out/ae_tests.log; cut -c1-330 gpurun_out/ae_cxx.jsonl; cut -c1-250 gpurun_out/ae_student_t_timing.jsonl; tail -3 gpurun_out/ae.err
