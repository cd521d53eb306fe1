#!/bin/bash
# elementwise densities after the warp-parallel reduction tail / hoisted scalar-nu terms: tests, s1/c1/d1 through C++, kernel durations
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_student_t.py tests/test_density_gpu.py tests/test_configs_gpu.py tests/test_glm_gpu.py tests/test_cxx_gpu.py -q -m gpu -k "student_t or Student or density or normal or c1 or distr" 2>&1 | tail -40 > gpurun_out/r_tests.log
for c in s1 c1 d1 d2; do timeout 300 build/cxx/stan_bench --config $c --reps 20 --cpu-reps 5 >> gpurun_out/r_cxx.jsonl 2>> gpurun_out/r_cxx.err; done
timeout 300 build/cxx/stan_bench --config s1 --n 10000000 --reps 20 --cpu-reps 5 >> gpurun_out/r_cxx.jsonl 2>> gpurun_out/r_cxx.err
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:student_t --csv --log-file gpurun_out/r_ncu_student_t.csv python tools/student_t_timing.py > gpurun_out/r_ncu.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"normal_lpdf|density_kernel" --csv --log-file gpurun_out/r_ncu_c1_d1.csv bash -c 'build/cxx/stan_bench --config c1 --reps 5 --no-cpu; build/cxx/stan_bench --config d1 --reps 5 --no-cpu; build/cxx/stan_bench --config d2 --reps 5 --no-cpu' > gpurun_out/r_ncu2.log 2>&1
cat gpurun_out/r_tests.log gpurun_out/r_cxx.jsonl; tail -3 gpurun_out/r_cxx.err
