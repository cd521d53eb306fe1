#!/bin/bash
# student_t_lpdf: GPU parity tests, C++ binaries, stan_bench s1, kernel durations under ncu (durations only; not bench values)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_student_t.py tests/test_cxx_gpu.py -q -m gpu -k "student_t or Student" 2>&1 | tail -40 > gpurun_out/q_tests.log
for n in 1000000 10000000; do timeout 300 build/cxx/stan_bench --config s1 --n $n --reps 20 --cpu-reps 5 >> gpurun_out/q_s1.jsonl 2>> gpurun_out/q_s1.err; done
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:student_t --csv --log-file gpurun_out/q_ncu_student_t.csv python tools/student_t_timing.py > gpurun_out/q_ncu.log 2>&1
cat gpurun_out/q_tests.log gpurun_out/q_s1.jsonl; tail -3 gpurun_out/q_s1.err; tail -5 gpurun_out/q_ncu.log
