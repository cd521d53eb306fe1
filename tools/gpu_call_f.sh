#!/bin/bash
# Round 2, GPU call F (1 GPU): full suite, C1 / d1 with the parallel chain(), ncu launch list of bench.py and
# ncu --set full captures of the dominant kernels.
O=gpurun_out/r02f; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
for c in c1 d1 d2 v2; do
  timeout 600 build/cxx/stan_bench --config $c > $O/cxx_$c.json 2> $O/cxx_$c.err; echo "stan_bench $c rc=$?" | tee -a $O/rc.txt
done
timeout 300 build/cxx/host_phases > $O/host_phases.json 2> $O/host_phases.err
unset LD_LIBRARY_PATH
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > $O/plain_bench.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c3_launches.csv $B > $O/ncu_launches.log 2>&1; echo "ncu launches rc=$?" | tee -a $O/rc.txt
python tools/ncu_one.py poisson_log 10000000 128 warp > $O/plain_one_c3.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:glm_warp_kernel -s 1 -c 1 -o $O/c3_glm_warp python tools/ncu_one.py poisson_log 10000000 128 warp > $O/ncu_c3.log 2>&1; echo "ncu c3 rc=$?" | tee -a $O/rc.txt
python tools/ncu_one.py bernoulli_logit 1000000 32 warp > $O/plain_one_c2.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:glm_warp_kernel -s 1 -c 1 -o $O/c2_glm_warp python tools/ncu_one.py bernoulli_logit 1000000 32 warp > $O/ncu_c2.log 2>&1; echo "ncu c2 rc=$?" | tee -a $O/rc.txt
python tools/ncu_dx.py > $O/plain_dx.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:dx_outer -s 1 -c 1 -o $O/dx_outer python tools/ncu_dx.py > $O/ncu_dx.log 2>&1; echo "ncu dx rc=$?" | tee -a $O/rc.txt
python tools/ncu_one.py poisson_log 1000000 1024 warp > $O/plain_one_k1024.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file $O/k1024_launches.csv python tools/ncu_one.py poisson_log 1000000 1024 warp > $O/ncu_k1024.log 2>&1; echo "ncu k1024 rc=$?" | tee -a $O/rc.txt
ls -la $O; grep -E "passed|failed" $O/pytest_gpu.log | tail -2; cat $O/rc.txt; head -c 600 $O/cxx_c1.json
