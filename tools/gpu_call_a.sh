#!/bin/bash
# Round 2, first GPU call (1 GPU): full GPU test-suite, the no-allocation test alone, bench both arms,
# the C++ bench configs and the host phase timer.  Outputs under gpurun_out/r02a/.
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
nproc > $O/nproc.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $O/rc.txt
( time timeout 600 python bench.py --steps 20 --warmup 5 ) > $O/bench_c3_g1.json 2> $O/bench_c3_g1.err; echo "bench rc=$?" | tee -a $O/rc.txt
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 ) > $O/bench_ref_c3_g1.json 2> $O/bench_ref_c3_g1.err; echo "bench ref rc=$?" | tee -a $O/rc.txt
timeout 300 python bench.py --workload c2 --steps 50 --warmup 5 > $O/bench_c2_g1.json 2> $O/bench_c2_g1.err; echo "bench c2 rc=$?" | tee -a $O/rc.txt
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
for c in c1 d1 d2 c2 v2 c5 c3; do
  timeout 600 build/cxx/stan_bench --config $c > $O/cxx_$c.json 2> $O/cxx_$c.err; echo "stan_bench $c rc=$?" | tee -a $O/rc.txt
done
timeout 300 build/cxx/host_phases > $O/host_phases.json 2> $O/host_phases.err; echo "host_phases rc=$?" | tee -a $O/rc.txt
timeout 300 python bench.py --workload c1 --steps 20 > $O/bench_c1.json 2> $O/bench_c1.err; echo "bench c1 rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --workload c5 --steps 20 > $O/bench_c5.json 2> $O/bench_c5.err; echo "bench c5 rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_gpu.log; tail -3 $O/pytest_noalloc_alone.log; cat $O/rc.txt
