#!/bin/bash
# Round 2, GPU call G (1 GPU): smoke(), full suite, A/B of the dependent launch on the widened families.
O=gpurun_out/r02g; mkdir -p $O
( time timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
for pdl in 1 0; do
  STAN_CUDA_PDL=$pdl timeout 900 python tools/ab.py --widened --cat-kernel mma > $O/ab_widened_pdl$pdl.jsonl 2> $O/ab_widened_pdl$pdl.err; echo "ab widened pdl=$pdl rc=$?" | tee -a $O/rc.txt
done
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_c3_g1.json 2> $O/bench_c3_g1.err; echo "bench rc=$?" | tee -a $O/rc.txt
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
timeout 300 build/cxx/stan_bench --config c1 > $O/cxx_c1.json 2>/dev/null; timeout 300 build/cxx/host_phases > $O/host_phases.json 2>/dev/null
tail -4 $O/smoke.log; grep -E "passed|failed" $O/pytest_gpu.log | tail -2; cat $O/rc.txt
paste -d'\n' $O/ab_widened_pdl1.jsonl $O/ab_widened_pdl0.jsonl | cut -c1-170
