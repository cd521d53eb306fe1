#!/bin/bash
# host-thread sweep of the N-operand paths (C1, d1, v2) through the C++ signatures
O=gpurun_out/r02i; mkdir -p $O
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
for c in c1 d1 v2; do for t in 1 2 4 8 16; do
  timeout 300 build/cxx/stan_bench --config $c --threads $t --no-cpu --reps 40 > $O/cxx_${c}_t$t.json 2>/dev/null
  echo "$c threads=$t $(python3 -c "import json;d=json.load(open('$O/cxx_${c}_t$t.json'));g=d['gpu'];print(g['median_ms'],g['p10_ms'],g['best_ms'])")" | tee -a $O/sweep.txt
done; done
for t in 1 16; do STAN_NUM_THREADS=$t timeout 300 build/cxx/host_phases > $O/host_phases_t$t.json 2>/dev/null; tail -3 $O/host_phases_t$t.json | head -2; done
