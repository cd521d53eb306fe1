#!/bin/bash
O=gpurun_out/r02j; mkdir -p $O
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
for c in c1 d1 d2 v2; do
  timeout 300 build/cxx/stan_bench --config $c --reps 50 > $O/cxx_$c.json 2>/dev/null; cut -c1-420 $O/cxx_$c.json; echo
done
for parts in 2 3 6 8; do STAN_CUDA_HOST_PARTS=$parts timeout 300 build/cxx/stan_bench --config c1 --no-cpu --reps 40 > $O/c1_parts$parts.json 2>/dev/null; echo "parts=$parts $(python3 -c "import json;print(json.load(open('$O/c1_parts$parts.json'))['gpu'])")"; done
STAN_NUM_THREADS=16 timeout 300 build/cxx/host_phases > $O/host_phases.json 2>/dev/null; tail -3 $O/host_phases.json
timeout 300 python -m pytest tests/test_cxx_gpu.py -m gpu -x -q 2>&1 | tail -2
