#!/bin/bash
O=gpurun_out/r02k; mkdir -p $O
( timeout 900 python -m pytest tests/test_glm_gpu.py tests/test_stress_gpu.py tests/test_wide_gpu.py tests/test_multi_gpu.py -m gpu -x -q -k "ordered or widened or golden" ) > $O/pytest_ord.log 2>&1; echo "pytest ord rc=$?" | tee -a $O/rc.txt
timeout 600 python tools/ab.py --widened --only ordered > $O/ab_ordered.jsonl 2>/dev/null; cut -c1-170 $O/ab_ordered.jsonl
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb
timeout 600 build/cxx/stan_bench --config c5 > $O/cxx_c5.json 2>/dev/null; cat $O/cxx_c5.json
tail -2 $O/pytest_ord.log; cat $O/rc.txt
