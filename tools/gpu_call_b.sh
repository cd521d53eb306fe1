#!/bin/bash
# Round 2, second GPU call (1 GPU): wide-matrix tests, cache tests, C++ parity binary, compute-sanitizer.
O=gpurun_out/r02b; mkdir -p $O
( time timeout 900 python -m pytest tests/test_wide_gpu.py tests/test_cache_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q ) > $O/pytest_wide.log 2>&1; echo "pytest wide rc=$?" | tee -a $O/rc.txt
( time timeout 900 python -m pytest tests/test_cxx_gpu.py -m gpu -x -q ) > $O/pytest_cxx.log 2>&1; echo "pytest cxx rc=$?" | tee -a $O/rc.txt
bash tools/sanitize.sh 1 $O/sanitize; echo "sanitize rc=$?" | tee -a $O/rc.txt
tail -15 $O/pytest_wide.log; tail -5 $O/pytest_cxx.log; cat $O/rc.txt
