#!/bin/bash
# gp_exp_quad_cov + mdivide_right_tri + everything dense; then the WHOLE gpu suite, smoke and a default bench run
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dense.py tests/test_cxx_gpu.py -q -m gpu -k "dense or gp or tri" 2>&1 | tail -12 > gpurun_out/ab_tests_dense.log
timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -8 > gpurun_out/ab_tests_all.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/ab_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/ab_bench.json 2> gpurun_out/ab_bench.err
cat gpurun_out/ab_tests_dense.log gpurun_out/ab_tests_all.log; tail -2 gpurun_out/ab_smoke.log; cat gpurun_out/ab_bench.json | cut -c1-900; tail -3 gpurun_out/ab_bench.err
