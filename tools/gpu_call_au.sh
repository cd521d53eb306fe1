#!/bin/bash
# final validation of the tree: GPU test-suite, smoke, bench (both arms)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/au_pytest.log
cat gpurun_out/au_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/au_smoke.log
timeout 600 python bench.py > gpurun_out/au_bench.json 2> gpurun_out/au_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/au_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/au_bench_ref.json 2> gpurun_out/au_bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/au_bench_ref.json
