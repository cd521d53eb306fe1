#!/bin/bash
# final validation of the tree
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/ao_pytest.log
cat gpurun_out/ao_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/ao_smoke.log
timeout 600 python bench.py > gpurun_out/ao_bench.json 2> gpurun_out/ao_bench.err; cut -c1-300 gpurun_out/ao_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/ao_bench_ref.json 2> gpurun_out/ao_bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/ao_bench_ref.json
for t in 4 8; do
  echo "copy threads $t" >> gpurun_out/ao_copy_threads.jsonl
  STAN_CUDA_COPY_THREADS=$t timeout 300 python tools/gemm_timing.py 2048x2048x2048 4096x4096x4096 2>/dev/null | cut -c1-900 >> gpurun_out/ao_copy_threads.jsonl
  STAN_CUDA_COPY_THREADS=$t timeout 300 python tools/chol_timing.py 2048 4096 2>/dev/null | cut -c1-400 >> gpurun_out/ao_copy_threads.jsonl
  for c in m1 k1 k2; do STAN_CUDA_COPY_THREADS=$t timeout 300 build/cxx/stan_bench --config $c --n 2048 2>&1 | tail -1 | cut -c1-330 >> gpurun_out/ao_copy_threads.jsonl; done
done
cat gpurun_out/ao_copy_threads.jsonl | cut -c1-330
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -s 4 -c 1 -o gpurun_out/ao_dgemm_4096 python tools/gemm_timing.py 4096x4096x4096 > gpurun_out/ao_ncu.log 2>&1
ncu -i gpurun_out/ao_dgemm_4096.ncu-rep --page details > gpurun_out/ao_dgemm_4096_details.txt 2>&1
grep -n "Duration\|Issue Slots Busy\|Registers Per\|Achieved Occ\|DRAM Throughput\|Compute (SM) Throughput" gpurun_out/ao_dgemm_4096_details.txt | head
