#!/bin/bash
# Cholesky: download of the factor overlapped with the sweep -- tests, A/B timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dense.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/an_pytest.log
cat gpurun_out/an_pytest.log
for b in cuda_cholesky_test ref_cholesky_decompose_test ref_mix_cholesky_decompose_test ref_rev_multiply_test ref_rev_mdivide_left_tri_test ref_mix_mdivide_left_tri_low_test ref_multi_normal_cholesky_test ref_rev_multi_normal_cholesky_test; do
  timeout 300 build/cxx/$b 2>&1 | tail -1
done
STAN_CUDA_CHOLESKY_OVERLAP=0 timeout 300 python tools/chol_timing.py 1024 2048 4096 > gpurun_out/an_chol_overlap0.jsonl 2>gpurun_out/an_err0.txt
timeout 300 python tools/chol_timing.py 1024 2048 4096 > gpurun_out/an_chol_overlap1.jsonl 2>gpurun_out/an_err1.txt
cut -c1-180 gpurun_out/an_chol_overlap0.jsonl; cut -c1-180 gpurun_out/an_chol_overlap1.jsonl
for n in 1024 2048 4096; do timeout 300 build/cxx/stan_bench --config k1 --n $n 2>&1 | tail -1 | cut -c1-400; done | tee gpurun_out/an_k1.jsonl
