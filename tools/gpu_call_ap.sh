#!/bin/bash
# triangle-only transfers for cholesky / adjoint / tri solve: tests + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dense.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/ap_pytest.log
cat gpurun_out/ap_pytest.log
for b in cuda_cholesky_test cuda_tri_solve_test ref_cholesky_decompose_test ref_mix_cholesky_decompose_test ref_mdivide_left_tri_test ref_mdivide_right_tri_test ref_mix_mdivide_left_tri_test ref_mix_mdivide_left_tri_low_test ref_multi_normal_cholesky_test ref_rev_multi_normal_cholesky_test; do
  timeout 300 build/cxx/$b 2>&1 | tail -1
done
timeout 300 python tools/chol_timing.py 1024 2048 4096 > gpurun_out/ap_chol.jsonl 2>gpurun_out/ap_err0.txt
timeout 300 python tools/tri_timing.py > gpurun_out/ap_tri.jsonl 2>gpurun_out/ap_err1.txt
cut -c1-200 gpurun_out/ap_chol.jsonl; cut -c1-200 gpurun_out/ap_tri.jsonl
for c in k1 k2; do timeout 300 build/cxx/stan_bench --config $c --n 2048 2>&1 | tail -1 | cut -c1-330; done | tee gpurun_out/ap_k.jsonl
