#!/bin/bash
# A/B of the triangle-only transfers in one run, interleaved
mkdir -p gpurun_out; rm -f gpurun_out/aq_ab.jsonl
for rep in 1 2; do for t in 0 1; do
  echo "tri copy $t (rep $rep)" >> gpurun_out/aq_ab.jsonl
  STAN_CUDA_TRI_COPY=$t timeout 300 python tools/chol_timing.py 1024 2048 4096 2>/dev/null | cut -c1-140 >> gpurun_out/aq_ab.jsonl
  STAN_CUDA_TRI_COPY=$t timeout 300 python tools/tri_timing.py 2>/dev/null | cut -c1-100 >> gpurun_out/aq_ab.jsonl
  STAN_CUDA_TRI_COPY=$t timeout 300 build/cxx/stan_bench --config k1 --n 2048 2>&1 | tail -1 | cut -c100-260 >> gpurun_out/aq_ab.jsonl
done; done
cat gpurun_out/aq_ab.jsonl
