#!/bin/bash
# gamma_lpdf / exponential_lpdf on the device: python + C++ tests, timing through the C++ signatures
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_locscale.py -q -m gpu 2>&1 | tail -8
for b in cuda_locscale_parity_test cuda_locscale_distr_test ref_gamma_test ref_exponential_test ref_gamma_log_test ref_exponential_log_test cuda_routing_static_test; do
  timeout 300 build/cxx/$b 2>&1 | tail -1
done
timeout 300 build/cxx/cuda_locscale_parity_test 2>&1 | grep -E "Failure|got |d/dalpha" | head -10
for c in l3 l4; do timeout 200 build/cxx/stan_bench --config $c 2>&1 | tail -1 | cut -c1-420; done | tee gpurun_out/aw_l.jsonl
