#!/bin/bash
# beta_lpdf / weibull_lpdf on the device: python + C++ tests, timing; plus the neighbours that share the kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_locscale.py tests/test_glm_gpu.py tests/test_configs_gpu.py -q -m gpu -k "locscale or normal or c1 or C1" 2>&1 | tail -6
for b in cuda_locscale_parity_test cuda_locscale_distr_test cuda_normal_distr_test ref_beta_test ref_weibull_test ref_beta_log_test ref_weibull_log_test ref_normal_test cuda_routing_static_test; do
  timeout 300 build/cxx/$b 2>&1 | tail -1
done
timeout 300 build/cxx/cuda_locscale_parity_test 2>&1 | grep -E "Failure|got " | head -10
for c in l5 l6 c1; do timeout 200 build/cxx/stan_bench --config $c 2>&1 | tail -1 | cut -c1-420; done | tee gpurun_out/ax_l.jsonl
