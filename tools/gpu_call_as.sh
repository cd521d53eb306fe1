#!/bin/bash
# poisson_lpmf / bernoulli_lpmf on the device: python + C++ tests, timing beside the reference
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_density_gpu.py -x -q -m gpu 2>&1 | tail -6 > gpurun_out/as_pytest.log
cat gpurun_out/as_pytest.log
for b in cuda_glm_parity_test cuda_int_density2_distr_test cuda_int_density_distr_test ref_poisson_test ref_bernoulli_test ref_bernoulli_log_test cuda_routing_static_test; do
  timeout 300 build/cxx/$b 2>&1 | tail -1
done
timeout 300 build/cxx/cuda_glm_parity_test --gtest_filter='CudaIntDensity*' 2>&1 | grep -E "FAILED|Failure|got" | head -20
timeout 300 python tools/time_density.py 2>&1 | tail -8 | cut -c1-300 | tee gpurun_out/as_time_density.jsonl
