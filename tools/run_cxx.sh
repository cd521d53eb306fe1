# C++ harness over the reference signatures (bench/stan_bench.cpp), one GPU: C1, C2, C3, C5
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb:$LD_LIBRARY_PATH
export STAN_CUDA_DEVICES=1
for c in c1 c2 c5; do
  timeout 400 build/cxx/stan_bench --config $c --reps 50 --cpu-reps 3 > gpurun_out/r01_cxx_$c.json 2>gpurun_out/cxx_$c.err
  cat gpurun_out/r01_cxx_$c.json
done
timeout 400 build/cxx/stan_bench --config c3 --reps 30 --cpu-reps 1 > gpurun_out/r01_cxx_c3.json 2>gpurun_out/cxx_c3.err
cat gpurun_out/r01_cxx_c3.json
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | grep "^{" | tee gpurun_out/r01_bench_ref_c3.json | cut -c1-600
