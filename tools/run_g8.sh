set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus 8 --steps 200 --warmup 10 --collective fused --no-cpu-baseline 2>gpurun_out/g8_fused.err | grep '^{' > gpurun_out/r01_bench_c3_g8_fused_warp.json
timeout 300 $TR bench.py --gpus 8 --steps 200 --warmup 10 --collective nccl --no-cpu-baseline 2>gpurun_out/g8_nccl.err | grep '^{' > gpurun_out/r01_bench_c3_g8_nccl_warp.json
timeout 400 $TR bench.py --gpus 8 --steps 100 --warmup 5 --workload c4 --no-cpu-baseline 2>gpurun_out/g8_c4.err | grep '^{' > gpurun_out/r01_bench_c4_g8_warp.json
export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb:$LD_LIBRARY_PATH
STAN_CUDA_DEVICES=8 STAN_NUM_THREADS=32 timeout 300 build/cxx/stan_bench --config c5 --reps 50 --cpu-reps 3 > gpurun_out/r01_cxx_c5_g8_warp.json 2>gpurun_out/cxx_c5.err
STAN_CUDA_DEVICES=8 STAN_NUM_THREADS=32 timeout 300 build/cxx/stan_bench --config c3 --reps 50 --no-cpu > gpurun_out/r01_cxx_c3_g8_warp.json 2>gpurun_out/cxx_c3.err
tail -3 gpurun_out/*.err
cat gpurun_out/r01_*_warp.json
