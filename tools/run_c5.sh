export LD_LIBRARY_PATH=$PWD/math_b200/lib:$PWD/oracle/_ref/tbb:$LD_LIBRARY_PATH
export STAN_CUDA_DEVICES=1
for t in 0 4 8 16 32; do
  echo "tiles_per_cta=$t"
  STAN_CUDA_SHARED_TILES_PER_CTA=$t timeout 300 build/cxx/stan_bench --config c5 --reps 50 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['gpu_map_rect_zero_copy_view'])"
done
